"""GPU parity: the tcgen05 fused density + colour MLPs through the C ABI vs the CPU oracle.

Tolerances: the kernel accumulates fp32 in TMEM, the oracle's mode 0 accumulates fp32 too and rounds at the
same fp16 points, so outputs agree to a few fp16 ulp; the stated bar is rel 2e-2 / abs 2e-3 on outputs, rel 5e-2 on
per-sample input gradients, and rel 1e-2 of the gradient scale on the fp32 weight-gradient sums.
"""
import ctypes as C

import numpy as np
import pytest

from util import rng_uniform

pytestmark = pytest.mark.gpu


def _weights(seed, scale=1.0):
    shapes = [(64, 32), (16, 64), (64, 32), (64, 64), (16, 64)]
    parts = []
    for i, (o, k) in enumerate(shapes):
        s = np.sqrt(6.0 / (o + k)) * scale
        parts.append(rng_uniform(o * k, seed + i, -s, s))
    return np.concatenate(parts).astype(np.float16)


def _inputs(n, seed):
    enc = rng_uniform((n, 32), seed, -1, 1).astype(np.float16)
    coords = rng_uniform((n, 7), seed + 1)
    d = coords[:, 4:7] * 2 - 1
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    coords[:, 4:7] = (d + 1) * 0.5
    return enc, np.ascontiguousarray(coords)


@pytest.mark.parametrize("n", [128, 1, 127, 129, 5000, 1 << 16])
def test_mlp_inference_matches_oracle(capi, orc, n):
    import gpuutil as G
    import torch
    lib = capi.load()
    w = _weights(11)
    enc, coords = _inputs(n, 21)
    ref = orc.network_fwd(w, enc, coords, n).astype(np.float32)
    out = G.zeros(n * 4, torch.int16)
    capi.check(lib.ngp_mlp_inference(G.stream(), G.p(G.dev(enc.view(np.uint16))), G.p(G.dev(coords)), n, G.p(G.dev(w.view(np.uint16))), G.p(out)))
    G.sync()
    got = G.host(out, np.float16).reshape(n, 4).astype(np.float32)
    err = np.abs(got - ref)
    assert np.all(err <= 2e-2 * np.abs(ref) + 2e-3), f"max err {err.max()} (ref scale {np.abs(ref).max()})"
    assert np.abs(ref).max() > 0.05     # the comparison is not vacuous


@pytest.mark.parametrize("n", [128, 300, 1 << 15])
def test_mlp_train_matches_oracle(capi, orc, n):
    import gpuutil as G
    import torch
    lib = capi.load()
    w = _weights(12)
    enc, coords = _inputs(n, 22)
    dL = (rng_uniform((n, 4), 23, -1, 1) * np.float32(2.0 ** -4)).astype(np.float16)
    ref_wg, ref_denc, ref_out = orc.network_fwd_bwd(w, enc, coords, dL, n)
    out = G.zeros(n * 4, torch.int16)
    denc = G.zeros(n * 32, torch.int16)
    wg = G.zeros(capi.NGP_MLP_PARAMS, torch.float32)
    capi.check(lib.ngp_mlp_train(G.stream(), G.p(G.dev(enc.view(np.uint16))), G.p(G.dev(coords)), n, G.p(G.dev(w.view(np.uint16))),
                                 G.p(G.dev(dL.view(np.uint16))), G.p(out), G.p(denc), G.p(wg)))
    G.sync()
    got_out = G.host(out, np.float16).reshape(n, 4).astype(np.float32)
    ro = ref_out.astype(np.float32)
    assert np.all(np.abs(got_out - ro) <= 2e-2 * np.abs(ro) + 2e-3)
    got_denc = G.host(denc, np.float16).reshape(n, 32).astype(np.float32)
    rd = ref_denc.astype(np.float32)
    scale = np.abs(rd).max()
    assert scale > 1e-3
    # a hidden unit whose pre-activation rounds to the other side of zero (fp32 sums in a different order) flips one
    # ReLU mask: allow 0.1 % of entries outside the per-element bar, all of them within 10 % of the gradient scale
    derr = np.abs(got_denc - rd)
    assert (derr <= 5e-2 * np.abs(rd) + 2e-3 * scale).mean() > 0.999, derr.max()
    assert derr.max() <= 0.1 * scale, derr.max()
    got_wg = G.host(wg)
    for name, lo, hi in [("W0", 0, 2048), ("W1", 2048, 3072), ("V0", 3072, 5120), ("V1", 5120, 9216), ("V2", 9216, 10240)]:
        r, g = ref_wg[lo:hi], got_wg[lo:hi]
        s = np.abs(r).max()
        assert s > 0, name
        assert np.abs(g - r).max() <= 1e-2 * s, (name, np.abs(g - r).max(), s)
    # rows 3..15 of V2 receive no gradient (their output grads are zero, ngp_network.h:420-425)
    assert np.all(got_wg[9216 + 3 * 64:] == 0)


def test_mlp_train_accumulates_into_wgrad(capi, orc):
    import gpuutil as G
    import torch
    lib = capi.load()
    n = 1024
    w = _weights(13)
    enc, coords = _inputs(n, 31)
    dL = (rng_uniform((n, 4), 33, -1, 1) * np.float32(2.0 ** -4)).astype(np.float16)
    args = (G.p(G.dev(enc.view(np.uint16))), G.p(G.dev(coords)), n, G.p(G.dev(w.view(np.uint16))), G.p(G.dev(dL.view(np.uint16))))
    wg1, wg2 = G.zeros(10240, torch.float32), G.zeros(10240, torch.float32)
    capi.check(lib.ngp_mlp_train(G.stream(), *args, None, None, G.p(wg1)))
    capi.check(lib.ngp_mlp_train(G.stream(), *args, None, None, G.p(wg2)))
    capi.check(lib.ngp_mlp_train(G.stream(), *args, None, None, G.p(wg2)))
    G.sync()
    a, b = G.host(wg1), G.host(wg2)
    assert np.allclose(b, 2 * a, rtol=1e-4, atol=1e-6 * np.abs(a).max())


def test_network_density_and_inference_chain(capi, orc):
    """hash encode -> MLP through ngp_network_inference / ngp_network_density equals the oracle's two stages"""
    import gpuutil as G
    import torch
    lib = capi.load()
    n = 4097
    L = orc.GridLevels(16, 19, 16)
    d, _ = capi.make_grid_desc()
    w = _weights(14)
    table = rng_uniform(L.n_params, 41, -1, 1).astype(np.float16)
    params = np.concatenate([w, table])
    _, coords = _inputs(n, 42)
    enc = orc.hash_encode_fwd(coords, 7, n, table.view(np.uint16), L)
    ref = orc.network_fwd(w, enc, coords, n).astype(np.float32)
    ref_den = orc.density_fwd(w, enc, n).astype(np.float32)
    ws_bytes = lib.ngp_network_workspace_bytes(n)
    ws = G.zeros(ws_bytes, torch.uint8)
    dp, dc = G.dev(params.view(np.uint16)), G.dev(coords)
    out = G.zeros(n * 4, torch.int16)
    capi.check(lib.ngp_network_inference(G.stream(), G.p(dc), n, None, G.p(dp), C.byref(d), G.p(out), G.p(ws), ws_bytes))
    den = G.zeros(n * 16, torch.int16)
    capi.check(lib.ngp_network_density(G.stream(), G.p(dc), 7, n, G.p(dp), C.byref(d), G.p(den), G.p(ws), ws_bytes))
    G.sync()
    got = G.host(out, np.float16).reshape(n, 4).astype(np.float32)
    assert np.all(np.abs(got - ref) <= 2e-2 * np.abs(ref) + 2e-3)
    gd = G.host(den, np.float16).reshape(n, 16).astype(np.float32)
    assert np.all(np.abs(gd - ref_den) <= 2e-2 * np.abs(ref_den) + 2e-3)
    # a device-side valid count below n leaves the tail untouched
    nv = torch.tensor([1000], dtype=torch.int32, device="cuda")
    out2 = torch.full((n * 4,), 0x7E00, dtype=torch.int16, device="cuda")
    capi.check(lib.ngp_network_inference(G.stream(), G.p(dc), n, G.p(nv), G.p(dp), C.byref(d), G.p(out2), G.p(ws), ws_bytes))
    G.sync()
    g2 = G.host(out2, np.int16).reshape(n, 4)
    assert np.array_equal(g2[:1000], G.host(out, np.int16).reshape(n, 4)[:1000])
    assert np.all(g2[1024:] == 0x7E00)


def test_network_train_reuse_equals_network_train(capi, orc):
    """gathering the encodings the inference pass left behind gives the same step as encoding the compacted batch again"""
    import gpuutil as G
    import torch
    lib = capi.load()
    m, n = 9000, 4096 + 17
    L = orc.GridLevels(16, 19, 16)
    d, _ = capi.make_grid_desc()
    w = _weights(15)
    table = rng_uniform(L.n_params, 43, -1, 1).astype(np.float16)
    params = np.concatenate([w, table])
    _, coords = _inputs(m, 44)
    idx = (rng_uniform(n, 45, 0, 1) * m).astype(np.uint32) % m
    idx[:64] = idx[0]                                       # repeated rows, as the rollover produces
    sub = np.ascontiguousarray(coords.reshape(m, 7)[idx])
    dl = (rng_uniform((n, 4), 46, -1, 1) * 0.01).astype(np.float16)
    dp, dc, dsub, ddl, didx = G.dev(params.view(np.uint16)), G.dev(coords), G.dev(sub), G.dev(dl.view(np.uint16)), G.dev(idx)
    ws_inf_bytes, ws_bytes = lib.ngp_network_workspace_bytes(m), lib.ngp_network_workspace_bytes(n)
    ws_inf, ws_a, ws_b = G.zeros(ws_inf_bytes, torch.uint8), G.zeros(ws_bytes, torch.uint8), G.zeros(ws_bytes, torch.uint8)
    out_inf = G.zeros(m * 4, torch.int16)
    capi.check(lib.ngp_network_inference(G.stream(), G.p(dc), m, None, G.p(dp), C.byref(d), G.p(out_inf), G.p(ws_inf), ws_inf_bytes))
    n_params = params.size
    ga, gb = G.zeros(n_params, torch.int16), G.zeros(n_params, torch.int16)
    oa, ob = G.zeros(n * 4, torch.int16), G.zeros(n * 4, torch.int16)
    capi.check(lib.ngp_network_train(G.stream(), G.p(dsub), n, G.p(dp), C.byref(d), G.p(ddl), G.p(oa), G.p(ga), G.p(ws_a), ws_bytes))
    capi.check(lib.ngp_network_train_reuse(G.stream(), G.p(dsub), n, G.p(dp), C.byref(d), G.p(ddl), G.p(ob), G.p(gb), G.p(ws_inf), m, G.p(didx),
                                           G.p(ws_b), ws_bytes))
    G.sync()
    assert np.array_equal(G.host(oa, np.int16), G.host(ob, np.int16)), "forward outputs must be bit-identical"
    a, b = G.host(ga, np.float16).astype(np.float32), G.host(gb, np.float16).astype(np.float32)
    assert np.array_equal(a[:10240], b[:10240]), "MLP weight gradients must be bit-identical (deterministic slot reduction)"
    assert np.abs(a[10240:]).max() > 0
    # the grid gradient is scattered with fp16 atomics whose order is not fixed: equal up to fp16 rounding of the partial sums
    assert np.allclose(a[10240:], b[10240:], rtol=2e-2, atol=2e-3 * np.abs(a[10240:]).max())
    # aliasing the two workspaces is refused
    assert lib.ngp_network_train_reuse(G.stream(), G.p(dsub), n, G.p(dp), C.byref(d), G.p(ddl), G.p(ob), G.p(gb), G.p(ws_b), m, G.p(didx), G.p(ws_b), ws_bytes) != 0


def test_network_input_gradient_matches_analytic_float64(capi, orc):
    """ngp_network_input_gradient (RenderMode::Normals): d sigma_raw / d position against a float64 restatement of
    grid.h:217-255 (dy/dx of the trilinear blend) chained through the density MLP; dt <- dt / 128 and direction <- 0 like the
    reference, which rescales the whole Coordinate matrix (object.h:149)."""
    import gpuutil as G
    import torch
    lib = capi.load()
    n = 3000
    L = orc.GridLevels(16, 19, 16)
    g = np.random.Generator(np.random.PCG64(11))
    w = ((g.random(capi.NGP_MLP_PARAMS, dtype=np.float32) - 0.5) * 0.6).astype(np.float16)
    table = ((g.random(L.n_params, dtype=np.float32) - 0.5) * 2).astype(np.float16)
    coords = g.random((n, 7), dtype=np.float32)
    params = np.concatenate([w, table])
    desc, _ = capi.make_grid_desc()
    d_coords, d_params = G.dev(coords.copy()), G.dev(params.view(np.uint16))
    ws_bytes = lib.ngp_network_workspace_bytes(n)
    ws = torch.zeros(ws_bytes, dtype=torch.uint8, device="cuda")
    capi.check(lib.ngp_network_input_gradient(G.stream(), G.p(d_coords), n, None, G.p(d_params), C.byref(desc), G.p(ws), ws_bytes))
    G.sync()
    got = G.host(d_coords).reshape(n, 7)
    assert np.array_equal(got[:, 3], coords[:, 3] * np.float32(1 / 128)), "dt is rescaled with the rest of the matrix"
    assert not got[:, 4:].any(), "the direction rows receive a zero gradient"

    # ---- float64 restatement ----
    _, idx = orc.hash_encode_fwd(coords, 7, n, table.view(np.uint16), L, want_indices=True)       # [n][16][8], corner bit 0 = x
    tab = table.astype(np.float64).reshape(-1, 2)
    enc = np.zeros((n, 32)); denc_dpos = np.zeros((n, 32, 3))
    for l in range(16):
        scale = np.float32(L.scales[l])
        p = (coords[:, :3] * scale + np.float32(0.5)).astype(np.float32)
        wgt = (p - np.floor(p)).astype(np.float64)                                                # [n][3]
        v = tab[int(L.offsets[l]) + idx[:, l, :].astype(np.int64)]                                # [n][8][2]
        for c in range(8):
            wc = np.ones(n)
            for d in range(3):
                wc = wc * (wgt[:, d] if (c >> d) & 1 else 1 - wgt[:, d])
            enc[:, 2 * l:2 * l + 2] += wc[:, None] * v[:, c]
        for gd in range(3):
            others = [d for d in range(3) if d != gd]
            for c in range(8):
                if (c >> gd) & 1:
                    continue
                wc = np.full(n, float(scale))
                for d in others:
                    wc = wc * (wgt[:, d] if (c >> d) & 1 else 1 - wgt[:, d])
                denc_dpos[:, 2 * l:2 * l + 2, gd] += wc[:, None] * (v[:, c | (1 << gd)] - v[:, c])
    W0 = w[:2048].astype(np.float64).reshape(64, 32); W1 = w[2048:3072].astype(np.float64).reshape(16, 64)
    pre = enc @ W0.T
    dsig_denc = ((pre > 0) * W1[0][None, :]) @ W0                                                 # [n][32]
    ref = np.einsum("nk,nkd->nd", dsig_denc, denc_dpos)
    err = np.abs(got[:, :3] - ref)
    tol = 5e-2 * np.abs(ref) + 2e-2 * np.abs(ref).mean()
    assert (err <= tol).mean() > 0.98, (err <= tol).mean()
    cos = (got[:, :3] * ref).sum(1) / (np.linalg.norm(got[:, :3], axis=1) * np.linalg.norm(ref, axis=1) + 1e-30)
    assert np.median(cos) > 0.999
