"""GPU parity: hash-grid encode forward/backward through the C ABI vs the CPU oracle (BASELINE.json configs[0]).

Bars: corner indices bit-exact; encodings bit-exact against the fp16-accumulating oracle (same operation
order, <= 2 fp16 ulp allowed); gradients within a tolerance scaled by the collision count against the oracle's fp32 sums.
"""
import ctypes as C

import numpy as np
import pytest

from util import rng_uniform, half_ulp_diff

pytestmark = pytest.mark.gpu


def _inputs(orc, n, stride, seed, table_amp, log2_T=19):
    L = orc.GridLevels(16, log2_T, 16)
    xyz = rng_uniform((n, stride), seed)
    table = rng_uniform(L.n_params, seed + 1, -table_amp, table_amp).astype(np.float16)
    return L, xyz, table


# (1 << 18, 3, ..., 19) is BASELINE configs[0] at full size (2^18 points, L=16 F=2 T=2^19); log2_T = 22 is the table of configs[4]
# (levels 0-7 dense, 8-15 hashed, 39 474 024 entries = 150.6 MB fp16: spills L2)
@pytest.mark.parametrize("n,stride,amp,log2_T", [(1 << 16, 3, 1.0, 19), (1 << 16, 7, 1e-4, 19), (1000, 3, 1.0, 19), (1, 7, 1.0, 19), (257, 3, 1.0, 19),
                                                 (1 << 18, 3, 1.0, 19), (1 << 18, 3, 1e-4, 19), (1 << 16, 3, 1.0, 22), (1 << 15, 7, 1.0, 14)])
def test_hash_encode_fwd_matches_oracle(capi, orc, n, stride, amp, log2_T):
    import gpuutil as G
    lib = capi.load()
    L, xyz, table = _inputs(orc, n, stride, 1337, amp, log2_T)
    d, n_entries = capi.make_grid_desc(log2_hashmap_size=log2_T)
    if log2_T == 22:
        assert n_entries == 39474024
    assert n_entries == L.n_entries
    ref, ref_idx = orc.hash_encode_fwd(xyz, stride, n, table.view(np.uint16), L, want_indices=True)
    dx, dt = G.dev(xyz), G.dev(table.view(np.uint16))
    out = G.zeros(n * 32, __import__("torch").int16)
    idx = G.zeros(n * 16 * 8, __import__("torch").int32)
    capi.check(lib.ngp_hash_encode_fwd(G.stream(), G.p(dx), stride, n, G.p(dt), C.byref(d), G.p(out), 32, G.p(idx)))
    G.sync()
    got_idx = G.host(idx, np.uint32).reshape(n, 16, 8)
    assert np.array_equal(got_idx, ref_idx), "hash / dense corner indices must be bit-exact"
    got = G.host(out, np.float16).reshape(n, 32)
    ulp = half_ulp_diff(got, ref)
    assert ulp.max() <= 2, f"encoding differs by {ulp.max()} fp16 ulp"
    # the version without the index dump must give the same bits
    out2 = G.zeros(n * 32, __import__("torch").int16)
    capi.check(lib.ngp_hash_encode_fwd(G.stream(), G.p(dx), stride, n, G.p(dt), C.byref(d), G.p(out2), 32, None))
    G.sync()
    assert np.array_equal(G.host(out2), G.host(out))


def test_hash_encode_fwd_boundary_points(capi, orc):
    """points on the faces of the unit cube index res_l on the dense levels and wrap"""
    import gpuutil as G
    import torch
    lib = capi.load()
    L = orc.GridLevels(16, 19, 16)
    pts = np.array([[0, 0, 0], [1, 1, 1], [1, 0, 0.5], [0.999999, 1, 0], [0.5, 0.5, 0.5]], np.float32)
    table = rng_uniform(L.n_params, 5, -1, 1).astype(np.float16)
    ref, ref_idx = orc.hash_encode_fwd(pts, 3, len(pts), table.view(np.uint16), L, want_indices=True)
    d, _ = capi.make_grid_desc()
    out, idx = G.zeros(len(pts) * 32, torch.int16), G.zeros(len(pts) * 128, torch.int32)
    capi.check(lib.ngp_hash_encode_fwd(G.stream(), G.p(G.dev(pts)), 3, len(pts), G.p(G.dev(table.view(np.uint16))), C.byref(d), G.p(out), 32, G.p(idx)))
    G.sync()
    assert np.array_equal(G.host(idx, np.uint32).reshape(-1, 16, 8), ref_idx)
    assert half_ulp_diff(G.host(out, np.float16).reshape(-1, 32), ref).max() <= 2


def test_hash_encode_empty_is_a_noop(capi):
    import gpuutil as G
    import torch
    lib = capi.load()
    d, _ = capi.make_grid_desc()
    x = G.zeros(8, torch.float32)
    capi.check(lib.ngp_hash_encode_fwd(G.stream(), G.p(x), 3, 0, G.p(x), C.byref(d), G.p(x), 32, None))


# (1 << 18, 3, 19): configs[0] at full size; (1 << 16, 3, 22): the L2-spilling table of configs[4]
@pytest.mark.parametrize("n,stride,log2_T", [(1 << 16, 3, 19), (4099, 7, 19), (1 << 18, 3, 19), (1 << 16, 3, 22)])
def test_hash_encode_bwd_matches_oracle(capi, orc, n, stride, log2_T):
    import gpuutil as G
    import torch
    lib = capi.load()
    L, xyz, _ = _inputs(orc, n, stride, 99, 1.0, log2_T)
    dy = (rng_uniform((n, 32), 7, -1, 1) * np.float32(2 ** -7)).astype(np.float16)
    ref = orc.hash_encode_bwd(xyz, stride, n, dy, L)                      # fp32 sums of fp16-rounded terms
    d, _ = capi.make_grid_desc(log2_hashmap_size=log2_T)
    grad = torch.full((L.n_params,), 0x3C00, dtype=torch.int16, device="cuda")   # garbage: zero_first must clear it
    capi.check(lib.ngp_hash_encode_bwd(G.stream(), G.p(G.dev(xyz)), stride, n, G.p(G.dev(dy.view(np.uint16))), 32, C.byref(d), G.p(grad), 1))
    G.sync()
    got = G.host(grad, np.float16).astype(np.float32)
    # untouched entries stay exactly zero (the optimizer's sparse skip depends on it)
    assert np.array_equal(got == 0, ref == 0) or np.count_nonzero((got == 0) != (ref == 0)) < 1e-4 * got.size
    # tolerance: each atomic add rounds to fp16 (rel 2^-11 of the running sum); k collisions -> ~sqrt(k) roundings
    err = np.abs(got - ref)
    tol = 4e-3 * np.abs(ref) + 2e-3 * np.abs(ref).max() * 0 + 1e-5
    coarse = L.offsets[5] * 2          # dense levels 0-4: thousands of addends per entry
    assert (err[coarse:] <= tol[coarse:] * 4).mean() > 0.9999
    rel_coarse = np.abs(got[:coarse] - ref[:coarse]).sum() / np.abs(ref[:coarse]).sum()
    assert rel_coarse < 2e-2, rel_coarse
    # linearity (size-independent property): sum of all gradients == sum over samples of dy * (sum of weights = 1) per level/feature
    per_slot = dy.astype(np.float32).reshape(n, 16, 2).sum(0)
    for l in range(16):
        s = got[L.offsets[l] * 2: L.offsets[l + 1] * 2].reshape(-1, 2).sum(0)
        assert np.allclose(s, per_slot[l], rtol=5e-2, atol=5e-2), (l, s, per_slot[l])


@pytest.mark.parametrize("n,step", [(5003, 1.6915e-3), (4096, 2e-2), (777, 1e-5)])
def test_hash_encode_bwd_on_ray_ordered_samples(capi, orc, n, step):
    """Consecutive samples of a ray share grid cells on the coarse levels: this is the input the in-warp merge of the backward kernel
    is built for (runs of every length, runs that cross warp boundaries, a ragged last warp).  The merged sums are fp32, so the
    result must sit at least as close to the oracle's fp32 sums as independent fp16 atomics would."""
    import gpuutil as G
    import torch
    lib = capi.load()
    L = orc.GridLevels(16, 19, 16)
    g = np.random.Generator(np.random.PCG64(31))
    per_ray = 37
    n_rays = (n + per_ray - 1) // per_ray
    o = g.random((n_rays, 3), dtype=np.float32) * 0.5 + 0.2
    dvec = g.normal(size=(n_rays, 3)).astype(np.float32)
    dvec /= np.linalg.norm(dvec, axis=1, keepdims=True)
    k = np.arange(per_ray, dtype=np.float32)[None, :, None]
    xyz = np.clip((o[:, None, :] + dvec[:, None, :] * k * np.float32(step)).reshape(-1, 3)[:n], 0, 1).astype(np.float32)
    dy = (rng_uniform((n, 32), 7, -1, 1) * np.float32(2 ** -7)).astype(np.float16)
    ref = orc.hash_encode_bwd(xyz, 3, n, dy, L)
    d, _ = capi.make_grid_desc()
    grad = G.zeros(L.n_params, torch.int16)
    capi.check(lib.ngp_hash_encode_bwd(G.stream(), G.p(G.dev(xyz)), 3, n, G.p(G.dev(dy.view(np.uint16))), 32, C.byref(d), G.p(grad), 1))
    G.sync()
    got = G.host(grad, np.float16).astype(np.float32)
    assert np.count_nonzero((got == 0) != (ref == 0)) <= 1e-4 * np.count_nonzero(ref)
    scale = np.abs(ref).max()
    assert np.abs(got - ref).max() <= 2e-2 * scale, np.abs(got - ref).max() / scale
    nz = ref != 0
    assert np.median(np.abs(got[nz] - ref[nz]) / np.abs(ref[nz])) < 1e-3
    per_slot = dy.astype(np.float32).reshape(n, 16, 2).sum(0)
    for l in range(16):
        ssum = got[L.offsets[l] * 2: L.offsets[l + 1] * 2].reshape(-1, 2).sum(0)
        assert np.allclose(ssum, per_slot[l], rtol=2e-2, atol=2e-2), (l, ssum, per_slot[l])


def test_hash_encode_bwd_accumulates_when_not_zeroed(capi, orc):
    import gpuutil as G
    import torch
    lib = capi.load()
    n = 2048
    L, xyz, _ = _inputs(orc, n, 3, 3, 1.0)
    dy = (rng_uniform((n, 32), 8, -1, 1) * np.float32(2 ** -7)).astype(np.float16)
    d, _ = capi.make_grid_desc()
    dx, ddy = G.dev(xyz), G.dev(dy.view(np.uint16))
    g1, g2 = G.zeros(L.n_params, torch.int16), G.zeros(L.n_params, torch.int16)
    capi.check(lib.ngp_hash_encode_bwd(G.stream(), G.p(dx), 3, n, G.p(ddy), 32, C.byref(d), G.p(g1), 1))
    capi.check(lib.ngp_hash_encode_bwd(G.stream(), G.p(dx), 3, n, G.p(ddy), 32, C.byref(d), G.p(g2), 1))
    capi.check(lib.ngp_hash_encode_bwd(G.stream(), G.p(dx), 3, n, G.p(ddy), 32, C.byref(d), G.p(g2), 0))
    G.sync()
    a, b = G.host(g1, np.float16).astype(np.float32), G.host(g2, np.float16).astype(np.float32)
    fine = slice(int(L.offsets[8]) * 2, None)   # few collisions: 2x within fp16 rounding
    assert np.allclose(b[fine], 2 * a[fine], rtol=2e-3, atol=1e-6)
