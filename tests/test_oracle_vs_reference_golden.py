"""CPU: the oracle against golden vectors produced by the REFERENCE'S OWN CUDA build.

tests/golden/ref_cuda_step.npz was written by tests/make_golden_from_ref.py on a B200 from oracle/_ref/ref_harness (the
unmodified reference compiled for sm_100): one occupancy refresh and one training step on 4 procedural 64x64 views, 256
rays, parameters from refcuda.start_params.  This pins the oracle -- and with it every GPU parity test that uses the oracle as
its checker -- to the reference itself.  Bars: integer / index / coordinate outcomes bit-exact; network outputs and loss
gradients within the fp16-accumulation tolerance stated in each test."""
import os

import numpy as np
import pytest

import refcuda

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_cuda_step.npz")


@pytest.fixture(scope="module")
def g():
    z = np.load(GOLDEN)
    d = {k: z[k] for k in z.files}
    n_views, res, n_rays, batch, k, nm, nc, n_tot = (int(v) for v in d["meta"])
    d.update(n_views=n_views, res=res, n_rays=n_rays, batch=batch, k=k, nm=nm, nc=nc, n_tot=n_tot)
    return d


def _order(rays, numsteps, coords):
    key = np.ascontiguousarray(rays, np.float32).view(np.uint32).reshape(-1, 6)
    first = np.ascontiguousarray(coords[numsteps[:, 1], :3]).view(np.uint32)
    return np.lexsort(tuple(first[:, c] for c in range(3)) + tuple(key[:, c] for c in range(6)))


def _runs(bases, counts, order):
    lens, starts = counts[order].astype(np.int64), bases[order].astype(np.int64)
    offs = np.concatenate([[0], np.cumsum(lens)[:-1]])
    return np.repeat(starts - offs, lens) + np.arange(int(lens.sum()))


def test_initial_parameters(orc, g):
    L = orc.GridLevels(16, 19, 16)
    p = orc.init_params(orc.trainer_rng(43), L.n_params)
    n = g["params_init_head"].size
    assert np.array_equal(p[:n], g["params_init_head"])
    assert np.array_equal(p[-g["params_init_tail"].size:], g["params_init_tail"])
    assert p.astype(np.float64).sum() == g["params_init_sum"][0] and np.abs(p.astype(np.float64)).sum() == g["params_init_sum"][1]


def test_untrained_grid_refresh_samples_and_first_bitfield(orc, g):
    cfg = orc.SamplerCfg(1)
    grid0 = orc.mark_untrained_grid(cfg, g["xforms"], g["res"], g["res"], g["fx"], g["fx"])
    assert np.array_equal(np.packbits(grid0 < 0), g["grid0_is_unseen"])
    assert set(np.unique(grid0)) <= {0.0, -1.0}
    bits0 = orc.grid_to_bitfield(cfg, grid0, orc.grid_mean(cfg, grid0))
    assert np.array_equal(bits0, g["bitfield0"])
    rng = orc.Pcg32(state=int(g["rng"][2]), inc=int(g["rng"][3]))
    pos, idx = orc.generate_grid_samples(cfg, g["n_tot"], rng, 0, grid0, -0.01)
    n = g["grid_indices_head"].size
    assert np.array_equal(idx[:n], g["grid_indices_head"])
    assert np.array_equal(pos[:n], g["grid_positions_head"])
    assert np.bitwise_xor.reduce(idx) == g["grid_indices_xor"]


def test_density_network_on_refresh_positions(orc, g):
    L = orc.GridLevels(16, 19, 16)
    params = refcuda.start_params(10240 + L.n_params)
    pos = np.ascontiguousarray(g["grid_positions_head"])
    n = pos.shape[0]
    enc = orc.hash_encode_fwd(pos, 3, n, params[10240:].view(np.uint16), L)
    got = orc.density_fwd(params[:10240], enc, n).astype(np.float32)
    want = g["grid_density_head"].view(np.float16).astype(np.float32)
    err = np.abs(got - want)
    assert np.abs(want).max() > 0.5
    assert (err <= 2e-2 * np.abs(want) + 4e-3).mean() > 0.999 and err.max() <= 0.05 * np.abs(want).max()


def test_training_march_per_ray(orc, g):
    cfg = orc.SamplerCfg(1)
    rng = orc.Pcg32(state=int(g["rng"][0]), inc=int(g["rng"][1]))
    s = orc.generate_training_samples(cfg, g["n_rays"], 0, 16 * g["batch"], g["xforms"], g["images"], g["res"], g["res"], g["fx"], g["fx"], g["bitfield1"], rng)
    k = g["k"]
    assert int(s["counters"][0]) == k and int(s["counters"][1]) == g["nm"]
    go, ro = _order(s["rays"][:k], s["numsteps"][:k], s["coords"]), _order(g["rays"], g["numsteps"], g["coords"])
    assert np.array_equal(s["rays"][:k][go], g["rays"][ro])
    assert np.array_equal(s["rgba"][:k][go], g["rays_rgba"][ro])
    assert np.array_equal(s["numsteps"][:k][go, 0], g["numsteps"][ro, 0]), "per-ray sample counts"
    a = s["coords"][_runs(s["numsteps"][:k, 1], s["numsteps"][:k, 0], go)]
    b = g["coords"][_runs(g["numsteps"][:, 1], g["numsteps"][:, 0], ro)]
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), "sample coordinates"
    assert len(np.unique(g["numsteps"][:, 0])) > 10


def test_network_outputs_on_marched_samples(orc, g):
    L = orc.GridLevels(16, 19, 16)
    params = refcuda.start_params(10240 + L.n_params)
    coords = np.ascontiguousarray(g["coords"])
    enc = orc.hash_encode_fwd(coords, 7, g["nm"], params[10240:].view(np.uint16), L)
    want = g["rgbsigma"].view(np.float16).astype(np.float32)
    for mode in (0, 1):             # fp32- and fp16-accumulating restatements both sit inside the bar
        got = orc.network_fwd(params[:10240], enc, coords, g["nm"], mode).astype(np.float32)
        err = np.abs(got - want)
        assert (err <= 2e-2 * np.abs(want) + 4e-3).mean() > 0.999 and err.max() <= 0.05 * np.abs(want).max(), (mode, err.max())


def test_compositing_loss_gradients_and_compaction(orc, g):
    cfg = orc.SamplerCfg(1)
    rng = orc.Pcg32(state=int(g["rng"][0]), inc=int(g["rng"][1]))
    ns = g["numsteps"].copy()
    c = orc.composite_train(g["n_rays"], g["k"], g["batch"], np.ascontiguousarray(g["rays_rgba"]), ns, np.ascontiguousarray(g["coords"]),
                            np.ascontiguousarray(g["rgbsigma"]).view(np.float16), cfg, float(g["grid_mean"][0]), rng)
    assert c["counter"] == g["nc"]
    assert np.array_equal(ns[:, 0], g["numsteps_compacted"][:, 0]), "samples kept per ray"
    assert (ns[:, 0] < g["numsteps"][:, 0]).any()
    assert np.allclose(c["loss"][:g["k"]], g["loss"], rtol=2e-3, atol=1e-9)
    order = np.arange(g["k"])
    gi, ri = _runs(ns[:, 1], ns[:, 0], order), _runs(g["numsteps_compacted"][:, 1], g["numsteps_compacted"][:, 0], order)
    assert np.array_equal(c["coords_compacted"][gi], g["coords_compacted"][ri])
    a, b = c["dloss"][gi].astype(np.float32), g["dloss"].view(np.float16)[ri].astype(np.float32)
    scale, err = np.abs(b).max(), np.abs(a - b)
    assert scale > 0 and (err <= 5e-2 * np.abs(b) + 1e-4 * scale).mean() > 0.9995 and err.max() <= 2e-2 * scale
