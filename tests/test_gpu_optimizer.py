"""GPU parity: parameter initialisation (pcg32 draw order) and the fused Adam + EMA step vs the oracle."""
import ctypes as C

import numpy as np
import pytest

from util import rng_uniform

pytestmark = pytest.mark.gpu


def test_init_params_matches_oracle_bit_exactly(capi, orc):
    import gpuutil as G
    import torch
    lib = capi.load()
    n_grid = 100003
    r_ref, r_gpu = orc.trainer_rng(43), orc.trainer_rng(43)
    ref = orc.init_params(r_ref, n_grid)
    w32, w16 = G.zeros(10240 + n_grid, torch.float32), G.zeros(10240 + n_grid, torch.int16)
    state = C.c_uint64(r_gpu.state)
    capi.check(lib.ngp_init_params(G.stream(), C.byref(state), r_gpu.inc, n_grid, G.p(w32), G.p(w16)))
    G.sync()
    assert state.value == r_ref.state, "the host generator must advance exactly like the reference's"
    assert np.array_equal(G.host(w32), ref)
    assert np.array_equal(G.host(w16, np.float16), ref.astype(np.float16))


@pytest.mark.parametrize("n", [10240 + 40000, 10240 + 40003])
def test_optimizer_steps_match_oracle(capi, orc, n):
    import gpuutil as G
    import torch
    lib = capi.load()
    p0 = rng_uniform(n, 1, -0.5, 0.5)
    opt = orc.Optimizer(p0)
    w32, w16, ema = G.dev(opt.w32), G.dev(opt.w16), G.dev(opt.ema16)
    m1, m2, ps = G.zeros(n + 8, torch.float32), G.zeros(n + 8, torch.float32), G.zeros(n + 8, torch.int32)
    cfg = capi.OptimizerCfg(1e-2, 0.9, 0.99, 1e-15, 1e-6, 128.0, 0.95)
    g = np.random.Generator(np.random.PCG64(4))
    for step in range(4):
        grads = (g.normal(size=n) * 0.05).astype(np.float16)
        grads[10240:][g.random(n - 10240) < 0.7] = 0          # most grid entries untouched: sparse skip path
        opt.step(grads)
        capi.check(lib.ngp_optimizer_step(G.stream(), C.byref(cfg), n, 10240, step, 1e-2, G.p(w32), G.p(w16), G.p(G.dev(grads)), G.p(m1), G.p(m2), G.p(ps), G.p(ema)))
    G.sync()
    assert np.array_equal(G.host(ps, np.uint32)[:n], opt.psteps)
    assert np.allclose(G.host(w32)[:n], opt.w32, rtol=1e-5, atol=1e-7)
    assert np.allclose(G.host(m1)[:n], opt.m1, rtol=1e-5, atol=1e-9)
    assert np.allclose(G.host(m2)[:n], opt.m2, rtol=1e-5, atol=1e-12)
    a, b = G.host(w16, np.float16)[:n].astype(np.float32), opt.w16.astype(np.float32)
    assert np.abs(a - b).max() <= 1e-3 * np.abs(b).max()
    a, b = G.host(ema, np.float16)[:n].astype(np.float32), opt.ema16.astype(np.float32)
    assert np.abs(a - b).max() <= 2e-3 * np.abs(b).max()


@pytest.mark.parametrize("n,cut", [(10240 + 40000, 10240 + 8000), (10240 + 40003, 10240), (10240 + 40003, 10240 + 40000)])
def test_optimizer_step_range_composes_to_the_full_step(capi, n, cut):
    """ngp_optimizer_step_range on [cut, n) and [0, cut) == ngp_optimizer_step on [0, n), bit for bit (elementwise update), and the
    consumed gradients are left zeroed."""
    import gpuutil as G
    import torch
    lib = capi.load()
    p0 = rng_uniform(n, 11, -0.5, 0.5)
    cfg = capi.OptimizerCfg(1e-2, 0.9, 0.99, 1e-15, 1e-6, 128.0, 0.95)
    g = np.random.Generator(np.random.PCG64(5))

    def state():
        return dict(w32=G.dev(p0), w16=G.dev(p0.astype(np.float16)), ema=G.dev(p0.astype(np.float16)), m1=G.zeros(n + 8, torch.float32),
                    m2=G.zeros(n + 8, torch.float32), ps=G.zeros(n + 8, torch.int32))
    a, b = state(), state()
    for step in range(3):
        grads = (g.normal(size=n) * 0.05).astype(np.float16)
        grads[10240:][g.random(n - 10240) < 0.7] = 0
        ga, gb = G.dev(grads), G.dev(grads)
        capi.check(lib.ngp_optimizer_step(G.stream(), C.byref(cfg), n, 10240, step, 1e-2, G.p(a["w32"]), G.p(a["w16"]), G.p(ga), G.p(a["m1"]), G.p(a["m2"]), G.p(a["ps"]), G.p(a["ema"])))
        for lo, hi in ((cut, n), (0, cut)):
            capi.check(lib.ngp_optimizer_step_range(G.stream(), C.byref(cfg), lo, hi, 10240, step, 1e-2, G.p(b["w32"]), G.p(b["w16"]), G.p(gb), G.p(b["m1"]), G.p(b["m2"]),
                                                    G.p(b["ps"]), G.p(b["ema"]), 1))
        G.sync()
        assert not G.host(gb, np.uint16).any(), "gradients must be zero after a zeroing step"
        assert np.array_equal(G.host(ga, np.uint16), grads.view(np.uint16)), "the plain step must not touch the gradients"
    for k in a:
        assert np.array_equal(G.host(a[k]).view(np.uint8), G.host(b[k]).view(np.uint8)), k
    assert lib.ngp_optimizer_step_range(G.stream(), C.byref(cfg), 4, n, 10240, 0, 1e-2, G.p(b["w32"]), G.p(b["w16"]), G.p(gb), G.p(b["m1"]), G.p(b["m2"]),
                                        G.p(b["ps"]), G.p(b["ema"]), 0) != 0                                          # begin not a multiple of 8
