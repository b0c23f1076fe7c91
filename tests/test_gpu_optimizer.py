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
