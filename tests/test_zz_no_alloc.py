"""ngp_composite_train_static -- the no-allocation variant of ngp_composite_train (scratch in a file-scope `static __device__`
array of the library, see include/ngp_b200.h) -- against the existing entry point, through the C ABI, on two calls in a row.

The file sorts last on purpose (pytest -x stops at the first failure).  It must pass alone in a fresh process and
after all the other tests: every buffer it needs is allocated up front, and the two entry points are compared on identical inputs.
The launches are the same kernels in the same order and nothing in them is order-dependent, so the results are compared bit for bit
(the other compositing tests allow float tolerances against the oracle / the reference; equality is the stricter bar)."""
import ctypes as C

import numpy as np
import pytest

from test_gpu_render import _rays, _render_cfg

pytestmark = pytest.mark.gpu


def _call(capi, lib, fn_name, static, n_rays, max_compacted, numsteps, coords, net, rgba, rng, ocfg):
    import gpuutil as G
    import torch
    c = _render_cfg(capi, ocfg)
    d_ns = G.dev(numsteps)
    cc, dloss = G.zeros(max_compacted * 7, torch.float32), G.zeros(max_compacted * 4, torch.int16)
    counter, loss, src = G.zeros(1, torch.int32), G.zeros(n_rays, torch.float32), G.zeros(max_compacted, torch.int32)
    ray_counter, d_mean = G.dev(np.array([n_rays], np.uint32)), G.dev(np.array([0.05], np.float32))
    args = [G.stream(), C.byref(c), n_rays, max_compacted, G.p(ray_counter), G.p(G.dev(rgba)), G.p(d_ns), G.p(G.dev(coords)), G.p(cc), G.p(counter), G.p(G.dev(net)),
            G.p(dloss), 4, G.p(d_mean), rng.state, rng.inc, G.p(loss)]
    if static:
        args += [G.p(src)]
    else:
        ws = G.zeros(lib.ngp_composite_workspace_bytes(n_rays) + 64, torch.uint8)
        args += [G.p(ws), G.p(src)]
    G.sync()                                            # every allocation and upload above is complete before the call under test
    rc = getattr(lib, fn_name)(*args)
    G.sync()
    return rc, dict(counter=G.host(counter, np.uint32).copy(), numsteps=G.host(d_ns, np.uint32).copy(), coords=G.host(cc).copy(), dloss=G.host(dloss, np.uint16).copy(),
                    loss=G.host(loss).copy(), src=G.host(src, np.uint32).copy())


def test_static_scratch_variant_matches_the_existing_entry_point_on_two_calls_in_a_row(capi, orc):
    lib = capi.load()
    ocfg = orc.SamplerCfg(1)
    rng = orc.Pcg32(99)
    # second call: fewer rays and other data, so anything the first call left in the shared array would show
    for call, (n_rays, seed, max_compacted) in enumerate([(3000, 5, 1 << 19), (1111, 6, 30000)]):
        numsteps, coords, net, rgba, total = _rays(n_rays, seed)
        rc_a, want = _call(capi, lib, "ngp_composite_train", False, n_rays, max_compacted, numsteps, coords, net, rgba, rng, ocfg)
        rc_b, got = _call(capi, lib, "ngp_composite_train_static", True, n_rays, max_compacted, numsteps, coords, net, rgba, rng, ocfg)
        assert rc_a == 0 and rc_b == 0, lib.ngp_last_error()
        assert int(want["counter"][0]) > 0
        for key in want:
            assert np.array_equal(got[key], want[key]), f"call {call}: {key} differs between the static-scratch variant and ngp_composite_train"


def test_static_scratch_variant_rejects_sizes_beyond_its_maximum(capi, orc):
    import gpuutil as G
    import torch
    lib = capi.load()
    c = _render_cfg(capi, orc.SamplerCfg(1))
    dummy = G.zeros(64, torch.float32)
    too_many = 65536 + 1                                  # NGP_COMPOSITE_STATIC_MAX_RAYS + 1
    rc = lib.ngp_composite_train_static(G.stream(), C.byref(c), too_many, 1024, G.p(dummy), G.p(dummy), G.p(dummy), G.p(dummy), G.p(dummy), G.p(dummy), G.p(dummy),
                                        G.p(dummy), 4, G.p(dummy), 1, 1, G.p(dummy), None)
    assert rc != 0
    msg = lib.ngp_last_error().decode()
    assert "NGP_COMPOSITE_STATIC_MAX_RAYS" in msg and "65536" in msg, msg
    G.sync()
