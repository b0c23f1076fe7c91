"""GPU parity: occupancy grid maintenance and ray marching through the C ABI vs the CPU oracle.

Bars (BASELINE.json): occupancy bits and per-ray sample counts bit-exact; sample coordinates bit-exact as well because both
sides use the same explicit operation order (fmaf where the kernels use __fmaf_rn, nothing else contracted)."""
import ctypes as C

import numpy as np
import pytest

from util import rng_uniform, make_cameras, make_density_grid

pytestmark = pytest.mark.gpu


def _cam(capi, W, H, fx):
    c = capi.CameraCfg()
    c.width, c.height, c.fx, c.fy, c.ppx, c.ppy = W, H, fx, fx, 0.5, 0.5
    return c


@pytest.mark.parametrize("aabb_scale", [1, 4])
def test_mark_untrained_grid_matches_oracle(capi, orc, aabb_scale):
    import gpuutil as G
    import torch
    lib = capi.load()
    cfg, ocfg = capi.make_sampler_cfg(aabb_scale), orc.SamplerCfg(aabb_scale)
    xf = make_cameras(5)
    W = H = 100
    fx = 120.0
    ref = orc.mark_untrained_grid(ocfg, xf, W, H, fx, fx)
    out = G.zeros(ref.size, torch.float32)
    cam = _cam(capi, W, H, fx)
    capi.check(lib.ngp_mark_untrained_grid(G.stream(), C.byref(cfg), C.byref(cam), 5, G.p(G.dev(xf)), G.p(out)))
    G.sync()
    got = G.host(out)
    assert np.array_equal(got, ref)
    assert 0 < (ref == 0).mean() < 1 or aabb_scale == 1


@pytest.mark.parametrize("thresh,step", [(-0.01, 0), (0.01, 3)])
def test_generate_grid_samples_matches_oracle(capi, orc, thresh, step):
    import gpuutil as G
    import torch
    lib = capi.load()
    cfg, ocfg = capi.make_sampler_cfg(1), orc.SamplerCfg(1)
    grid = make_density_grid()
    rng = orc.Pcg32(77)
    n = 50000
    rpos, ridx = orc.generate_grid_samples(ocfg, n, rng, step, grid, thresh)
    pos, idx = G.zeros(n * 3, torch.float32), G.zeros(n, torch.int32)
    capi.check(lib.ngp_generate_grid_samples(G.stream(), C.byref(cfg), n, rng.state, rng.inc, step, G.p(G.dev(grid)), thresh, G.p(pos), G.p(idx)))
    G.sync()
    assert np.array_equal(G.host(idx, np.uint32), ridx)
    assert np.array_equal(G.host(pos).reshape(n, 3), rpos)


@pytest.mark.parametrize("n_mult,thresh,step", [(0.02, -0.01, 0), (0.5, 0.01, 3), (1.0, -0.01, 17), (2.0, -0.01, 5), (1.5 + 17 / (1 << 21), 0.01, 40), (0.25, 0.01, 257)])
def test_generate_grid_samples_sorted_is_a_permutation_in_cell_order(capi, orc, n_mult, thresh, step):
    """ngp_generate_grid_samples_sorted: the identical multiset of (position, index) pairs, in Morton order of the first candidate
    cell.  Checked against the plain-order kernel (itself checked against the oracle above) and, for the smallest case, the oracle."""
    import gpuutil as G
    import torch
    lib = capi.load()
    cfg, ocfg = capi.make_sampler_cfg(1), orc.SamplerCfg(1)
    cells = 128 ** 3
    grid = make_density_grid()
    if thresh < 0:
        grid = np.abs(grid)                       # every first candidate is accepted: the output must be sorted by cell
    rng = orc.Pcg32(91)
    n = int(round(n_mult * cells))
    d_grid = G.dev(grid)
    pos, idx = G.zeros(n * 3, torch.float32), G.zeros(n, torch.int32)
    spos, sidx = G.zeros(n * 3 + 3, torch.float32), G.zeros(n + 1, torch.int32)
    capi.check(lib.ngp_generate_grid_samples(G.stream(), C.byref(cfg), n, rng.state, rng.inc, step, G.p(d_grid), thresh, G.p(pos), G.p(idx)))
    capi.check(lib.ngp_generate_grid_samples_sorted(G.stream(), C.byref(cfg), n, rng.state, rng.inc, step, G.p(d_grid), thresh, G.p(spos), G.p(sidx)))
    G.sync()
    a_idx, a_pos = G.host(idx, np.uint32), G.host(pos).reshape(n, 3).view(np.uint32)
    b_idx, b_pos = G.host(sidx, np.uint32), G.host(spos).view(np.uint32)
    assert b_idx[n] == 0 and not b_pos[3 * n:].any(), "wrote past the n-th slot"
    b_idx, b_pos = b_idx[:n], b_pos[:3 * n].reshape(n, 3)

    def canon(i, p):
        o = np.lexsort((p[:, 2], p[:, 1], p[:, 0], i))
        return i[o], p[o]
    ai, ap = canon(a_idx, a_pos)
    bi, bp = canon(b_idx, b_pos)
    assert np.array_equal(ai, bi) and np.array_equal(ap, bp)
    if thresh < 0:
        assert (np.diff((b_idx % cells).astype(np.int64)) >= 0).all(), "not in cell order"
        assert not np.array_equal(a_idx, b_idx)
    if n <= 50000:
        rpos, ridx = orc.generate_grid_samples(ocfg, n, rng, step, grid, thresh)
        ri, rp = canon(ridx, rpos.view(np.uint32))
        assert np.array_equal(ri, bi) and np.array_equal(rp, bp)


@pytest.mark.parametrize("n_mult,n_parts,thresh,step", [(0.25, 2, -0.01, 300), (0.25, 8, 0.01, 301), (1.0, 3, -0.01, 16), (0.37, 5, 0.01, 7), (0.0001, 8, -0.01, 2)])
def test_grid_sample_parts_tile_the_full_call(capi, orc, n_mult, n_parts, thresh, step):
    """ngp_generate_grid_samples_sorted_part (the sharded occupancy refresh of the data-parallel path): the parts write disjoint,
    contiguous slot ranges that tile [0, n), and their union is the full call's output bit for bit."""
    import gpuutil as G
    import torch
    lib = capi.load()
    cfg = capi.make_sampler_cfg(1)
    cells = 128 ** 3
    grid = make_density_grid()
    rng = orc.Pcg32(17)
    n = int(round(n_mult * cells))
    d_grid = G.dev(grid)
    full_pos, full_idx = G.zeros(n * 3, torch.float32), G.zeros(n, torch.int32)
    capi.check(lib.ngp_generate_grid_samples_sorted(G.stream(), C.byref(cfg), n, rng.state, rng.inc, step, G.p(d_grid), thresh, G.p(full_pos), G.p(full_idx)))
    G.sync()
    want_pos, want_idx = G.host(full_pos).view(np.uint32).reshape(n, 3), G.host(full_idx, np.uint32)
    cursor = 0
    for part in range(n_parts):
        pos, idx = G.dev(np.full(n * 3, np.nan, np.float32)), G.dev(np.full(n, 0xFFFFFFFF, np.uint32))
        lo, hi = C.c_uint32(123), C.c_uint32(456)
        capi.check(lib.ngp_generate_grid_samples_sorted_part(G.stream(), C.byref(cfg), n, rng.state, rng.inc, step, G.p(d_grid), thresh, G.p(pos), G.p(idx),
                                                             part, n_parts, C.byref(lo), C.byref(hi)))
        G.sync()
        assert lo.value == cursor and hi.value >= lo.value, (part, lo.value, hi.value, cursor)
        cursor = hi.value
        got_pos, got_idx = G.host(pos).view(np.uint32).reshape(n, 3), G.host(idx, np.uint32)
        assert np.array_equal(got_idx[lo.value:hi.value], want_idx[lo.value:hi.value]) and np.array_equal(got_pos[lo.value:hi.value], want_pos[lo.value:hi.value])
        untouched = np.ones(n, bool)
        untouched[lo.value:hi.value] = False
        assert (got_idx[untouched] == 0xFFFFFFFF).all(), "a part wrote outside the slot range it reported"
    assert cursor == n
    assert lib.ngp_generate_grid_samples_sorted_part(G.stream(), C.byref(cfg), n, rng.state, rng.inc, step, G.p(d_grid), thresh, G.p(full_pos), G.p(full_idx), 3, 3, None, None) != 0


def test_splat_then_ema_equals_update_grid_and_partial_maxima_combine(capi, orc):
    """ngp_splat_grid + ngp_ema_grid are ngp_update_grid's two halves; splatting two halves of the samples into separate grids and
    taking the element-wise max (what the all-reduce between the halves does) gives the same grid bit for bit."""
    import gpuutil as G
    import torch
    lib = capi.load()
    cfg = capi.make_sampler_cfg(1)
    grid = make_density_grid()
    n = 300001
    g = np.random.Generator(np.random.PCG64(5))
    indices = g.integers(0, grid.size, n).astype(np.uint32)
    density = (g.random((n, 16), dtype=np.float32) * 8 - 6).astype(np.float16)
    d_idx, d_den = G.dev(indices), G.dev(density)
    want_grid, want_cur = G.dev(grid), G.zeros(grid.size, torch.float32)
    capi.check(lib.ngp_update_grid(G.stream(), C.byref(cfg), n, G.p(d_idx), G.p(d_den), 16, 3, 0.95, G.p(want_cur), G.p(want_grid)))
    cur_a, cur_b, got_grid = G.zeros(grid.size, torch.float32), G.zeros(grid.size, torch.float32), G.dev(grid)
    cut = 123457
    capi.check(lib.ngp_splat_grid(G.stream(), C.byref(cfg), cut, G.p(d_idx), G.p(d_den), 16, 3, G.p(cur_a)))
    capi.check(lib.ngp_splat_grid(G.stream(), C.byref(cfg), n - cut, C.c_void_p(d_idx.data_ptr() + 4 * cut), C.c_void_p(d_den.data_ptr() + 32 * cut), 16, 3, G.p(cur_b)))
    G.sync()
    combined = torch.maximum(cur_a, cur_b)
    capi.check(lib.ngp_ema_grid(G.stream(), C.byref(cfg), 0.95, G.p(combined), G.p(got_grid)))
    G.sync()
    assert np.array_equal(G.host(combined).view(np.uint32), G.host(want_cur).view(np.uint32))
    assert np.array_equal(G.host(got_grid).view(np.uint32), G.host(want_grid).view(np.uint32))


def test_update_grid_and_bitfield_match_oracle(capi, orc):
    import gpuutil as G
    import torch
    lib = capi.load()
    cfg, ocfg = capi.make_sampler_cfg(1), orc.SamplerCfg(1)
    grid = make_density_grid()
    n = 200000
    g = np.random.Generator(np.random.PCG64(3))
    indices = g.integers(0, grid.size, n).astype(np.uint32)
    density = (g.random((n, 16), dtype=np.float32) * 8 - 6).astype(np.float16)
    ref_grid = grid.copy()
    orc.update_grid(ocfg, indices, density, 16, 0.95, ref_grid)
    d_grid, d_cur = G.dev(grid), G.zeros(grid.size, torch.float32)
    capi.check(lib.ngp_update_grid(G.stream(), C.byref(cfg), n, G.p(G.dev(indices)), G.p(G.dev(density)), 16, 3, 0.95, G.p(d_cur), G.p(d_grid)))
    G.sync()
    got = G.host(d_grid)
    # __expf vs expf: relative 2^-21 or so on the splatted values
    assert np.allclose(got, ref_grid, rtol=2e-6, atol=0)
    assert np.array_equal(got < 0, ref_grid < 0)
    # bitfield from the SAME grid values on both sides: bit-exact
    mean = orc.grid_mean(ocfg, got)
    ref_bits = orc.grid_to_bitfield(ocfg, got, mean)
    d_mean, d_bits = G.zeros(256, torch.float32), G.zeros(grid.size // 8, torch.uint8)
    capi.check(lib.ngp_update_grid_bitfield(G.stream(), C.byref(cfg), G.p(d_grid), 0.01, G.p(d_mean), G.p(d_bits)))
    G.sync()
    assert abs(float(G.host(d_mean)[0]) - mean) <= 1e-6 * abs(mean)
    bits = G.host(d_bits)
    if float(G.host(d_mean)[0]) > 0.01 and mean > 0.01:      # threshold = min(0.01, mean): identical on both sides
        assert np.array_equal(bits, ref_bits)
    assert np.count_nonzero(bits != ref_bits) <= 2
    # coarser cascade contains the max-pooled finer one: every set bit of cascade 0 shows up in the centre octant of cascade 1
    assert bits[grid.size // 16:].any()


def _scene(orc, n_images=4, W=64, H=64):
    xf = make_cameras(n_images, seed=11)
    images = rng_uniform((n_images, H, W, 4), 21).astype(np.float16)
    grid = make_density_grid(seed=13)
    ocfg = orc.SamplerCfg(1)
    bits = orc.grid_to_bitfield(ocfg, grid, 1.0)
    return xf, images, bits


# cone 1/256: variable step (sequential lattice path); max_steps 4096: rays of >64 windows (emit pass marches again)
@pytest.mark.parametrize("n_rays,n_coords,cone,max_steps", [(4096, 1 << 20, 0.0, 1024), (1000, 1 << 20, 0.0, 1024), (4096, 20000, 0.0, 1024),
                                                            (1, 1 << 16, 0.0, 1024), (2048, 1 << 20, 1.0 / 256, 1024), (1024, 1 << 21, 0.0, 4096)])
def test_generate_training_samples_matches_oracle(capi, orc, n_rays, n_coords, cone, max_steps):
    import gpuutil as G
    import torch
    lib = capi.load()
    cfg, ocfg = capi.make_sampler_cfg(1, max_steps=max_steps, cone_angle=cone), orc.SamplerCfg(1, max_steps=max_steps, cone_angle=cone)
    W = H = 64
    fx = 70.0
    xf, images, bits = _scene(orc)
    rng = orc.Pcg32(43)
    ref = orc.generate_training_samples(ocfg, n_rays, 17, n_coords, xf, images.view(np.uint16), W, H, fx, fx, bits, rng)
    cam = _cam(capi, W, H, fx)
    rays, rgba = G.zeros(n_rays * 6, torch.float32), G.zeros(n_rays * 4, torch.float32)
    numsteps, coords = G.zeros(n_rays * 2, torch.int32), G.zeros(n_coords * 7, torch.float32)
    counters, per_ray = G.zeros(2, torch.int32), G.zeros(n_rays, torch.int32)
    ws = G.zeros(lib.ngp_scan_workspace_bytes(n_rays) + 64, torch.uint8)
    capi.check(lib.ngp_generate_training_samples(G.stream(), C.byref(cfg), C.byref(cam), n_rays, 17, n_coords, 4, G.p(G.dev(xf)), G.p(G.dev(images)),
                                                 G.p(G.dev(bits)), 1, rng.state, rng.inc, G.p(rays), G.p(rgba), G.p(numsteps), G.p(coords), G.p(counters),
                                                 G.p(per_ray), G.p(ws)))
    G.sync()
    assert np.array_equal(G.host(per_ray, np.uint32), ref["per_ray_n"]), "per-ray sample counts must be bit-exact"
    cnt = G.host(counters, np.uint32)
    assert list(cnt) == list(ref["counters"])
    kept = int(cnt[0])
    assert kept > 0
    assert np.array_equal(G.host(numsteps, np.uint32).reshape(-1, 2)[:kept], ref["numsteps"][:kept])
    assert np.array_equal(G.host(rays).reshape(-1, 6)[:kept], ref["rays"][:kept])
    assert np.array_equal(G.host(rgba).reshape(-1, 4)[:kept], ref["rgba"][:kept])
    n_written = int(ref["numsteps"][kept - 1].sum())
    assert np.array_equal(G.host(coords).reshape(-1, 7)[:n_written], ref["coords"][:n_written]), "sample coordinates differ"
    if n_coords == 20000:
        assert cnt[1] > n_coords and kept < n_rays      # the budget really clipped rays


def test_generate_training_samples_keeps_empty_rays_for_the_envmap(capi, orc):
    """train_envmap (occupancy_sampler.cu:408): rays without a single sample keep a slot in the ray list (count 0), every other
    ray keeps exactly the samples, the pixel and the order it has without the flag"""
    import gpuutil as G
    import torch
    lib = capi.load()
    n_rays, n_coords, W, H, fx = 4096, 1 << 20, 64, 64, 70.0
    xf, images, bits = _scene(orc)
    rng = orc.Pcg32(43)
    cam = _cam(capi, W, H, fx)
    out = []
    for keep in (0, 1):
        cfg = capi.make_sampler_cfg(1)
        cfg.keep_empty_rays = keep
        rays, rgba = G.zeros(n_rays * 6, torch.float32), G.zeros(n_rays * 4, torch.float32)
        numsteps, coords = G.zeros(n_rays * 2, torch.int32), G.zeros(n_coords * 7, torch.float32)
        counters, per_ray = G.zeros(2, torch.int32), G.zeros(n_rays, torch.int32)
        ws = G.zeros(lib.ngp_scan_workspace_bytes(n_rays) + 64, torch.uint8)
        capi.check(lib.ngp_generate_training_samples(G.stream(), C.byref(cfg), C.byref(cam), n_rays, 17, n_coords, 4, G.p(G.dev(xf)), G.p(G.dev(images)),
                                                     G.p(G.dev(bits)), 1, rng.state, rng.inc, G.p(rays), G.p(rgba), G.p(numsteps), G.p(coords), G.p(counters),
                                                     G.p(per_ray), G.p(ws)))
        G.sync()
        cnt = G.host(counters, np.uint32)
        kept = int(cnt[0])
        out.append(dict(kept=kept, total=int(cnt[1]), per_ray=G.host(per_ray, np.uint32).copy(), ns=G.host(numsteps, np.uint32).reshape(-1, 2)[:kept].copy(),
                        rays=G.host(rays).reshape(-1, 6)[:kept].copy(), rgba=G.host(rgba).reshape(-1, 4)[:kept].copy(), coords=G.host(coords).reshape(-1, 7)[:int(cnt[1])].copy()))
    a, b = out
    assert np.array_equal(a["per_ray"], b["per_ray"]) and a["total"] == b["total"]
    assert (a["per_ray"] == 0).any(), "the test scene must have rays that miss everything"
    assert b["kept"] == n_rays and a["kept"] == int((a["per_ray"] > 0).sum())
    assert np.array_equal(b["ns"][:, 0], b["per_ray"]), "with the flag the ray list is the launch order, empty rays included"
    full = b["ns"][:, 0] > 0
    assert np.array_equal(b["ns"][full], a["ns"]) and np.array_equal(b["rays"][full], a["rays"]) and np.array_equal(b["rgba"][full], a["rgba"])
    assert np.array_equal(a["coords"], b["coords"])
    assert np.all(np.abs(np.linalg.norm(b["rays"][~full][:, 3:], axis=1) - 1) < 1e-5), "empty rays still carry their direction"


def test_coarse_bitfield_is_the_or_over_blocks_of_512_cells(capi, orc):
    import gpuutil as G
    import torch
    lib = capi.load()
    cfg = capi.make_sampler_cfg(4)                      # 3 cascades
    g = np.random.Generator(np.random.PCG64(4))
    n_bytes = 128 ** 3 // 8 * 3
    bits = np.zeros(n_bytes, np.uint8)
    hot = g.integers(0, n_bytes, 700)
    bits[hot] = (1 << g.integers(0, 8, 700)).astype(np.uint8)
    assert lib.ngp_coarse_bitfield_bytes(C.byref(cfg)) == n_bytes // 512
    d_coarse = G.dev(np.full(n_bytes // 512, 0xAA, np.uint8))
    capi.check(lib.ngp_build_coarse_bitfield(G.stream(), C.byref(cfg), G.p(G.dev(bits)), G.p(d_coarse)))
    G.sync()
    want = np.packbits(bits.reshape(-1, 64).any(axis=1), bitorder="little")
    assert np.array_equal(G.host(d_coarse, np.uint8), want) and want.any() and not want.all()
    bad = capi.make_sampler_cfg(1, grid_size=16)
    assert lib.ngp_build_coarse_bitfield(G.stream(), C.byref(bad), G.p(d_coarse), G.p(d_coarse)) != 0


@pytest.mark.parametrize("coarse", [False, True])
@pytest.mark.parametrize("spp,n_steps", [(0, 1), (3, 8), (1, 64)])
def test_inference_rays_and_coords_match_oracle(capi, orc, spp, n_steps, coarse):
    """coarse: with ngp_sampler_cfg.coarse_bitfield set the march crosses whole empty 8x8x8 blocks in one hop where it can prove the
    landing point -- payloads and samples must still be bit-exact against the oracle's cell-by-cell march (the test grid is
    blocky at exactly that granularity and 92 % empty, so the block hops carry most of the march)"""
    import gpuutil as G
    import torch
    lib = capi.load()
    cfg, ocfg = capi.make_sampler_cfg(1), orc.SamplerCfg(1)
    W, H, fx = 96, 64, 80.0
    xf, _, bits = _scene(orc)
    cam_m = xf[1]
    ref_p = orc.init_rays(ocfg, W, H, fx, fx, cam_m, spp, bits)
    cam = _cam(capi, W, H, fx)
    payloads = G.zeros(W * H * 10, torch.int32)
    d_bits = G.dev(bits)
    if coarse:
        d_coarse = G.zeros(lib.ngp_coarse_bitfield_bytes(C.byref(cfg)), torch.uint8)
        capi.check(lib.ngp_build_coarse_bitfield(G.stream(), C.byref(cfg), G.p(d_bits), G.p(d_coarse)))
        cfg.coarse_bitfield = d_coarse.data_ptr()
    cm = (C.c_float * 12)(*cam_m.tolist())
    capi.check(lib.ngp_init_rays(G.stream(), C.byref(cfg), C.byref(cam), cm, spp, 1, G.p(d_bits), G.p(payloads)))
    G.sync()
    got_p = G.host(payloads, np.uint32).reshape(-1, 10)
    alive = ref_p[:, 0] != 0
    assert np.array_equal(got_p[:, 0] != 0, alive)
    assert alive.any() and not alive.all()
    assert np.array_equal(got_p[alive], ref_p[alive]), "payloads of live rays must be bit-exact"
    # per-iteration sample generation on the same payloads
    ref_coords = orc.generate_coords_from_rays(ocfg, ref_p, n_steps, bits)
    coords = G.zeros(W * H * n_steps * 7, torch.float32)
    capi.check(lib.ngp_generate_coords_from_rays(G.stream(), C.byref(cfg), W * H, None, n_steps, G.p(d_bits), G.p(payloads), G.p(coords)))
    G.sync()
    got_p2 = G.host(payloads, np.uint32).reshape(-1, 10)
    assert np.array_equal(got_p2[alive], ref_p[alive]), "n_steps / t after generate_coords must be bit-exact"
    gc = G.host(coords).reshape(W * H, n_steps, 7)
    rc = ref_coords.reshape(W * H, n_steps, 7)
    ns = ref_p[:, 9]
    for j in range(n_steps):
        m = alive & (ns > j)
        assert np.array_equal(gc[m, j], rc[m, j])
