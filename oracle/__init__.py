"""TEST INFRASTRUCTURE ONLY -- ctypes front end of the CPU oracle (oracle/*.cpp).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
Pinned to golden vectors from the reference's own CUDA build (tests/golden/, see oracle/oracle_common.h).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_DIR, "libngp_oracle.so")


def build(force=False):
    srcs = [os.path.join(_DIR, f) for f in os.listdir(_DIR) if f.endswith((".cpp", ".h", ".inl"))]
    if force or not os.path.exists(_LIB) or any(os.path.getmtime(s) > os.path.getmtime(_LIB) for s in srcs):
        subprocess.check_call(["make", "-C", _DIR, "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB)
        _lib.orc_grid_levels.restype = C.c_uint32
        _lib.orc_grid_index.restype = C.c_uint32
        _lib.orc_auto_per_level_scale.restype = C.c_float
        _lib.orc_grid_mean.restype = C.c_float
        _lib.orc_composite_train.restype = C.c_uint32
        _lib.orc_ld_random_val.restype = C.c_float
        _lib.orc_pcg32_next_uint.restype = C.c_uint32
        _lib.orc_pcg32_next_float.restype = C.c_float
    return _lib


def P(a):
    """pointer to a C-contiguous numpy array (or None)"""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"], "oracle arrays must be contiguous"
    return a.ctypes.data_as(C.c_void_p)


u32, f32, u64, i32 = C.c_uint32, C.c_float, C.c_uint64, C.c_int


class Pcg32:
    """pcg32 stream driven through the oracle (tcnn pcg32.h)."""

    def __init__(self, initstate=None, initseq=1, state=None, inc=None):
        if state is not None:
            self.state, self.inc = int(state), int(inc)
        else:
            s, i = u64(), u64()
            lib().orc_pcg32_seed(u64(initstate), u64(initseq), C.byref(s), C.byref(i))
            self.state, self.inc = s.value, i.value

    def copy(self):
        return Pcg32(state=self.state, inc=self.inc)

    def advance(self, delta=1 << 32):
        s = u64(self.state)
        lib().orc_pcg32_advance(C.byref(s), u64(self.inc), C.c_int64(delta))
        self.state = s.value

    def next_uint(self):
        s = u64(self.state)
        v = lib().orc_pcg32_next_uint(C.byref(s), u64(self.inc))
        self.state = s.value
        return v

    def next_float(self):
        s = u64(self.state)
        v = lib().orc_pcg32_next_float(C.byref(s), u64(self.inc))
        self.state = s.value
        return v


def trainer_rng(seed=43):
    s, i = u64(), u64()
    lib().orc_trainer_rng(u32(seed), C.byref(s), C.byref(i))
    return Pcg32(state=s.value, inc=i.value)


class GridLevels:
    def __init__(self, n_levels=16, log2_hashmap_size=19, base_resolution=16, per_level_scale=None, desired_resolution=2048.0):
        if per_level_scale is None:
            # ngp_network.h:623-632 (float math)
            per_level_scale = lib().orc_auto_per_level_scale(f32(desired_resolution), u32(base_resolution), u32(n_levels)) if n_levels > 1 else 1.0
        self.n_levels = n_levels
        self.per_level_scale = np.float32(per_level_scale)
        self.offsets = np.zeros(n_levels + 1, np.uint32)
        self.scales = np.zeros(n_levels, np.float32)
        self.resolutions = np.zeros(n_levels, np.uint32)
        self.n_entries = lib().orc_grid_levels(u32(n_levels), u32(log2_hashmap_size), u32(base_resolution), f32(self.per_level_scale),
                                               P(self.offsets), P(self.scales), P(self.resolutions))
        self.n_params = int(self.n_entries) * 2


def hash_encode_fwd(xyz, stride, n, table, L, want_indices=False):
    out = np.zeros((n, 2 * L.n_levels), np.uint16)
    idx = np.zeros((n, L.n_levels, 8), np.uint32) if want_indices else None
    lib().orc_hash_encode_fwd(P(xyz), u32(stride), u32(n), P(table), u32(L.n_levels), P(L.offsets), P(L.scales), P(L.resolutions), P(out), P(idx))
    return (out.view(np.float16), idx) if want_indices else out.view(np.float16)


def hash_encode_bwd(xyz, stride, n, dy, L):
    grad = np.zeros(L.n_params, np.float32)
    lib().orc_hash_encode_bwd(P(xyz), u32(stride), u32(n), P(dy.view(np.uint16)), u32(L.n_levels), P(L.offsets), P(L.scales), P(L.resolutions), P(grad))
    return grad


def network_fwd(weights, enc, coords, n, mode=0):
    out = np.zeros((n, 4), np.uint16)
    lib().orc_network_fwd(P(weights.view(np.uint16)), P(enc.view(np.uint16)), P(coords), u32(coords.shape[1]), u32(n), i32(mode), P(out))
    return out.view(np.float16)


def density_fwd(weights, enc, n, mode=0):
    out = np.zeros((n, 16), np.uint16)
    lib().orc_density_fwd(P(weights.view(np.uint16)), P(enc.view(np.uint16)), u32(n), i32(mode), P(out))
    return out.view(np.float16)


def network_fwd_bwd(weights, enc, coords, dL, n, mode=0):
    wgrad = np.zeros(10240, np.float32)
    denc = np.zeros((n, 32), np.uint16)
    out = np.zeros((n, 4), np.uint16)
    lib().orc_network_fwd_bwd(P(weights.view(np.uint16)), P(enc.view(np.uint16)), P(coords), u32(coords.shape[1]), P(dL.view(np.uint16)),
                              u32(n), i32(mode), P(wgrad), P(denc), P(out))
    return wgrad, denc.view(np.float16), out.view(np.float16)


def sh4(dirs):
    n = dirs.shape[0]
    out = np.zeros((n, 16), np.uint16)
    lib().orc_sh4(P(dirs), u32(dirs.shape[1]), u32(n), P(out))
    return out.view(np.float16)


# ---- sampler ------------------------------------------------------------------------------------------------

class SamplerCfg:
    """OccupancySampler ctor arithmetic (occupancy_sampler.cu:603-631) in float32"""

    def __init__(self, aabb_scale=1, grid_size=128, max_steps=1024, cone_angle=0.0, near_distance=0.2):
        n_cascades = 2
        while (1 << (n_cascades - 1)) < aabb_scale:
            n_cascades += 1
        sqrt3 = np.float32(1.73205080757)
        self.aabb_min = np.float32(0.5) - np.float32(0.5) * np.float32(aabb_scale)
        self.aabb_max = np.float32(0.5) + np.float32(0.5) * np.float32(aabb_scale)
        self.grid_size, self.n_cascades, self.max_steps = grid_size, n_cascades, max_steps
        self.cone_angle = np.float32(cone_angle)
        self.min_step = sqrt3 / np.float32(max_steps)
        self.max_step = sqrt3 * np.float32(1 << (n_cascades - 1)) / np.float32(grid_size)
        self.near_distance = np.float32(near_distance)
        self.cells = grid_size ** 3

    def f(self):
        return np.array([self.aabb_min, self.aabb_max, self.cone_angle, self.min_step, self.max_step, self.near_distance], np.float32)

    def u(self):
        return np.array([self.grid_size, self.n_cascades, self.max_steps], np.uint32)


def mark_untrained_grid(cfg, xforms, W, H, fx, fy):
    grid = np.zeros(cfg.cells * cfg.n_cascades, np.float32)
    xf = np.ascontiguousarray(xforms, np.float32)
    lib().orc_mark_untrained_grid(u32(cfg.n_cascades), u32(cfg.grid_size), u32(xf.shape[0]), i32(W), i32(H), f32(fx), f32(fy), P(xf), P(grid))
    return grid


def generate_grid_samples(cfg, n, rng, step, grid, thresh):
    pos = np.zeros((n, 3), np.float32)
    idx = np.zeros(n, np.uint32)
    lib().orc_generate_grid_samples(u32(n), u64(rng.state), u64(rng.inc), u32(step), f32(cfg.aabb_min), f32(cfg.aabb_max), P(grid),
                                    u32(cfg.grid_size), u32(cfg.n_cascades), f32(thresh), P(pos), P(idx))
    return pos, idx


def update_grid(cfg, indices, density_out, stride, decay, grid):
    """grid is updated in place; returns grid_current"""
    cur = np.zeros_like(grid)
    lib().orc_update_grid(u32(len(indices)), P(indices), P(density_out.view(np.uint16)), u32(stride), f32(cfg.min_step), f32(decay), u32(grid.size), P(cur), P(grid))
    return cur


def grid_mean(cfg, grid):
    return float(lib().orc_grid_mean(P(grid), u32(cfg.cells)))


def grid_to_bitfield(cfg, grid, mean, min_optical_thickness=0.01):
    bits = np.zeros(cfg.cells * cfg.n_cascades // 8, np.uint8)
    lib().orc_grid_to_bitfield(P(grid), u32(cfg.grid_size), u32(cfg.n_cascades), f32(mean), f32(min_optical_thickness), P(bits))
    return bits


def generate_training_samples(cfg, n_rays, n_rays_shift, n_coords, xforms, images, W, H, fx, fy, bits, rng, snap=True, ppx=0.5, ppy=0.5):
    """images: uint16 (fp16 bits) [n_images][H][W][4].  Returns dict of the reference kernel's outputs."""
    xf = np.ascontiguousarray(xforms, np.float32)
    rays = np.zeros((n_rays, 6), np.float32)
    rgba = np.zeros((n_rays, 4), np.float32)
    numsteps = np.zeros((n_rays, 2), np.uint32)
    coords = np.zeros((max(n_coords, 1), 7), np.float32)
    counters = np.zeros(2, np.uint32)
    per_ray = np.zeros(n_rays, np.uint32)
    lib().orc_generate_training_samples(u32(n_rays), u32(n_rays_shift), u32(n_coords), u32(xf.shape[0]), i32(W), i32(H), f32(fx), f32(fy), f32(ppx), f32(ppy),
                                        P(xf), P(images), P(cfg.f()), P(cfg.u()), P(bits), i32(1 if snap else 0), u64(rng.state), u64(rng.inc),
                                        P(rays), P(rgba), P(numsteps), P(coords), P(counters), P(per_ray))
    return dict(rays=rays, rgba=rgba, numsteps=numsteps, coords=coords, counters=counters, per_ray_n=per_ray)


def init_rays(cfg, W, H, fx, fy, cam, spp, bits, snap=True, ppx=0.5, ppy=0.5):
    payloads = np.zeros((W * H, 10), np.uint32)
    c = np.ascontiguousarray(cam, np.float32)
    lib().orc_init_rays(i32(W), i32(H), f32(fx), f32(fy), f32(ppx), f32(ppy), P(c), u32(spp), i32(1 if snap else 0), P(cfg.f()), P(cfg.u()), P(bits), P(payloads))
    return payloads


def generate_coords_from_rays(cfg, payloads, n_steps, bits):
    n = payloads.shape[0]
    coords = np.zeros((n * n_steps, 7), np.float32)
    lib().orc_generate_coords_from_rays(u32(n), u32(n_steps), P(cfg.f()), P(cfg.u()), P(bits), P(payloads), P(coords))
    return coords


# ---- render -------------------------------------------------------------------------------------------------

def composite_train(n_rays, n_kept, max_compacted, rgba_gt, numsteps, coords, net_out, cfg, mean_density, rng, min_optical_thickness=0.01,
                    T_thresh=1e-4, loss_scale=128.0, bg=(0, 0, 0), rgb_act=2, density_act=3, loss_type=4, color_space=0, random_bg=1, train_in_linear=0, padded_width=4):
    """numsteps is updated in place (compacted count, compacted base).  Returns dict."""
    cc = np.zeros((max_compacted, 7), np.float32)
    dloss = np.zeros((max_compacted, padded_width), np.uint16)
    loss = np.zeros(n_rays, np.float32)
    rgb_ray = np.zeros((max(n_kept, 1), 3), np.float32)
    cf = np.array([cfg.min_step, mean_density, min_optical_thickness, T_thresh, loss_scale, bg[0], bg[1], bg[2]], np.float32)
    cu = np.array([cfg.n_cascades, rgb_act, density_act, loss_type, color_space, random_bg, train_in_linear], np.uint32)
    counter = lib().orc_composite_train(u32(n_rays), u32(n_kept), u32(max_compacted), P(rgba_gt), P(numsteps), P(coords), P(cc), P(net_out.view(np.uint16)),
                                        P(dloss), u32(padded_width), P(cf), P(cu), u64(rng.state), u64(rng.inc), P(loss), P(rgb_ray))
    return dict(counter=int(counter), coords_compacted=cc, dloss=dloss.view(np.float16), loss=loss, rgb_ray=rgb_ray)


def composite_train_envmap(n_rays, n_kept, max_compacted, rgba_gt, numsteps, coords, net_out, cfg, mean_density, rng, rays, envmap, env_res_x, env_res_y,
                           env_loss_type=6, min_optical_thickness=0.01, T_thresh=1e-4, loss_scale=128.0, bg=(0, 0, 0), rgb_act=2, density_act=3, loss_type=4,
                           color_space=0, random_bg=1, train_in_linear=0, padded_width=4):
    """composite_train with a trainable environment map behind the volume (ray_marcher.cu:96-102,190-212; utils/envmap.h).
    rays: [n_kept][6] f32 {o, d}; envmap: fp16 [res_y][res_x][4].  Returns the dict of composite_train + envmap_grad (fp32 sums)."""
    import ctypes as C
    cc = np.zeros((max_compacted, 7), np.float32)
    dloss = np.zeros((max_compacted, padded_width), np.uint16)
    loss = np.zeros(n_rays, np.float32)
    rgb_ray = np.zeros((max(n_kept, 1), 3), np.float32)
    egrad = np.zeros(env_res_x * env_res_y * 4, np.float32)
    cf = np.array([cfg.min_step, mean_density, min_optical_thickness, T_thresh, loss_scale, bg[0], bg[1], bg[2]], np.float32)
    cu = np.array([cfg.n_cascades, rgb_act, density_act, loss_type, color_space, random_bg, train_in_linear], np.uint32)
    rays = np.ascontiguousarray(rays, np.float32)
    env = np.ascontiguousarray(envmap).view(np.uint16)
    fn = lib().orc_composite_train_envmap
    fn.restype = C.c_uint32
    counter = fn(u32(n_rays), u32(n_kept), u32(max_compacted), P(rgba_gt), P(numsteps), P(coords), P(cc), P(net_out.view(np.uint16)),
                 P(dloss), u32(padded_width), P(cf), P(cu), u64(rng.state), u64(rng.inc), P(loss), P(rgb_ray),
                 P(rays), P(env), P(egrad), C.c_int32(env_res_x), C.c_int32(env_res_y), C.c_int32(env_loss_type))
    return dict(counter=int(counter), coords_compacted=cc, dloss=dloss.view(np.float16), loss=loss, rgb_ray=rgb_ray, envmap_grad=egrad)


def rollover(B, n_in, grads, coords, stride=4):
    g = np.ascontiguousarray(grads).view(np.uint16)
    lib().orc_rollover(u32(B), u32(n_in), P(g), u32(stride), P(coords))
    return g.view(np.float16)


def composite_infer(payloads, rgba, coords, net_out, n_steps, n_steps_offset, cfg, render_mode=1, T_thresh=1e-4, depth_scale=1.0, cam_fwd=(0, 0, 1),
                    rgb_act=2, density_act=3, padded_width=4):
    cf = np.array([cfg.min_step, T_thresh, depth_scale, cfg.aabb_min, cfg.aabb_max, cam_fwd[0], cam_fwd[1], cam_fwd[2]], np.float32)
    cu = np.array([cfg.n_cascades, rgb_act, density_act, render_mode], np.uint32)
    lib().orc_composite_infer(u32(payloads.shape[0]), u32(n_steps), u32(n_steps_offset), P(payloads), P(rgba), P(coords), P(net_out.view(np.uint16)),
                              u32(padded_width), P(cf), P(cu))


def shade(rgba, payloads, frame, n_hit, render_mode=1, train_in_linear=0):
    lib().orc_shade(u32(n_hit), P(rgba), P(payloads), P(frame), i32(render_mode), i32(train_in_linear))


def accumulate(frame, accum, sample_count, color_space):
    lib().orc_accumulate(u32(frame.shape[0]), P(frame), P(accum), f32(sample_count), i32(color_space))


def tonemap(accum, exposure, bg_rgba, color_space, out_color_space):
    out = np.zeros_like(accum)
    b = np.array(bg_rgba, np.float32)
    lib().orc_tonemap(u32(accum.shape[0]), f32(exposure), P(b), P(accum), i32(color_space), i32(out_color_space), P(out))
    return out


# ---- parameters / optimizer ---------------------------------------------------------------------------------

MLP_SHAPES = np.array([[64, 32], [16, 64], [64, 32], [64, 64], [16, 64]], np.uint32)


def init_params(rng, n_grid):
    params = np.zeros(10240 + n_grid, np.float32)
    s = u64(rng.state)
    lib().orc_init_params(C.byref(s), u64(rng.inc), P(MLP_SHAPES), u32(5), u64(n_grid), P(params))
    rng.state = s.value
    return params


class Optimizer:
    """Ema(ExponentialDecay(Adam)) state over an arena of n params, base.json hyper-parameters by default"""

    def __init__(self, params_fp32, n_matrix=10240, lr=1e-2, b1=0.9, b2=0.99, eps=1e-15, l2=1e-6, loss_scale=128.0, ema_decay=0.95,
                 decay_base=0.33, decay_start=20000, decay_interval=10000):
        n = params_fp32.size
        self.n, self.n_matrix = n, n_matrix
        self.w32 = params_fp32.astype(np.float32).copy()
        self.w16 = self.w32.astype(np.float16)
        self.ema16 = self.w16.copy()
        self.m1, self.m2 = np.zeros(n, np.float32), np.zeros(n, np.float32)
        self.psteps = np.zeros(n, np.uint32)
        self.hyper = np.array([lr, b1, b2, eps, l2, loss_scale, ema_decay, decay_base], np.float32)
        self.sched = np.array([decay_start, decay_interval], np.uint32)
        self.step_count = C.c_uint32(0)
        self.lr_factor = C.c_float(1.0)

    def step(self, g16):
        lib().orc_optimizer_step(u64(self.n), u64(self.n_matrix), P(self.hyper), P(self.sched), C.byref(self.step_count), C.byref(self.lr_factor),
                                 P(self.w32), P(self.w16.view(np.uint16)), P(np.ascontiguousarray(g16).view(np.uint16)), P(self.m1), P(self.m2), P(self.psteps),
                                 P(self.ema16.view(np.uint16)))
