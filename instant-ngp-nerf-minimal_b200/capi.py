"""ctypes binding of the C ABI declared in include/ngp_b200.h (libngp_b200.so).

This is the reference-side binding a Python maintainer would use; the C++ Testbed / pyngp module links the same library.
Every compute entry point takes raw device pointers (torch tensors are only the allocator here) and fails loudly when the
CUDA library is missing -- there is no CPU fallback.
"""
import ctypes as C
import os

_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_DIR, "libngp_b200.so")

NGP_MAX_LEVELS = 32
NGP_MLP_PARAMS = 10240

u32, i32, f32, u64, vp, sz = C.c_uint32, C.c_int32, C.c_float, C.c_uint64, C.c_void_p, C.c_size_t


class GridDesc(C.Structure):
    _fields_ = [("n_levels", u32), ("offsets", u32 * (NGP_MAX_LEVELS + 1)), ("scales", f32 * NGP_MAX_LEVELS), ("resolutions", u32 * NGP_MAX_LEVELS)]


class SamplerCfg(C.Structure):
    _fields_ = [("aabb_min", f32), ("aabb_max", f32), ("grid_size", u32), ("n_cascades", u32), ("max_steps", u32),
                ("cone_angle", f32), ("min_step", f32), ("max_step", f32), ("near_distance", f32), ("keep_empty_rays", u32), ("coarse_bitfield", vp)]


class CameraCfg(C.Structure):
    _fields_ = [("width", i32), ("height", i32), ("fx", f32), ("fy", f32), ("ppx", f32), ("ppy", f32)]


class RenderCfg(C.Structure):
    _fields_ = [("n_cascades", u32), ("min_step", f32), ("rgb_activation", i32), ("density_activation", i32), ("loss_type", i32),
                ("color_space", i32), ("train_with_random_bg_color", i32), ("train_in_linear_color", i32),
                ("transmittance_threshold", f32), ("min_optical_thickness", f32), ("loss_scale", f32), ("background_color", f32 * 3),
                ("envmap", vp), ("envmap_gradient", vp), ("rays", vp), ("envmap_res_x", i32), ("envmap_res_y", i32), ("envmap_loss_type", i32), ("reg_scale", f32)]


NGP_MAX_PEERS = 8


class PeerSet(C.Structure):
    _fields_ = [("world", u32), ("rank", u32), ("grads", vp * NGP_MAX_PEERS), ("params", vp * NGP_MAX_PEERS), ("flags", vp * NGP_MAX_PEERS)]


class OptimizerCfg(C.Structure):
    _fields_ = [("learning_rate", f32), ("beta1", f32), ("beta2", f32), ("epsilon", f32), ("l2_reg", f32), ("loss_scale", f32), ("ema_decay", f32)]


# every exported symbol of include/ngp_b200.h with its signature (restype, argtypes)
SIGNATURES = {
    "ngp_last_error": (C.c_char_p, []),
    "ngp_abi_version": (i32, []),
    "ngp_launch_count": (u64, []),
    "ngp_kernel_timing_enable": (i32, [i32, u32]),
    "ngp_kernel_timing_collect": (i32, [u32, vp, vp, vp]),
    "ngp_kernel_timing_timeline": (i32, [u32, vp, vp, vp]),
    "ngp_peer_stamps_read": (i32, [vp]),
    "ngp_l2_probe": (i32, [vp, i32, vp, u64, u32, vp]),
    "ngp_grid_desc_init": (u32, [C.POINTER(GridDesc), u32, u32, u32, f32]),
    "ngp_auto_per_level_scale": (f32, [f32, u32, u32]),
    "ngp_grid_desc_device_scales": (i32, [C.POINTER(GridDesc), u32, f32, vp, vp]),
    "ngp_hash_encode_fwd": (i32, [vp, vp, u32, u32, vp, C.POINTER(GridDesc), vp, u32, vp]),
    "ngp_hash_encode_bwd": (i32, [vp, vp, u32, u32, vp, u32, C.POINTER(GridDesc), vp, i32]),
    "ngp_network_workspace_bytes": (sz, [u32]),
    "ngp_network_inference": (i32, [vp, vp, u32, vp, vp, C.POINTER(GridDesc), vp, vp, sz]),
    "ngp_network_density": (i32, [vp, vp, u32, u32, vp, C.POINTER(GridDesc), vp, vp, sz]),
    "ngp_network_input_gradient": (i32, [vp, vp, u32, vp, vp, C.POINTER(GridDesc), vp, sz]),
    "ngp_network_train": (i32, [vp, vp, u32, vp, C.POINTER(GridDesc), vp, vp, vp, vp, sz]),
    "ngp_network_train_reuse": (i32, [vp, vp, u32, vp, C.POINTER(GridDesc), vp, vp, vp, vp, u32, vp, vp, sz]),
    "ngp_mlp_inference": (i32, [vp, vp, vp, u32, vp, vp]),
    "ngp_mlp_train": (i32, [vp, vp, vp, u32, vp, vp, vp, vp, vp]),
    "ngp_mark_untrained_grid": (i32, [vp, C.POINTER(SamplerCfg), C.POINTER(CameraCfg), u32, vp, vp]),
    "ngp_generate_grid_samples": (i32, [vp, C.POINTER(SamplerCfg), u32, u64, u64, u32, vp, f32, vp, vp]),
    "ngp_generate_grid_samples_sorted": (i32, [vp, C.POINTER(SamplerCfg), u32, u64, u64, u32, vp, f32, vp, vp]),
    "ngp_generate_grid_samples_sorted_part": (i32, [vp, C.POINTER(SamplerCfg), u32, u64, u64, u32, vp, f32, vp, vp, u32, u32, C.POINTER(u32), C.POINTER(u32)]),
    "ngp_splat_grid": (i32, [vp, C.POINTER(SamplerCfg), u32, vp, vp, u32, i32, vp]),
    "ngp_ema_grid": (i32, [vp, C.POINTER(SamplerCfg), f32, vp, vp]),
    "ngp_update_grid": (i32, [vp, C.POINTER(SamplerCfg), u32, vp, vp, u32, i32, f32, vp, vp]),
    "ngp_coarse_bitfield_bytes": (C.c_size_t, [C.POINTER(SamplerCfg)]),
    "ngp_build_coarse_bitfield": (i32, [vp, C.POINTER(SamplerCfg), vp, vp]),
    "ngp_update_grid_bitfield": (i32, [vp, C.POINTER(SamplerCfg), vp, f32, vp, vp]),
    "ngp_scan_workspace_bytes": (sz, [u32]),
    "ngp_generate_training_samples": (i32, [vp, C.POINTER(SamplerCfg), C.POINTER(CameraCfg), u32, u32, u32, u32, vp, vp, vp, i32, u64, u64,
                                            vp, vp, vp, vp, vp, vp, vp]),
    "ngp_init_rays": (i32, [vp, C.POINTER(SamplerCfg), C.POINTER(CameraCfg), C.POINTER(f32), u32, i32, vp, vp]),
    "ngp_init_rays_rows": (i32, [vp, C.POINTER(SamplerCfg), C.POINTER(CameraCfg), C.POINTER(f32), u32, i32, vp, vp, u32, u32]),
    "ngp_init_rays_rows_envmap": (i32, [vp, C.POINTER(SamplerCfg), C.POINTER(CameraCfg), C.POINTER(f32), u32, i32, vp, vp, u32, u32, vp, i32, i32, vp]),
    "ngp_generate_coords_from_rays": (i32, [vp, C.POINTER(SamplerCfg), u32, vp, u32, vp, vp, vp]),
    "ngp_composite_workspace_bytes": (sz, [u32]),
    "ngp_composite_train": (i32, [vp, C.POINTER(RenderCfg), u32, u32, vp, vp, vp, vp, vp, vp, vp, vp, u32, vp, u64, u64, vp, vp, vp]),
    "ngp_composite_train_static": (i32, [vp, C.POINTER(RenderCfg), u32, u32, vp, vp, vp, vp, vp, vp, vp, vp, u32, vp, u64, u64, vp, vp]),
    "ngp_rollover": (i32, [vp, u32, vp, vp, u32, vp, vp]),
    "ngp_composite_infer": (i32, [vp, C.POINTER(RenderCfg), u32, vp, u32, u32, vp, vp, vp, vp, u32, i32, f32, f32, C.POINTER(f32), f32]),
    "ngp_composite_infer_dev": (i32, [vp, C.POINTER(RenderCfg), u32, vp, u32, vp, vp, vp, vp, u32, i32, f32, f32, C.POINTER(f32), f32]),
    "ngp_composite_infer_compact_dev": (i32, [vp, C.POINTER(RenderCfg), u32, vp, u32, vp, vp, vp, vp, u32, i32, f32, f32, C.POINTER(f32), f32, vp, vp, vp, vp, vp]),
    "ngp_render_round": (i32, [vp, vp, u32, u32, u32, u32, u32, u32, u32]),
    "ngp_generate_coords_from_rays_dev": (i32, [vp, C.POINTER(SamplerCfg), u32, vp, u32, vp, vp, vp]),
    "ngp_compact_rays": (i32, [vp, u32, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "ngp_shade": (i32, [vp, u32, vp, vp, vp, vp, i32, i32]),
    "ngp_accumulate": (i32, [vp, u32, vp, vp, f32, i32]),
    "ngp_tonemap": (i32, [vp, u32, f32, C.POINTER(f32), vp, i32, i32, i32, vp]),
    "ngp_init_params": (i32, [vp, C.POINTER(u64), u64, u64, vp, vp]),
    "ngp_optimizer_step": (i32, [vp, C.POINTER(OptimizerCfg), u64, u64, u32, f32, vp, vp, vp, vp, vp, vp, vp]),
    "ngp_peer_publish_gradients": (i32, [vp, C.POINTER(PeerSet), u32]),
    "ngp_optimizer_step_sharded": (i32, [vp, C.POINTER(OptimizerCfg), u64, u64, u64, u32, f32, vp, vp, vp, vp, vp, C.POINTER(PeerSet), u32, vp]),
    "ngp_peer_wait_weights": (i32, [vp, C.POINTER(PeerSet), u32]),
    "ngp_from_rgba32": (i32, [vp, u64, vp, vp]),
    "ngp_reduce_sum": (i32, [vp, vp, u64, vp]),
    "ngp_render_procedural_views": (i32, [vp, u32, u32, i32, i32, f32, f32, vp, vp]),
}

_lib = None


class NgpError(RuntimeError):
    pass


def load():
    """Load libngp_b200.so; raises if it has not been built (no fallback path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NgpError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (nvcc, sm_100a)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)          # AttributeError here = header/library mismatch
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc):
    if rc != 0:
        raise NgpError(load().ngp_last_error().decode())


def ptr(t):
    """device pointer of a torch tensor (or None)"""
    return None if t is None else C.c_void_p(t.data_ptr())


def make_grid_desc(n_levels=16, log2_hashmap_size=19, base_resolution=16, per_level_scale=None, desired_resolution=2048.0, device_scales=False):
    """device_scales: per-level scales as the reference's kernels evaluate them on the GPU (needs a device); the default host
    values are what the CPU oracle uses"""
    lib = load()
    if per_level_scale is None:
        per_level_scale = lib.ngp_auto_per_level_scale(desired_resolution, base_resolution, n_levels)
    d = GridDesc()
    n_entries = lib.ngp_grid_desc_init(C.byref(d), n_levels, log2_hashmap_size, base_resolution, per_level_scale)
    if n_entries == 0:
        raise NgpError(lib.ngp_last_error().decode())
    if device_scales:
        import torch                                     # device memory only: the scratch the set-up call writes its 32 scales to
        scratch = torch.zeros(32, dtype=torch.float32, device="cuda")
        check(lib.ngp_grid_desc_device_scales(C.byref(d), base_resolution, per_level_scale, None, C.c_void_p(scratch.data_ptr())))
    return d, n_entries


def make_sampler_cfg(aabb_scale=1, grid_size=128, max_steps=1024, cone_angle=0.0, near_distance=0.2):
    """OccupancySampler ctor arithmetic (occupancy_sampler.cu:603-631), float32 throughout"""
    import numpy as np
    n_cascades = 2
    while (1 << (n_cascades - 1)) < aabb_scale:
        n_cascades += 1
    sqrt3 = np.float32(1.73205080757)
    c = SamplerCfg()
    c.aabb_min = float(np.float32(0.5) - np.float32(0.5) * np.float32(aabb_scale))
    c.aabb_max = float(np.float32(0.5) + np.float32(0.5) * np.float32(aabb_scale))
    c.grid_size, c.n_cascades, c.max_steps = grid_size, n_cascades, max_steps
    c.cone_angle = cone_angle
    c.min_step = float(sqrt3 / np.float32(max_steps))
    c.max_step = float(sqrt3 * np.float32(1 << (n_cascades - 1)) / np.float32(grid_size))
    c.near_distance = near_distance
    return c


def kernel_timing_timeline(max_launches=8192):
    """[(kernel name, start ms, end ms)] per recorded launch, relative to the first one; synchronises the device"""
    import numpy as np
    lib = load()
    names = C.create_string_buffer(max_launches * 64)
    a = np.zeros(max_launches, np.float32)
    b = np.zeros(max_launches, np.float32)
    k = lib.ngp_kernel_timing_timeline(max_launches, names, a.ctypes.data_as(vp), b.ctypes.data_as(vp))
    if k < 0:
        raise NgpError(lib.ngp_last_error().decode())
    return [(names.raw[i * 64:(i + 1) * 64].split(b"\0")[0].decode(), float(a[i]), float(b[i])) for i in range(k)]


def kernel_timing_collect(max_kernels=64):
    """{kernel name: (total ms, launches)} recorded since ngp_kernel_timing_enable(1, n); synchronises the device"""
    import numpy as np
    lib = load()
    names = C.create_string_buffer(max_kernels * 64)
    ms = np.zeros(max_kernels, np.float32)
    cnt = np.zeros(max_kernels, np.uint32)
    k = lib.ngp_kernel_timing_collect(max_kernels, names, ms.ctypes.data_as(vp), cnt.ctypes.data_as(vp))
    if k < 0:
        raise NgpError(lib.ngp_last_error().decode())
    return {names.raw[i * 64:(i + 1) * 64].split(b"\0")[0].decode(): (float(ms[i]), int(cnt[i])) for i in range(k)}
