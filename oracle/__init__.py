"""TEST INFRASTRUCTURE ONLY -- ctypes front end of the CPU oracle (oracle/*.cpp).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
Parity unpinned: see oracle/oracle_common.h.
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
