"""helpers shared by the tests: seeded inputs drawn from the oracle's pcg32 so CPU and GPU see identical bits"""
import numpy as np


def pcg_floats(n, seed):
    """n floats in [0,1) from pcg32(seed), vectorised restatement of next_float for test inputs"""
    mult, mask = 0x5851F42D4C957F2D, (1 << 64) - 1
    inc = (1 << 1) | 1
    state = 0
    state = (state * mult + inc) & mask
    state = (state + seed) & mask
    state = (state * mult + inc) & mask
    out = np.empty(n, np.float32)
    for i in range(n):
        old = state
        state = (old * mult + inc) & mask
        xs = (((old >> 18) ^ old) >> 27) & 0xFFFFFFFF
        rot = old >> 59
        r = ((xs >> rot) | (xs << ((-rot) & 31))) & 0xFFFFFFFF
        out[i] = np.uint32((r >> 9) | 0x3F800000).view(np.float32) - np.float32(1.0)
    return out


def rng_uniform(shape, seed, lo=0.0, hi=1.0, dtype=np.float32):
    """fast seeded uniform inputs (numpy PCG64) -- used where only CPU/GPU agreement matters"""
    g = np.random.Generator(np.random.PCG64(seed))
    return (g.random(shape, dtype=np.float32) * np.float32(hi - lo) + np.float32(lo)).astype(dtype)


def half_ulp_diff(a, b):
    """distance in fp16 representable steps between two float16 arrays (sign-magnitude aware)"""
    ia = a.view(np.int16).astype(np.int32)
    ib = b.view(np.int16).astype(np.int32)
    ia = np.where(ia < 0, -(ia & 0x7FFF), ia)
    ib = np.where(ib < 0, -(ib & 0x7FFF), ib)
    return np.abs(ia - ib)
