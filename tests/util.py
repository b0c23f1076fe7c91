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


def make_cameras(n, radius=1.3, seed=5):
    """n camera-to-world matrices (3x4, column-major flattened to 12 floats) in the unit-cube frame, looking at (0.5,0.5,0.5)"""
    g = np.random.Generator(np.random.PCG64(seed))
    out = np.zeros((n, 12), np.float32)
    for i in range(n):
        v = g.normal(size=3)
        v /= np.linalg.norm(v)
        pos = 0.5 + radius * v
        fwd = (0.5 - pos) / np.linalg.norm(0.5 - pos)
        up = np.array([0.0, 0.0, 1.0]) if abs(fwd[2]) < 0.9 else np.array([0.0, 1.0, 0.0])
        right = np.cross(fwd, up)
        right /= np.linalg.norm(right)
        down = np.cross(fwd, right)
        m = np.stack([right, down, fwd, pos], axis=1)       # 3x4, camera looks along +z (OpenCV-style, as after spec_opengl_to_opencv)
        out[i] = m.T.reshape(-1)                             # column-major
    return out


def make_density_grid(n_cascades=2, G=128, seed=9, fill=0.08):
    """float occupancy grid [n_cascades*G^3] in Morton order: a few blobs above threshold, some -1 (never seen) cells"""
    g = np.random.Generator(np.random.PCG64(seed))
    grid = np.zeros(n_cascades * G ** 3, np.float32)
    coarse = g.random((n_cascades, 16, 16, 16)) < fill           # blocky occupancy at 8-cell granularity
    idx = np.arange(G ** 3, dtype=np.uint32)

    def compact(v):
        v = v & 0x49249249
        v = (v | (v >> 2)) & 0xC30C30C3
        v = (v | (v >> 4)) & 0x0F00F00F
        v = (v | (v >> 8)) & 0xFF0000FF
        v = (v | (v >> 16)) & 0x0000FFFF
        return v
    x, y, z = compact(idx), compact(idx >> 1), compact(idx >> 2)
    for c in range(n_cascades):
        occ = coarse[c][x // 8, y // 8, z // 8]
        vals = np.where(occ, 0.02 + g.random(G ** 3, dtype=np.float32), g.random(G ** 3, dtype=np.float32) * 0.005).astype(np.float32)
        vals[g.random(G ** 3) < 0.01] = -1.0
        grid[c * G ** 3:(c + 1) * G ** 3] = vals
    return grid
