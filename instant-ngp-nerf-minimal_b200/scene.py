"""Benchmark fixture helpers: the camera rig of the procedural nerf_synthetic-shaped scene and the pose conversion the
reference applies when it loads transforms_*.json (internal/utils/common_device.cu:16-30)."""
import numpy as np

CAMERA_ANGLE_X = 0.6911112      # Blender nerf_synthetic camera_angle_x


def _pcg32_floats(seed, n):
    mult, mask = 0x5851F42D4C957F2D, (1 << 64) - 1
    inc, state = 3, 0
    state = (state * mult + inc) & mask
    state = (state + seed) & mask
    state = (state * mult + inc) & mask
    out = []
    for _ in range(n):
        old = state
        state = (old * mult + inc) & mask
        xs = (((old >> 18) ^ old) >> 27) & 0xFFFFFFFF
        rot = old >> 59
        r = ((xs >> rot) | (xs << ((-rot) & 31))) & 0xFFFFFFFF
        out.append(float(np.uint32((r >> 9) | 0x3F800000).view(np.float32) - np.float32(1.0)))
    return out


def sphere_cameras(n, seed=2024, radius=4.0311):
    """OpenGL camera-to-world matrices [n][3][4] on the upper hemisphere, golden-angle spiral (same rig as host/dataset.cpp)"""
    jitter = _pcg32_floats(seed, n)
    out = np.zeros((n, 3, 4), np.float32)
    for i in range(n):
        theta = np.float32(2.39996323) * np.float32(i) + np.float32(0.1) * np.float32(jitter[i])
        phi = np.arcsin(np.float32(0.05) + np.float32(0.9) * (np.float32(i) + np.float32(0.5)) / np.float32(n))
        c = np.array([radius * np.cos(phi) * np.cos(theta), radius * np.cos(phi) * np.sin(theta), radius * np.sin(phi)], np.float64)
        f = -c / np.linalg.norm(c)
        r = np.cross(f, [0, 0, 1.0])
        r /= np.linalg.norm(r)
        u = np.cross(r, f)
        out[i, :, 0], out[i, :, 1], out[i, :, 2], out[i, :, 3] = r, u, -f, c
    return out


def opengl_to_ngp(ogl, scale=0.33, offset=(0.5, 0.5, 0.5)):
    """spec_opengl_to_opencv on [n][3][4] matrices -> [n][12] column-major floats (what the kernels read)"""
    m = np.array(ogl, np.float32).copy()
    m[:, :, 1] *= -1
    m[:, :, 2] *= -1
    m[:, :, 3] = m[:, :, 3] * np.float32(scale) + np.array(offset, np.float32)
    m = m[:, [1, 2, 0], :]
    return np.ascontiguousarray(m.transpose(0, 2, 1).reshape(-1, 12))
