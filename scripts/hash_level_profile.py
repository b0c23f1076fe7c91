#!/usr/bin/env python
"""Per-level cost of the hash-grid scatter (ngp_hash_encode_bwd_levels) on ray-ordered samples: which levels the backward
kernel's time goes to.  usage: python scripts/hash_level_profile.py [n_rays=8192] [samples_per_ray=32]"""
import ctypes as C
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("instant-ngp-nerf-minimal_b200")
capi = pkg.capi


def main():
    import torch
    lib = capi.load()
    n_rays = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    spr = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    g = np.random.Generator(np.random.PCG64(1))
    o = g.random((n_rays, 3), dtype=np.float32) * 0.6 + 0.2
    d = g.normal(size=(n_rays, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    t = (np.arange(spr, dtype=np.float32) * np.float32(np.sqrt(3) / 1024))[None, :, None]
    pos = np.clip(o[:, None, :] + d[:, None, :] * t, 0, 1).reshape(-1, 3)
    n = pos.shape[0]
    coords = np.zeros((n, 7), np.float32)
    coords[:, :3] = pos
    dy = (g.normal(size=(n, 32)) * 2 ** -7).astype(np.float16)
    desc, n_entries = capi.make_grid_desc()
    dc, ddy = torch.from_numpy(coords).cuda(), torch.from_numpy(dy.view(np.int16)).cuda()
    grad = torch.zeros(n_entries * 2, dtype=torch.int16, device="cuda")
    table = torch.from_numpy((g.random(n_entries * 2, dtype=np.float32) * 2e-4 - 1e-4).astype(np.float16).view(np.int16)).cuda()
    out = torch.zeros(n * 32, dtype=torch.int16, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    vp = lambda x: C.c_void_p(x.data_ptr())

    def timed(fn, reps=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps * 1e3

    print(f"n = {n} samples ({n_rays} rays x {spr})")
    full = timed(lambda: capi.check(lib.ngp_hash_encode_bwd(s, vp(dc), 7, n, vp(ddy), 32, C.byref(desc), vp(grad), 1)))
    print(f"bwd all levels + memset: {full:7.1f} us")
    tot = 0.0
    for l in range(16):
        us = timed(lambda: capi.check(lib.ngp_hash_encode_bwd_levels(s, vp(dc), 7, n, vp(ddy), 32, C.byref(desc), vp(grad), l, l + 1)))
        tot += us
        print(f"  level {l:2d} res {desc.resolutions[l]:5d} entries {desc.offsets[l + 1] - desc.offsets[l]:7d}: {us:6.1f} us")
    print(f"  sum of single-level launches: {tot:7.1f} us")
    for a, b in ((0, 4), (4, 8), (8, 12), (12, 16), (0, 8), (8, 16)):
        us = timed(lambda: capi.check(lib.ngp_hash_encode_bwd_levels(s, vp(dc), 7, n, vp(ddy), 32, C.byref(desc), vp(grad), a, b)))
        print(f"  levels [{a:2d},{b:2d}): {us:6.1f} us")
    fwd = timed(lambda: capi.check(lib.ngp_hash_encode_fwd(s, vp(dc), 7, n, vp(table), C.byref(desc), vp(out), 32, None)))
    print(f"fwd all levels: {fwd:7.1f} us")


if __name__ == "__main__":
    main()
