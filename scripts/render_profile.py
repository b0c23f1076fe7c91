#!/usr/bin/env python
"""Train briefly, then time Testbed.render at 1920x1080 and 800x800 with the library's per-kernel CUDA-event table.
usage: python scripts/render_profile.py [train_calls=60] [out.json]"""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("instant-ngp-nerf-minimal_b200")
capi, pyngp = pkg.capi, pkg.load_pyngp()
lib = capi.load()


def main():
    calls = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "render_profile.json")
    tb = pyngp.Testbed()
    tb.load_model_config(os.path.join(ROOT, "instant-ngp-nerf-minimal_b200", "configs", "base.json"))
    tb.load_procedural_training_data(100, 800, 800, 2024, 0.33, [0.5, 0.5, 0.5])
    for _ in range(calls):
        tb.train(16, 1 << 18)
    tb.set_camera_to_training_view(3)
    tb.background_color = [0, 0, 0, 1]
    res = {}
    for (w, h) in [(1920, 1080), (800, 800)]:
        tb.render(w, h, 1, True)                                   # warm-up (allocations)
        t0 = time.perf_counter()
        n = 5
        for _ in range(n):
            img = tb.render(w, h, 1, True)
        dt = (time.perf_counter() - t0) / n
        l0 = lib.ngp_launch_count()
        capi.check(lib.ngp_kernel_timing_enable(1, 65536))
        tb.render(w, h, 1, True)
        kt = capi.kernel_timing_collect()
        capi.check(lib.ngp_kernel_timing_enable(0, 0))
        launches = lib.ngp_launch_count() - l0
        res[f"{w}x{h}"] = {"ms_per_frame": dt * 1e3, "mrays_per_s": w * h / dt / 1e6, "launches": int(launches), "alpha_mean": float(img[..., 3].mean()),
                           "kernel_ms": {k: round(v[0], 3) for k, v in sorted(kt.items(), key=lambda kv: -kv[1][0])},
                           "kernel_launches": {k: int(v[1]) for k, v in kt.items()}}
        print(f"{w}x{h}: {dt * 1e3:.2f} ms/frame = {w * h / dt / 1e6:.1f} Mrays/s, {launches} launches, kernel sum {sum(v[0] for v in kt.values()):.2f} ms")
        for k, v in sorted(kt.items(), key=lambda kv: -kv[1][0])[:8]:
            print(f"   {k:40s} {v[0]:8.3f} ms  x{v[1]}")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    json.dump(res, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
