"""Train the procedural scene through pyngp and report loss, throughput and PSNR on a training view.
usage: python scripts/train_eval.py [n_train_calls] [n_views] [res]"""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("instant-ngp-nerf-minimal_b200")
pyngp = pkg.load_pyngp()


def lin2srgb(x):
    x = np.clip(x, 0, 1)
    return np.where(x < 0.0031308, 12.92 * x, 1.055 * np.power(x, 0.41666) - 0.055)


def psnr_view(tb, i, res):
    tb.set_camera_to_training_view(i)
    tb.background_color = [0, 0, 0, 1]
    img = tb.render(res, res, 1, False)
    gt = tb.training_image(i)
    a, b = lin2srgb(img[..., :3]), lin2srgb(gt[..., :3])
    mse = float(np.mean((a - b) ** 2))
    return -10 * np.log10(max(mse, 1e-12)), img, gt


def main():
    n_calls = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    n_views = int(sys.argv[2]) if len(sys.argv) > 2 else 100
    res = int(sys.argv[3]) if len(sys.argv) > 3 else 800
    tb = pyngp.Testbed()
    tb.load_model_config(os.path.join(ROOT, "instant-ngp-nerf-minimal_b200", "configs", "base.json"))
    t0 = time.perf_counter()
    tb.load_procedural_training_data(n_views, res, res, 2024, 0.33, [0.5, 0.5, 0.5])
    print(f"scene: {n_views} views {res}x{res} generated in {time.perf_counter() - t0:.2f}s", flush=True)
    gt = tb.training_image(0)
    print("gt view0: mean alpha %.3f mean rgb %s" % (gt[..., 3].mean(), gt[..., :3].mean((0, 1))))
    tot_samples, tot_ms = 0, 0.0
    for c in range(n_calls):
        t0 = time.perf_counter()
        tb.train(16, 1 << 18)
        wall = (time.perf_counter() - t0) * 1e3
        b = tb.training_buffer
        if c >= 2:
            tot_samples += b.measured_batch_size * 16
            tot_ms += b.training_ms + b.training_prep_ms
        if c < 5 or c % 10 == 9 or c == n_calls - 1:
            print(f"call {c:4d} step {b.i_step:5d} loss {b.loss:.5f} rays {b.n_rays_per_batch:6d} before {b.measured_batch_size_before_compaction:8d} "
                  f"after {b.measured_batch_size:7d} prep {b.training_prep_ms:7.2f} ms train {b.training_ms:7.2f} ms wall {wall:7.2f} ms", flush=True)
    if tot_ms:
        print(f"throughput: {tot_samples / tot_ms * 1e3 / 1e6:.2f} M samples/s (device time, after warm-up)")
    t0 = time.perf_counter()
    p, img, gt = psnr_view(tb, 0, res)
    print(f"PSNR view 0: {p:.2f} dB; render {time.perf_counter() - t0:.3f}s; rendered alpha mean {img[..., 3].mean():.3f}")
    t0 = time.perf_counter()
    img = tb.render(1920, 1080, 1, True)
    dt = time.perf_counter() - t0
    print(f"1920x1080 render: {dt * 1e3:.1f} ms = {1920 * 1080 / dt / 1e6:.1f} Mrays/s")


if __name__ == "__main__":
    main()
