#!/usr/bin/env python
"""PSNR parity after a fixed step count (BASELINE.json configs[1]): train the same procedural scene for the same number of
steps with (a) our sm_100a path through pyngp and (b) the reference's own CUDA build (oracle/_ref/ref_harness), render the same
held-out poses with both, and compare each against the analytic ground truth.  Also reports both render rates.

    python tests/psnr_parity.py [steps=5008] [views=100] [res=800] [n_test=8] [out.json]"""
import ctypes as C
import importlib
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "instant-ngp-nerf-minimal_b200"
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")


def srgb_to_linear(x):
    return np.where(x <= 0.04045, x / 12.92, np.power((x + 0.055) / 1.055, 2.4))


def linear_to_srgb(x):
    x = np.clip(x, 0, 1)
    return np.where(x < 0.0031308, 12.92 * x, 1.055 * np.power(x, 0.41666) - 0.055)


def psnr(a, b):
    return float(-10 * np.log10(max(np.mean((a - b) ** 2), 1e-12)))


def main():
    import torch
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 5008
    views = int(sys.argv[2]) if len(sys.argv) > 2 else 100
    res = int(sys.argv[3]) if len(sys.argv) > 3 else 800
    n_test = int(sys.argv[4]) if len(sys.argv) > 4 else 8
    out_path = sys.argv[5] if len(sys.argv) > 5 else os.path.join(ROOT, "gpurun_out", "psnr_parity.json")
    calls = -(-steps // 16)
    pkg = importlib.import_module(PKG)
    capi, scene = pkg.capi, importlib.import_module(PKG + ".scene")
    lib, pyngp = capi.load(), pkg.load_pyngp()
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    fx = 0.5 * res / np.tan(0.5 * scene.CAMERA_ANGLE_X)
    fov_val = float(scene.CAMERA_ANGLE_X)          # run.py:101 assigns camera_angle_x (radians) to Testbed.fov; the setter takes it raw

    def render_gt(ogl):
        xf = torch.from_numpy(scene.opengl_to_ngp(ogl)).cuda()
        img = torch.empty((ogl.shape[0], res, res, 4), dtype=torch.uint8, device="cuda")
        capi.check(lib.ngp_render_procedural_views(s, 2024, ogl.shape[0], res, res, fx, fx, C.c_void_p(xf.data_ptr()), C.c_void_p(img.data_ptr())))
        torch.cuda.synchronize()
        return xf.cpu().numpy(), img

    ogl_train = scene.sphere_cameras(views, 2024)
    ogl_test = scene.sphere_cameras(n_test, 777)                      # held-out poses on the same hemisphere
    xf_train, rgba8_train = render_gt(ogl_train)
    xf_test, rgba8_test = render_gt(ogl_test)
    gt8 = rgba8_test.cpu().numpy().astype(np.float32) / 255.0
    gt = linear_to_srgb(srgb_to_linear(gt8[..., :3]) * gt8[..., 3:4])   # composited over black, as both renderers do with bg (0,0,0,1)

    # ---- ours ----
    tb = pyngp.Testbed()
    tb.load_model_config(os.path.join(ROOT, PKG, "configs", "base.json"))
    tb.load_training_data_from_memory(rgba8_train.cpu().numpy(), ogl_train, scene.CAMERA_ANGLE_X, 0.33, [0.5, 0.5, 0.5])
    ms, curve = 0.0, []
    for c in range(calls):
        tb.train(16, 1 << 18)
        b = tb.training_buffer
        ms += b.training_ms + b.training_prep_ms
        curve.append((16 * (c + 1), float(b.loss)))
    tb.fov = fov_val
    tb.background_color = [0, 0, 0, 1]
    ours_psnr, t_render = [], 0.0
    for i in range(n_test):
        tb.set_nerf_camera_matrix(np.ascontiguousarray(ogl_test[i, :3, :4], np.float32))
        t0 = time.perf_counter()
        img = tb.render(res, res, 1, True)
        if i > 0:
            t_render += time.perf_counter() - t0
        ours_psnr.append(psnr(np.clip(img[..., :3], 0, 1), gt[i]))
    ours = {"train_ms_per_step": ms / (16 * calls), "loss": curve[-1][1], "psnr_db": ours_psnr, "psnr_mean_db": float(np.mean(ours_psnr)),
            "render_mrays_per_s": res * res * (n_test - 1) / t_render / 1e6 if n_test > 1 else None,
            "loss_curve": curve[::max(1, len(curve) // 20)]}
    del tb
    torch.cuda.empty_cache()

    # ---- the reference's own CUDA build ----
    ref = {"unavailable": "oracle/_ref/ref_harness not built"}
    if os.path.exists(HARNESS):
        d = tempfile.mkdtemp(prefix="ngp_psnr_")
        try:
            half = torch.empty(rgba8_train.shape, dtype=torch.int16, device="cuda")
            capi.check(lib.ngp_from_rgba32(s, rgba8_train.numel() // 4, C.c_void_p(rgba8_train.data_ptr()), C.c_void_p(half.data_ptr())))
            torch.cuda.synchronize()
            half.cpu().numpy().tofile(os.path.join(d, "images.f16"))
            xf_train.astype(np.float32).tofile(os.path.join(d, "xforms.f32"))
            xf_test.astype(np.float32).tofile(os.path.join(d, "test_xforms.f32"))
            with open(os.path.join(d, "meta.txt"), "w") as f:
                f.write(f"{views} {res} {res} {float(np.float32(fx))!r} {float(np.float32(fx))!r} 0.33\n")
            shutil.copy(os.path.join(ROOT, PKG, "configs", "base.json"), os.path.join(d, "config.json"))
            r = subprocess.run([HARNESS, "trainrender", d, str(calls), str(res), str(res), "1", repr(fov_val)], capture_output=True, text=True, timeout=1500)
            if r.returncode != 0:
                ref = {"unavailable": f"ref_harness failed: {r.stderr[-400:]}"}
            else:
                ref = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
                ref_psnr = []
                for i in range(n_test):
                    img = np.fromfile(os.path.join(d, f"ref_render_{i}.f32"), np.float32).reshape(res, res, 4)
                    ref_psnr.append(psnr(np.clip(img[..., :3], 0, 1), gt[i]))
                ref["psnr_db"], ref["psnr_mean_db"] = ref_psnr, float(np.mean(ref_psnr))
                lc = [l.split() for l in open(os.path.join(d, "ref_loss_curve.txt"))]
                ref["loss_curve"] = [(int(a), float(b)) for a, b, *_ in lc][::max(1, len(lc) // 20)]
        finally:
            shutil.rmtree(d, ignore_errors=True)

    out = {"steps": 16 * calls, "views": views, "res": res, "n_test_poses": n_test, "ours": ours, "reference_cuda": ref}
    os.makedirs(os.path.dirname(out_path), exist_ok=True)
    json.dump(out, open(out_path, "w"), indent=1)
    print(json.dumps({k: (v if k not in ("ours", "reference_cuda") else {kk: vv for kk, vv in v.items() if kk != "loss_curve"}) for k, v in out.items()}))


if __name__ == "__main__":
    main()
