"""Train / evaluate driver with the command line and the evaluation protocol of the reference's
experiment/nerf_synthetic/python/run.py (argument names :14-29, training loop :66-76, evaluation :84-169) and the image metrics of
its common.py (sRGB transfer :22-28, read_image :53-70, SSIM :134-151, mse2psnr :190), on top of pyngp.

  python -m instant-ngp-nerf-minimal_b200.run  is not importable (the package name has dashes); run it as a script:
  python instant-ngp-nerf-minimal_b200/run.py --scene <dir>/transforms_train.json --network <cfg>.json --n_steps 5000 \
         --save_snapshot out/model.msgpack --test_transforms <dir>/transforms_test.json

LPIPS needs a downloaded VGG network and is reported as None (there is no network access here); the xls report becomes a csv."""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def srgb_to_linear(img):
    return np.where(img > 0.04045, np.power((img + 0.055) / 1.055, 2.4), img / 12.92)


def linear_to_srgb(img):
    return np.where(img > 0.0031308, 1.055 * (np.maximum(img, 0.0031308) ** (1.0 / 2.4)) - 0.055, 12.92 * img)


def luminance(a):
    a = np.maximum(0, a) ** 0.4545454545
    return 0.2126 * a[:, :, 0] + 0.7152 * a[:, :, 1] + 0.0722 * a[:, :, 2]


def _convolve1d_reflect(a, k, axis):
    """scipy.ndimage.convolve1d(a, k, axis) with its default mode='reflect' (d c b a | a b c d | d c b a), written out so the
    metric has no dependency beyond numpy"""
    r = len(k) // 2
    a = np.moveaxis(a, axis, 0)
    pad = np.concatenate([a[:r][::-1], a, a[-r:][::-1]], axis=0)
    out = sum(k[len(k) - 1 - j] * pad[j:j + a.shape[0]] for j in range(len(k)))
    return np.moveaxis(out, 0, axis)


def ssim_map(a, b):
    """common.py:134-151: SSIM of the luminance with a 5-tap Gaussian, per pixel"""
    k = np.array([0.120078, 0.233881, 0.292082, 0.233881, 0.120078])

    def blur(x):
        return _convolve1d_reflect(_convolve1d_reflect(x, k, 0), k, 1)
    a, b = luminance(a), luminance(b)
    mA, mB = blur(a), blur(b)
    sA, sB, sAB = blur(a * a) - mA ** 2, blur(b * b) - mB ** 2, blur(a * b) - mA * mB
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    return (2.0 * mA * mB + c1) / (mA * mA + mB * mB + c1) * (2.0 * sAB + c2) / (sA + sB + c2)


def mse2psnr(x):
    return -10.0 * np.log(x) / np.log(10.0)


def compute_error(metric, img, ref):
    """common.py:207-217 for the metrics run.py uses: mean of the per-pixel map (non-finite entries count as 0)"""
    img = np.maximum(np.where(np.isfinite(img), img, 0.0), 0.0)
    if metric == "MSE":
        m = (img - ref) ** 2
    elif metric == "SSIM":
        m = ssim_map(np.clip(img, 0.0, 1.0), np.clip(ref, 0.0, 1.0))
    else:
        raise ValueError(f"Unknown metric: {metric}.")
    m = np.where(np.isfinite(m), m, 0.0)
    if m.ndim == 3:
        m = m.mean(axis=2)
    return float(m.mean())


def read_image(path):
    """common.py:53-70 for PNG / JPEG: float RGBA, colour sRGB -> linear, alpha premultiplied"""
    import PIL.Image
    img = PIL.Image.open(path, "r")
    img = img.convert("RGB" if os.path.splitext(path)[1] == ".jpg" else "RGBA")
    img = np.asarray(img).astype(np.float32) / 255.0
    if img.shape[2] == 4:
        img[..., 0:3] = srgb_to_linear(img[..., 0:3])
        img[..., 0:3] *= img[..., 3:4]
    else:
        img = srgb_to_linear(img)
    return img


def write_image(path, img):
    """common.py:72-90 for PNG: un-premultiply, linear -> sRGB, 8 bit"""
    import PIL.Image
    img = np.array(img, np.float32)
    if img.shape[2] == 4:
        img[..., 0:3] = np.divide(img[..., 0:3], img[..., 3:4], out=np.zeros_like(img[..., 0:3]), where=img[..., 3:4] != 0)
        img[..., 0:3] = linear_to_srgb(img[..., 0:3])
    else:
        img = linear_to_srgb(img)
    PIL.Image.fromarray((np.clip(img, 0.0, 1.0) * 255.0 + 0.5).astype(np.uint8)).save(path)


def evaluate(testbed, ngp, test_transforms_path, output_dir=None, spp=8, max_frames=None, render_depth=True):
    """run.py:84-169: every frame of transforms_test.json is rendered (linear, spp samples, black background, fov on axis 0) and
    compared with its image in clipped sRGB.  Returns {"psnr", "ssim", "lpips": None, "time_s", "frames": [...]}."""
    with open(test_transforms_path) as f:
        test_transforms = json.load(f)
    data_dir = os.path.dirname(test_transforms_path)
    if output_dir:
        os.makedirs(os.path.join(output_dir, "image"), exist_ok=True)
        os.makedirs(os.path.join(output_dir, "depth"), exist_ok=True)
    testbed.background_color = [0.0, 0.0, 0.0, 1.0]
    testbed.rendering_buffer.fov_axis = 0
    testbed.fov = test_transforms["camera_angle_x"]
    rows = []
    for frame in test_transforms["frames"][:max_frames]:
        p = frame["file_path"] + ("" if os.path.splitext(frame["file_path"])[1] else ".png")
        ref_image = read_image(os.path.join(data_dir, p))
        if testbed.color_space == ngp.ColorSpace.SRGB and ref_image.shape[2] == 4:
            # NeRF blends with the background in sRGB space (run.py:108-120)
            ref_image[..., :3] = np.divide(ref_image[..., :3], ref_image[..., 3:4], out=np.zeros_like(ref_image[..., :3]), where=ref_image[..., 3:4] != 0)
            ref_image[..., :3] = linear_to_srgb(ref_image[..., :3])
            ref_image[..., :3] *= ref_image[..., 3:4]
            ref_image += (1.0 - ref_image[..., 3:4]) * np.array(testbed.background_color, np.float32)
            ref_image[..., :3] = srgb_to_linear(ref_image[..., :3])
        testbed.set_nerf_camera_matrix(np.array(frame["transform_matrix"], np.float32)[:3, :])
        t0 = time.perf_counter()
        image = np.array(testbed.render(ref_image.shape[1], ref_image.shape[0], spp, False))
        elapsed = time.perf_counter() - t0
        depth = None
        if render_depth:
            testbed.rendering_buffer.render_mode = ngp.RenderMode.Depth
            depth = np.array(testbed.render(ref_image.shape[1], ref_image.shape[0], spp, False))
            testbed.rendering_buffer.render_mode = ngp.RenderMode.Shade
        A = np.clip(linear_to_srgb(image[..., :3]), 0.0, 1.0)
        R = np.clip(linear_to_srgb(ref_image[..., :3]), 0.0, 1.0)
        rows.append({"file": os.path.basename(p), "psnr": float(mse2psnr(compute_error("MSE", A, R))), "ssim": compute_error("SSIM", A, R), "lpips": None, "time_s": elapsed})
        if output_dir:
            write_image(os.path.join(output_dir, "image", os.path.basename(p)), image[..., :3])
            if depth is not None:
                d8 = (np.clip(np.nan_to_num(depth[..., 0]) / 6.0, 0.0, 1.0) * 255.0 + 0.5).astype(np.uint8)
                import PIL.Image
                PIL.Image.fromarray(d8).save(os.path.join(output_dir, "depth", os.path.splitext(os.path.basename(p))[0] + ".png"))
    out = {"psnr": float(np.mean([r["psnr"] for r in rows])), "ssim": float(np.mean([r["ssim"] for r in rows])), "lpips": None,
           "time_s": float(np.mean([r["time_s"] for r in rows])), "frames": rows}
    if output_dir:
        with open(os.path.join(output_dir, "eval_stat.csv"), "w") as f:
            f.write("file,PSNR,SSIM,LPIPS,Time (s)\n")
            for r in rows:
                f.write(f"{r['file']},{r['psnr']},{r['ssim']},,{r['time_s']}\n")
            f.write(f"mean,{out['psnr']},{out['ssim']},,{out['time_s']}\n")
    return out


def parse_args():
    parser = argparse.ArgumentParser()
    parser.add_argument("--scene", "--training_data", default="", help="The scene to load. A path to the training data.")
    parser.add_argument("--network", default="", help="Path to the network config.")
    parser.add_argument("--load_snapshot", default="", help="Load this snapshot before training. recommended extension: .msgpack")
    parser.add_argument("--save_snapshot", default="", help="Save this snapshot after training. recommended extension: .msgpack")
    parser.add_argument("--nerf_compatibility", action="store_true", help="Matches parameters with original NeRF. Can cause slowness and worse results on some scenes.")
    parser.add_argument("--test_transforms", default="", help="Path to a nerf style transforms json from which we will compute PSNR")
    parser.add_argument("--n_steps", type=int, default=-1, help="Number of steps to train.")
    return parser.parse_args()


def main():
    args = parse_args()
    sys.path.insert(0, ROOT)
    pkg = importlib.import_module("instant-ngp-nerf-minimal_b200")
    ngp = pkg.load_pyngp()
    testbed = ngp.Testbed()
    if args.load_snapshot:
        print("Loading snapshot ", args.load_snapshot)
        testbed.load_snapshot(args.load_snapshot)
    else:
        assert os.path.exists(args.network)
        testbed.load_model_config(args.network)
        testbed.load_training_data(args.scene, 0.33, [0.5, 0.5, 0.5])
    if args.nerf_compatibility:
        print("NeRF compatibility mode enabled")
        testbed.color_space = ngp.ColorSpace.SRGB
    n_steps = args.n_steps
    if n_steps < 0 and not args.load_snapshot:
        n_steps = 100000
    if n_steps > 0:
        assert args.load_snapshot or args.save_snapshot
        output_dir = os.path.dirname(args.load_snapshot if args.load_snapshot else args.save_snapshot)
        start_t = time.perf_counter()
        while testbed.training_buffer.i_step < n_steps:
            testbed.train(16, 2 ** 18)
            if testbed.training_buffer.i_step % 1024 == 0:
                print(f"step {testbed.training_buffer.i_step} loss {testbed.training_buffer.loss:.6f}", flush=True)
        with open(os.path.join(output_dir, "train_timing.txt"), "w") as f:
            f.writelines(f"{time.perf_counter() - start_t} s")
    if args.save_snapshot:
        print("Saving snapshot ", args.save_snapshot)
        testbed.save_snapshot(args.save_snapshot, False)
    if args.test_transforms:
        print("Evaluating test transforms from ", args.test_transforms)
        assert args.load_snapshot or args.save_snapshot
        output_dir = os.path.dirname(args.load_snapshot if args.load_snapshot else args.save_snapshot)
        r = evaluate(testbed, ngp, args.test_transforms, output_dir)
        print(f"PSNR={r['psnr']:.3f} SSIM={r['ssim']:.3f} LPIPS(VGG)=n/a FPS={1.0 / r['time_s']:.2f}")


if __name__ == "__main__":
    main()
