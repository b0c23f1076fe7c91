"""GPU parity of the real-dataset path against the reference's own loader.

The reference's Testbed::load_training_data (internal/testbed.cu:95-125) -> NeRFSynthetic::load_from_json
(internal/dataset/nerf_synthetic.cu:31-134: transforms json, stb_image PNG decode, spec_opengl_to_opencv, fov -> focal length,
from_rgba32 on the GPU) runs inside oracle/_ref/ref_harness on a directory the test writes; pyngp.load_training_data reads the
same directory.  Camera matrices, focal length and every fp16 pixel must be BIT-EXACT.  Then the evaluation protocol of
experiment/nerf_synthetic/python/run.py:84-169 (package run.py) on held-out views, and PSNR parity with the reference's build
after the same number of training steps (+-0.3 dB)."""
import importlib
import json
import os
import tempfile

import numpy as np
import pytest

import pngutil
import refcuda

pytestmark = pytest.mark.gpu
PKG = "instant-ngp-nerf-minimal_b200"


@pytest.fixture(scope="module")
def pyngp():
    pkg = importlib.import_module(PKG)
    pkg.capi.load()
    return pkg.load_pyngp()


def test_loader_matches_the_references_loader_bit_for_bit(capi, pyngp):
    """one image per PNG flavour the decoder distinguishes (all five filter types, RGB / gray / gray+alpha / palette / 16 bit /
    interlaced), poses from the camera rig; everything the loader produces is compared with the reference's dump"""
    if not refcuda.available():
        pytest.skip("oracle/_ref/ref_harness is not built")
    scene = importlib.import_module(PKG + ".scene")
    d = tempfile.mkdtemp(prefix="ngp_loaddata_")
    w, h = 40, 28                            # deliberately not square: resolution.x is the width (nerf_synthetic.cu:108)
    flavours = [c for c in pngutil.cases(seed=9, w=w, h=h)]
    ogl = scene.sphere_cameras(len(flavours), 31)
    os.makedirs(os.path.join(d, "train"))
    frames = []
    for i, (name, kw) in enumerate(flavours):
        pngutil.write_png(os.path.join(d, "train", f"{name}.png"), **kw)
        # file_path without extension (Blender's convention, nerf_synthetic.cu:66-76) for every other frame, with it for the rest
        fp = f"./train/{name}" + (".png" if i % 2 else "")
        frames.append({"file_path": fp, "transform_matrix": [[float(v) for v in row] for row in ogl[i]] + [[0.0, 0.0, 0.0, 1.0]]})
    jpath = os.path.join(d, "transforms_train.json")
    json.dump({"camera_angle_x": scene.CAMERA_ANGLE_X, "frames": frames}, open(jpath, "w"))
    refcuda.write_config(d)
    refcuda.run("loaddata", d, jpath, 0.33)
    meta = open(os.path.join(d, "ref_ld_meta.txt")).read().split("\n")
    n, rw, rh, fx, fy, ppx, ppy = [float(x) for x in meta[0].split()]
    assert (int(n), int(rw), int(rh)) == (len(flavours), w, h)

    tb = pyngp.Testbed()
    tb.load_model_config(os.path.join(d, "config.json"))
    tb.load_training_data(jpath, 0.33, [0.5, 0.5, 0.5])
    assert tb.n_training_images == len(flavours)
    assert np.array_equal(np.array(tb.training_xforms, np.float32).view(np.uint32), np.fromfile(os.path.join(d, "ref_ld_xforms.f32"), np.uint32).reshape(-1, 12)), "camera matrices"
    assert abs(tb.training_focal_length[0] - fx) <= 1e-6 * fx and abs(tb.training_focal_length[1] - fy) <= 1e-6 * fy     # the dump prints 9 digits
    assert list(tb.training_filenames) == [m for m in meta[1:] if m]
    ref_img = np.fromfile(os.path.join(d, "ref_ld_images.f16"), np.float16).reshape(len(flavours), h, w, 4)
    for i, (name, kw) in enumerate(flavours):
        ours = np.array(tb.training_image(i))
        assert ours.shape == (h, w, 4)
        assert np.array_equal(ours.astype(np.float16).view(np.uint16), ref_img[i].view(np.uint16)), f"pixels of {name}"
        # and the decode itself, through the expectation the CPU tests use
        rgba8 = pngutil.expected_rgba8(kw["samples"], kw["ctype"], kw.get("depth", 8), kw.get("palette"), kw.get("trns"))
        assert np.array_equal(ours[..., 3].astype(np.float32), (rgba8[..., 3].astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float16).astype(np.float32))
    # sampler initialisation from the loaded cameras (OccupancySampler::init -> mark_untrained_density_grid): the snapshot carries the grid
    import msgpack
    snap = os.path.join(d, "after_load.msgpack")
    tb.save_snapshot(snap, False)
    grid = np.frombuffer(msgpack.unpackb(open(snap, "rb").read(), raw=False, strict_map_key=False)["snapshot"]["density_grid"], np.float32)
    assert np.array_equal(grid, np.fromfile(os.path.join(d, "ref_ld_grid0_f32.bin"), np.float32))
    with pytest.raises(RuntimeError):
        tb.load_training_data(os.path.join(d, "missing.json"), 0.33, [0.5, 0.5, 0.5])
    with pytest.raises(RuntimeError):
        tb.load_training_data(os.path.join(d, "train", flavours[0][0] + ".png"), 0.33, [0.5, 0.5, 0.5])      # "The data path must be a .json file."


@pytest.fixture(scope="module")
def png_scene(capi):
    scene = importlib.import_module(PKG + ".scene")
    d = tempfile.mkdtemp(prefix="ngp_pngscene_")
    res, n_train, n_test = 200, 16, 4
    ogl_train, ogl_test = scene.sphere_cameras(n_train, 2024), scene.sphere_cameras(n_test, 99)
    train8, test8 = refcuda.render_views(capi, scene, ogl_train, res), refcuda.render_views(capi, scene, ogl_test, res)
    jtrain = refcuda.write_png_scene(d, "train", ogl_train, train8, scene)
    jtest = refcuda.write_png_scene(d, "test", ogl_test, test8, scene)
    refcuda.write_config(d)
    return dict(dir=d, train=jtrain, test=jtest, res=res, ogl_train=ogl_train, ogl_test=ogl_test, train8=train8, test8=test8, scene=scene)


STEPS = 512


def test_eval_loop_and_psnr_parity_with_the_reference_build(capi, pyngp, png_scene):
    """run.py's protocol end to end: load_training_data(transforms_train.json) -> train(16, 2^18) x 32 -> evaluate(transforms_test.json):
    PSNR / SSIM on held-out views.  The reference's CUDA build trains the same images for the same 512 steps and renders the same
    poses (ref_harness snapshot); both sets of pictures go through the same metric code: PSNR within +-0.3 dB."""
    run = importlib.import_module(PKG + ".run")
    s = png_scene
    tb = pyngp.Testbed()
    tb.load_model_config(os.path.join(s["dir"], "config.json"))
    tb.load_training_data(s["train"], 0.33, [0.5, 0.5, 0.5])
    while tb.training_buffer.i_step < STEPS:
        tb.train(16, 1 << 18)
    out_dir = os.path.join(s["dir"], "eval")
    r = run.evaluate(tb, pyngp, s["test"], out_dir, spp=1)
    assert len(r["frames"]) == 4 and all(np.isfinite(f["psnr"]) and 0 < f["ssim"] <= 1 for f in r["frames"])
    assert os.path.exists(os.path.join(out_dir, "image", "r_0.png")) and os.path.exists(os.path.join(out_dir, "depth", "r_0.png")) and os.path.exists(os.path.join(out_dir, "eval_stat.csv"))
    assert r["psnr"] > 20.0 and r["ssim"] > 0.7, r
    if not refcuda.available():
        pytest.skip("oracle/_ref/ref_harness is not built: no reference side for the parity half")
    # the reference on the same images (handed over decoded: its PNG path is pinned by the loader test above)
    d = tempfile.mkdtemp(prefix="ngp_psnr_")
    import ctypes as C
    import torch
    lib = capi.load()
    half = torch.empty(s["train8"].shape, dtype=torch.int16, device="cuda")
    d8 = torch.from_numpy(s["train8"]).cuda()
    capi.check(lib.ngp_from_rgba32(C.c_void_p(torch.cuda.current_stream().cuda_stream), d8.numel() // 4, C.c_void_p(d8.data_ptr()), C.c_void_p(half.data_ptr())))
    torch.cuda.synchronize()
    half.cpu().numpy().tofile(os.path.join(d, "images.f16"))
    np.ascontiguousarray(s["scene"].opengl_to_ngp(s["ogl_train"]), np.float32).tofile(os.path.join(d, "xforms.f32"))
    res = s["res"]
    fx = float(np.float32(0.5 * res / np.tan(0.5 * s["scene"].CAMERA_ANGLE_X)))
    open(os.path.join(d, "meta.txt"), "w").write(f"{len(s['ogl_train'])} {res} {res} {fx!r} {fx!r} 0.33\n")
    refcuda.write_config(d)
    np.ascontiguousarray(s["scene"].opengl_to_ngp(s["ogl_test"]), np.float32).tofile(os.path.join(d, "test_xforms.f32"))
    info = refcuda.run("snapshot", d, STEPS // 16, res, res, s["scene"].CAMERA_ANGLE_X, "1:0:1:0")
    psnr_ref, ssim_ref = [], []
    for p in range(4):
        img = np.fromfile(os.path.join(d, f"ref_render_0_{p}.f32"), np.float32).reshape(res, res, 4)
        gt = run.read_image(os.path.join(s["dir"], "test", f"r_{p}.png"))
        A, R = np.clip(run.linear_to_srgb(img[..., :3]), 0, 1), np.clip(run.linear_to_srgb(gt[..., :3]), 0, 1)
        psnr_ref.append(run.mse2psnr(run.compute_error("MSE", A, R)))
        ssim_ref.append(run.compute_error("SSIM", A, R))
    print(f"PSNR after {STEPS} steps: ours {r['psnr']:.3f} dB / SSIM {r['ssim']:.4f}; reference build {np.mean(psnr_ref):.3f} dB / SSIM {np.mean(ssim_ref):.4f}; "
          f"loss ours {tb.training_buffer.loss:.5f} reference {info['loss']:.5f}")
    assert abs(r["psnr"] - float(np.mean(psnr_ref))) <= 0.3, (r["psnr"], psnr_ref)
    # SSIM after 512 steps moves by +-0.05 from run to run on EITHER side (atomics-ordered training, a mostly empty background where
    # the 5x5 windows see noise only): measured 0.758 / 0.818 / 0.807 (ours) and 0.812 / 0.772 / 0.722 (reference) in three
    # back-to-back runs.  The bar SURVEY.md 8c sets is the PSNR one above; SSIM only has to be no worse than the reference's spread.
    assert r["ssim"] >= float(np.mean(ssim_ref)) - 0.1 and abs(r["ssim"] - float(np.mean(ssim_ref))) <= 0.15, (r["ssim"], ssim_ref)
