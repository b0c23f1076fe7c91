#!/usr/bin/env python
"""Is the psnr_db gap between the two bench arms at step 352 a TRAINING difference or a RENDERING difference?
The reference build trains the bench dataset for 352 steps, saves a snapshot and renders the bench's held-out poses (seed 99);
our renderer then renders the same poses FROM THE REFERENCE'S SNAPSHOT.  Prints per-view PSNR against the ground truth for both
renders and the PSNR between the two renders.  (tests/ infrastructure is used: GPU box only.)"""
import importlib, json, os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import refcuda
import bench
PKG = "instant-ngp-nerf-minimal_b200"
pkg = importlib.import_module(PKG)
capi = pkg.capi; lib = capi.load(); pyngp = pkg.load_pyngp(); scene = importlib.import_module(PKG + ".scene"); run = importlib.import_module(PKG + ".run")


class A: pass


def main():
    import ctypes as C, torch
    args = A(); args.views, args.res = 100, 800
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 352
    ogl = scene.sphere_cameras(args.views, 2024)
    rgba8 = torch.from_numpy(refcuda.render_views(capi, scene, ogl, args.res)).cuda()
    fx = 0.5 * args.res / np.tan(0.5 * scene.CAMERA_ANGLE_X)
    d = bench.stage_reference_dataset(args, capi, lib, scene.opengl_to_ngp(ogl), fx, rgba8, scene)
    test = bench.held_out_poses(scene)
    info = refcuda.run("snapshot", d, steps // 16, args.res, args.res, scene.CAMERA_ANGLE_X, "1:0:1:0")
    gt = bench.ground_truth_linear(args, capi, scene, lib, test)
    ref_imgs = [np.fromfile(os.path.join(d, f"ref_render_0_{p}.f32"), np.float32).reshape(args.res, args.res, 4) for p in range(len(test))]
    tb = pyngp.Testbed()
    tb.load_snapshot(os.path.join(d, "ref_snapshot.msgpack"))
    tb.background_color = [0.0, 0.0, 0.0, 1.0]
    tb.fov = scene.CAMERA_ANGLE_X
    ours_imgs = []
    for m in test:
        tb.set_nerf_camera_matrix(m)
        ours_imgs.append(np.array(tb.render(args.res, args.res, 1, False)))
    print("reference render of its own model :", [round(v, 3) for v in bench.psnr_db(ref_imgs, gt)[1]])
    print("our render of the reference model :", [round(v, 3) for v in bench.psnr_db(ours_imgs, gt)[1]])
    print("between the two renders (dB)       :", [round(v, 2) for v in bench.psnr_db(ours_imgs, ref_imgs)[1]])
    # and our own training, same steps, same poses
    tb2 = pyngp.Testbed()
    tb2.load_model_config(os.path.join(ROOT, PKG, "configs", "base.json"))
    tb2.load_training_data_from_memory(rgba8.cpu().numpy(), ogl, scene.CAMERA_ANGLE_X, 0.33, [0.5, 0.5, 0.5])
    while tb2.training_buffer.i_step < steps:
        tb2.train(16, 1 << 18)
    tb2.background_color = [0.0, 0.0, 0.0, 1.0]
    tb2.fov = scene.CAMERA_ANGLE_X
    mine = []
    for m in test:
        tb2.set_nerf_camera_matrix(m)
        mine.append(np.array(tb2.render(args.res, args.res, 1, False)))
    print("our render of OUR model           :", [round(v, 3) for v in bench.psnr_db(mine, gt)[1]], "loss", tb2.training_buffer.loss, "ref loss", info.get("loss"))


if __name__ == "__main__":
    main()
