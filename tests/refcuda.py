"""TEST INFRASTRUCTURE: runs the reference's own CUDA implementation (oracle/_ref/ref_harness, the unmodified reference
compiled for sm_100 by oracle/build_ref.sh) on a small synthetic dataset and loads what it dumped.  GPU box only; nothing
here reads /root/reference at run time."""
import os
import shutil
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
PKG = "instant-ngp-nerf-minimal_b200"

N_VIEWS, RES, SEED, BATCH = 12, 200, 2024, 1 << 18


def available():
    return os.path.exists(HARNESS) and os.access(HARNESS, os.X_OK)


def write_dataset(d, capi, scene, n_views=N_VIEWS, res=RES, seed=SEED):
    """procedural views rendered by the device fixture renderer, converted with ngp_from_rgba32 (both tested against the CPU
    oracle elsewhere) -> the raw files ref_harness reads"""
    import ctypes as C
    import torch
    lib = capi.load()
    ogl = scene.sphere_cameras(n_views, seed)
    xf = np.ascontiguousarray(scene.opengl_to_ngp(ogl), np.float32)
    fx = np.float32(0.5 * res / np.tan(0.5 * scene.CAMERA_ANGLE_X))
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    d_xf = torch.from_numpy(xf).cuda()
    rgba8 = torch.empty((n_views, res, res, 4), dtype=torch.uint8, device="cuda")
    capi.check(lib.ngp_render_procedural_views(s, seed, n_views, res, res, float(fx), float(fx), C.c_void_p(d_xf.data_ptr()), C.c_void_p(rgba8.data_ptr())))
    half = torch.empty((n_views, res, res, 4), dtype=torch.int16, device="cuda")
    capi.check(lib.ngp_from_rgba32(s, n_views * res * res, C.c_void_p(rgba8.data_ptr()), C.c_void_p(half.data_ptr())))
    torch.cuda.synchronize()
    images = half.cpu().numpy().view(np.float16)
    xf.tofile(os.path.join(d, "xforms.f32"))
    images.tofile(os.path.join(d, "images.f16"))
    with open(os.path.join(d, "meta.txt"), "w") as f:
        f.write(f"{n_views} {res} {res} {float(fx)!r} {float(fx)!r} 0.33\n")
    shutil.copy(os.path.join(ROOT, PKG, "configs", "base.json"), os.path.join(d, "config.json"))
    return dict(xforms=xf, images=images, fx=float(fx), W=res, H=res, n_views=n_views)


def start_params(n_params):
    """O(1) parameters the reference is started from: xavier-uniform MLP matrices (density head x24, so log-densities reach
    +-10 and rays terminate early), hash table U(-1, 1).  Deterministic (numpy PCG64), so tests can regenerate them."""
    g = np.random.Generator(np.random.PCG64(20240))
    parts = []
    for i, (o, k) in enumerate([(64, 32), (16, 64), (64, 32), (64, 64), (16, 64)]):
        s = np.sqrt(6.0 / (o + k)) * (24.0 if i == 1 else 1.0)
        parts.append((g.random(o * k, dtype=np.float32) * 2 - 1) * s)
    parts.append(g.random(n_params - 10240, dtype=np.float32) * 2 - 1)
    return np.concatenate(parts).astype(np.float16)


def run_dump(d, batch=BATCH, n_rays=0, timeout=900):
    r = subprocess.run([HARNESS, "dump", d, str(batch), str(n_rays)], capture_output=True, text=True, timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError(f"ref_harness dump failed ({r.returncode}):\n{r.stdout[-2000:]}\n{r.stderr[-4000:]}")
    return r.stdout


class Dump:
    """lazy loader of <dir>/ref_*.bin"""

    def __init__(self, d, dataset):
        self.dir, self.dataset = d, dataset
        self.meta = {}
        for line in open(os.path.join(d, "ref_step.txt")):
            k, v = line.split()
            self.meta[k] = int(v)

    def arr(self, name, dtype, shape=None):
        a = np.fromfile(os.path.join(self.dir, f"ref_{name}.bin"), dtype=dtype)
        return a if shape is None else a.reshape(shape)


def make(capi, scene, params=None, batch=BATCH, n_rays=0, n_views=N_VIEWS, res=RES):
    """params (optional): fp16 parameter arena the reference starts from instead of its seed-43 initialisation"""
    d = tempfile.mkdtemp(prefix="ngp_refcuda_")
    ds = write_dataset(d, capi, scene, n_views, res)
    if params is not None:
        np.ascontiguousarray(params, np.float16).tofile(os.path.join(d, "params_init.f16"))
    run_dump(d, batch, n_rays)
    return Dump(d, ds)
