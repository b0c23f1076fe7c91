"""TEST INFRASTRUCTURE: runs the reference's own CUDA implementation (oracle/_ref/ref_harness, the unmodified reference
compiled for sm_100 by oracle/build_ref.sh) on a small synthetic dataset and loads what it dumped.  GPU box only; nothing
here reads /root/reference at run time."""
import json
import os
import re
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


def write_config(d, **edits):
    """base.json (the reference's own default model) with some values replaced, e.g. **{"samlper.aabb_scale": 4}"""
    text = re.sub(r"^\s*//.*$", "", open(os.path.join(ROOT, PKG, "configs", "base.json")).read(), flags=re.M)
    cfg = json.loads(text)
    for path, value in edits.items():
        node = cfg
        keys = path.split(".")
        for k in keys[:-1]:
            node = node[k]
        node[keys[-1]] = value
    out = os.path.join(d, "config.json")
    json.dump(cfg, open(out, "w"))
    return out


def write_test_poses(d, scene, n=3, seed=77):
    """held-out cameras on the same rig: OpenGL matrices for pyngp.set_nerf_camera_matrix, NGP-frame ones for the harness"""
    ogl = scene.sphere_cameras(n, seed)
    np.ascontiguousarray(scene.opengl_to_ngp(ogl), np.float32).tofile(os.path.join(d, "test_xforms.f32"))
    return ogl


def run(*args, timeout=900):
    """ref_harness <args...>; returns the last JSON line it printed (or {} when it printed none)"""
    r = subprocess.run([HARNESS] + [str(a) for a in args], capture_output=True, text=True, timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError(f"ref_harness {args[0]} failed ({r.returncode}):\n{r.stdout[-2000:]}\n{r.stderr[-4000:]}")
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    return json.loads(lines[-1]) if lines else {}


def render_views(capi, scene, ogl, res, seed=SEED):
    """the procedural scene seen from the given OpenGL cameras: RGBA8 (sRGB, straight alpha) [n][res][res][4], as a PNG would hold it"""
    import ctypes as C
    import torch
    lib = capi.load()
    xf = torch.from_numpy(np.ascontiguousarray(scene.opengl_to_ngp(ogl), np.float32)).cuda()
    fx = float(np.float32(0.5 * res / np.tan(0.5 * scene.CAMERA_ANGLE_X)))
    rgba8 = torch.empty((len(ogl), res, res, 4), dtype=torch.uint8, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    capi.check(lib.ngp_render_procedural_views(s, seed, len(ogl), res, res, fx, fx, C.c_void_p(xf.data_ptr()), C.c_void_p(rgba8.data_ptr())))
    torch.cuda.synchronize()
    return rgba8.cpu().numpy()


def write_png_scene(d, split, ogl, rgba8, scene):
    """a nerf_synthetic-style directory: <d>/transforms_<split>.json + <d>/<split>/r_<i>.png (extension omitted in file_path, as Blender writes it)"""
    import PIL.Image
    os.makedirs(os.path.join(d, split), exist_ok=True)
    frames = []
    for i, (m, img) in enumerate(zip(ogl, rgba8)):
        PIL.Image.fromarray(img, "RGBA").save(os.path.join(d, split, f"r_{i}.png"))
        frames.append({"file_path": f"./{split}/r_{i}", "rotation": 0.0125, "transform_matrix": [[float(v) for v in row] for row in m] + [[0.0, 0.0, 0.0, 1.0]]})
    path = os.path.join(d, f"transforms_{split}.json")
    json.dump({"camera_angle_x": scene.CAMERA_ANGLE_X, "frames": frames}, open(path, "w"), indent=1)
    return path


def write_dataset(d, capi, scene, n_views=N_VIEWS, res=RES, seed=SEED, config=None):
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
    if config:
        write_config(d, **config)
    else:
        shutil.copy(os.path.join(ROOT, PKG, "configs", "base.json"), os.path.join(d, "config.json"))
    return dict(xforms=xf, images=images, fx=float(fx), W=res, H=res, n_views=n_views, ogl=ogl, rgba8=rgba8.cpu().numpy())


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


def make(capi, scene, params=None, batch=BATCH, n_rays=0, n_views=N_VIEWS, res=RES, config=None, envmap=None):
    """params (optional): fp16 parameter arena the reference starts from instead of its seed-43 initialisation;
    config: base.json edits; envmap (optional): fp16 texels the trainable environment map starts from"""
    d = tempfile.mkdtemp(prefix="ngp_refcuda_")
    ds = write_dataset(d, capi, scene, n_views, res, config=config)
    if params is not None:
        np.ascontiguousarray(params, np.float16).tofile(os.path.join(d, "params_init.f16"))
    if envmap is not None:
        np.ascontiguousarray(envmap, np.float16).tofile(os.path.join(d, "envmap_init.f16"))
    run_dump(d, batch, n_rays)
    return Dump(d, ds)
