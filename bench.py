#!/usr/bin/env python
"""bench.py -- training samples/s of the instant-NGP NeRF step on a nerf_synthetic-shaped procedural scene
(BASELINE.json configs[1]: 100 views 800x800, 2^18-sample batches, base.json model), 1..8 B200.

  python bench.py --gpus N --steps K --warmup W            our sm_100a path through the pyngp Testbed (one rank per GPU)
  python bench.py --impl reference --steps K --warmup W    the reference's algorithm on the host cores (CPU oracle, rank 0)

A "step" is one training step (march -> network -> composite -> forward/backward -> optimizer); occupancy-grid refreshes
happen every 16 steps as in the reference's run.py loop and are inside the timed region.  One JSON line on stdout.

BASELINE.json's metric is "train samples/s & render Mrays/s ... PSNR parity" over five configs; the contract keys carry configs[1]
(training), and the same line carries the rest as objects, measured in the same run by both arms:
  render   configs[3]: 1920x1080 novel views, spp 1, Mrays/s (N > 1: row bands, one per rank, no collective)
  psnr_db  4 held-out 800x800 poses after exactly warmup + steps training steps (run.py's protocol: linear render, clipped sRGB)
  c1       configs[0]: hash-grid encode fwd / bwd on 2^18 random points, microseconds
  c5       configs[4]: T = 2^22, 2^20-sample batches, aabb_scale 16 (5 cascades): ms/step and samples/s on a bounded number of calls"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "instant-ngp-nerf-minimal_b200"
METRIC = "nerf_train_samples_per_s"
UNIT = "samples/s"
ENQUEUE_MS = 0.0                     # host time spent issuing work, summed over train() calls (diagnostic)
CALL_STEPS = 16                      # run.py:71 trains in calls of 16 steps; the occupancy grid is refreshed once per call


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=256)
    p.add_argument("--warmup", type=int, default=96)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--batch", type=int, default=1 << 18)
    p.add_argument("--views", type=int, default=100)
    p.add_argument("--res", type=int, default=800)
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-extras", action="store_true", help="skip the render / psnr_db / c1 / c5 objects")
    p.add_argument("--c5-steps", type=int, default=32, help="timed steps of the configs[4] leg (after 32 warm-up steps)")
    p.add_argument("--timeline", default=None, help="write a per-launch timeline (json per rank) of two further train() calls with the march pipelined")
    p.add_argument("--dp-mode", default="peer", choices=["peer", "nccl"],
                   help="N > 1: fused reduce-scatter + Adam + all-gather over NVLink peer memory (default), or NCCL all-reduce + replicated optimizer")
    p.add_argument("--no-ref-cuda", action="store_true", help="(kept so older command lines still parse: the reference's CUDA build is timed by --impl reference only)")
    p.add_argument("--reference-kind", default="auto", choices=["auto", "cuda", "cpu"],
                   help="--impl reference: the reference's CUDA build when oracle/_ref/ref_harness exists (auto/cuda), or the CPU oracle port")
    return p.parse_args()


def config_dict(args, n_gpus):
    return {"workload": f"configs[1]: nerf_synthetic-shaped procedural scene, {args.views} views {args.res}x{args.res}, base.json model (L=16 F=2 T=2^19, 64-wide MLPs), "
                        f"{args.batch}-sample batches, occupancy refresh every {CALL_STEPS} steps",
            "global_batch": args.batch * n_gpus, "per_gpu_batch": args.batch, "parallelism": (f"dp{n_gpus} ({'peer-memory sharded optimizer' if args.dp_mode == 'peer' else 'nccl all-reduce'})") if n_gpus > 1 else "single",
            "l2_policy": "inputs larger than L2 (512 MB fp16 image set sampled at random pixels; 37-117 MB of samples streamed per step); "
                         "the 23 MB hash table is L2-resident by design, as in the real workload"}


class ClockSampler:
    """SM clock + throttle reasons sampled WHILE the timed region runs.  NVML in a thread (pyngp releases the GIL inside train());
    spawning nvidia-smi every 100 ms measurably disturbs a step that lasts 0.7 ms, so it is only the fallback."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index
        self.sm, self.mx, self.reasons, self.stop_flag, self.thread, self.nvml = [], [], set(), False, None, None
        self.max_query_ms = 0.0
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml, self.handle = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self.mx.append(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))     # static
            self._query()                          # first use of each query initialises driver state: keep that out of the timed region
            self.sm, self.reasons, self.max_query_ms = [], set(), 0.0
        except Exception:
            self.nvml = None

    def _query(self):
        n = self.nvml
        t0 = time.perf_counter()
        self.sm.append(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
        r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
        self.max_query_ms = max(self.max_query_ms, (time.perf_counter() - t0) * 1e3)
        for bit, name in ((n.nvmlClocksThrottleReasonHwSlowdown, "hw_slowdown"), (n.nvmlClocksThrottleReasonHwThermalSlowdown, "hw_thermal_slowdown"),
                          (n.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_thermal_slowdown"), (n.nvmlClocksThrottleReasonSwPowerCap, "sw_power_cap")):
            if r & bit:
                self.reasons.add(name)

    def sample_now(self):
        """one sample taken by the caller's thread: immediately before and immediately after the timed region, so a region shorter than
        the polling period still carries its own clocks"""
        if self.nvml is not None and not os.environ.get("NGP_NO_CLOCKS"):
            try:
                self._query()
            except Exception:
                pass

    def _summary(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": float(max(self.mx)) if self.mx else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml", "max_query_ms": round(self.max_query_ms, 3)}

    def _poll(self):
        # first sample 30 ms into the region, then every 200 ms (the recipe's nvidia-smi -lms 200): every query takes a driver lock
        # that can hold up the launches of all processes on the box for milliseconds, and a 0.6 ms step notices
        period = float(os.environ.get("NGP_CLOCK_PERIOD_S", "0.2"))
        t_next = time.perf_counter() + 0.03
        while not self.stop_flag:
            if time.perf_counter() >= t_next:
                try:
                    self._query()
                except Exception:
                    pass
                t_next += period
            time.sleep(0.005)

    def start(self):
        if os.environ.get("NGP_NO_CLOCKS") or (int(os.environ.get("RANK", "0")) != 0 and not os.environ.get("NGP_CLOCKS_ALL_RANKS")):
            # rank 0 samples its GPU; a query takes a driver lock that holds up launches of every process on the box, and in a
            # lock-stepped data-parallel run each hold-up is paid by all ranks (measured at 2 GPUs: 0.805 -> 0.737 ms/step)           #             self.nvml, self.thread, self.proc = None, None, None
            return
        if self.nvml is not None:
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: [self.rows.append(l) for l in self.proc.stdout], daemon=True).start()
        except OSError:
            self.proc = None

    def stop(self):
        if self.thread is not None:
            self.stop_flag = True
            self.thread.join(timeout=1.0)
            return self._summary()
        if self.nvml is not None and self.sm:
            return self._summary()
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        time.sleep(0.05)
        sm, mx, reasons = [], [], set()
        for l in self.rows:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 6 or not f[0].isdigit():
                continue
            sm.append(int(f[0])); mx.append(int(f[1]))
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(max(mx)) if mx else None, "reasons": sorted(reasons), "samples": len(sm),
                "source": "nvidia-smi"}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1400.0, "fallback"


# ALGORITHMIC bytes (or flops) per STEP of each hot kernel, from the units it processes (per-unit figures
# from SURVEY.md 8d): m = samples marched per step, b = samples in the compacted batch, r = rays, p = params,
# g = occupancy-refresh positions per step (one refresh of 2^22 / 2^21 positions every 16 steps)
def kernel_work(name, m, b, r, p, g=0.0, world=1, peer=False):
    if name == "adam_ema_kernel" and world > 1 and peer:
        # the fused exchange kernel: this rank steps p / world parameters (36 B of local fp32 state + 6 B of EMA each), reads their
        # gradients out of `world` arenas and stores the new fp16 weights into `world` arenas over NVLink (2 B each way per arena)
        return ("hbm", (p / world) * (42 + 4 * (world - 1)))
    table = {
        "composite_count_kernel": ("hbm", m * (8 + 4) + r * 40),
        "composite_grad_kernel": ("hbm", b * (8 + 28 + 8 + 28 + 4) + r * 40),
        "march_count_kernel": ("hbm", r * (64 + 256)),
        "march_emit_kernel": ("hbm", m * 28 + r * (64 + 256)),
        # the table (24.4 MB at T = 2^19) is L2-resident by design: 512 of the 588 B per point are L2 gathers / reductions
        "hash_encode_fwd_kernel": ("l2", (m + g) * 588),                  # the marched samples + the refresh positions; the batch reuses pass 1
        "hash_encode_bwd_kernel": ("l2", b * 588 + 2 * (p - 10240)),
        "adam_ema_kernel": ("hbm", p * 42),
        "mlp_kernel": ("tensor", b * 54400),                              # training launch: forward + dgrad + wgrad on the compacted batch
        "mlp_fwd_kernel": ("tensor", m * 18816 + g * 6144),               # inference launch on the marched samples + density-only refresh
    }
    return table.get(name)


REF_HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")


def make_scene_on_device(args, capi, scene, lib):
    """procedural views rendered once by the fixture renderer (never timed): (ogl matrices, ngp-frame matrices, device rgba8)"""
    import ctypes as C
    import torch
    ogl = scene.sphere_cameras(args.views, 2024)
    xf_np = scene.opengl_to_ngp(ogl)
    xf = torch.from_numpy(xf_np).cuda()
    fx = 0.5 * args.res / np.tan(0.5 * scene.CAMERA_ANGLE_X)
    rgba8 = torch.empty((args.views, args.res, args.res, 4), dtype=torch.uint8, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    capi.check(lib.ngp_render_procedural_views(s, 2024, args.views, args.res, args.res, fx, fx, C.c_void_p(xf.data_ptr()), C.c_void_p(rgba8.data_ptr())))
    torch.cuda.synchronize()
    return ogl, xf_np, fx, rgba8


HELD_OUT, RENDER_W, RENDER_H = 4, 1920, 1080
C5_EDITS = {"samlper.aabb_scale": 16, "network.encoding.log2_hashmap_size": 22}
C5_BATCH, C5_WARMUP = 1 << 20, 32
C1_WORKLOAD = "configs[0]: hash-grid encode fwd / bwd (incl. the gradient-table clear), 2^18 uniformly random points, L=16 F=2 T=2^19, table L2-resident, median of 100"
C5_WORKLOAD = "configs[4]: T=2^22 per level (150.6 MB fp16 table, spills L2), 2^20-sample batches, aabb_scale 16 (5 cascades), same 100 views"


def held_out_poses(scene):
    """4 cameras of the same rig that are not training views (another jitter seed)"""
    return scene.sphere_cameras(HELD_OUT, 99)


def write_config(path, **edits):
    """base.json with some values replaced (the file carries // comments, which the reference's json reader accepts)"""
    import re
    text = re.sub(r"^\s*//.*$", "", open(os.path.join(ROOT, PKG, "configs", "base.json")).read(), flags=re.M)
    cfg = json.loads(text)
    for key, value in edits.items():
        node = cfg
        keys = key.split(".")
        for k in keys[:-1]:
            node = node[k]
        node[keys[-1]] = value
    json.dump(cfg, open(path, "w"))
    return path


def ground_truth_linear(args, capi, scene, lib, ogl):
    """the held-out views as run.py's read_image would return their PNGs: float RGBA, sRGB -> linear, alpha premultiplied"""
    import ctypes as C
    import torch
    run = importlib.import_module(PKG + ".run")
    xf = torch.from_numpy(np.ascontiguousarray(scene.opengl_to_ngp(ogl), np.float32)).cuda()
    fx = 0.5 * args.res / np.tan(0.5 * scene.CAMERA_ANGLE_X)
    rgba8 = torch.empty((len(ogl), args.res, args.res, 4), dtype=torch.uint8, device="cuda")
    capi.check(lib.ngp_render_procedural_views(C.c_void_p(torch.cuda.current_stream().cuda_stream), 2024, len(ogl), args.res, args.res, fx, fx,
                                               C.c_void_p(xf.data_ptr()), C.c_void_p(rgba8.data_ptr())))
    torch.cuda.synchronize()
    img = rgba8.cpu().numpy().astype(np.float32) / 255.0
    img[..., :3] = run.srgb_to_linear(img[..., :3]) * img[..., 3:4]
    return img


def psnr_db(images, gt):
    """run.py:139-143: mean over the views of PSNR between clipped-sRGB render and clipped-sRGB ground truth (black background)"""
    run = importlib.import_module(PKG + ".run")
    vals = []
    for a, r in zip(images, gt):
        A, R = np.clip(run.linear_to_srgb(a[..., :3]), 0.0, 1.0), np.clip(run.linear_to_srgb(r[..., :3]), 0.0, 1.0)
        vals.append(float(run.mse2psnr(run.compute_error("MSE", A, R))))
    return float(np.mean(vals)), vals


def stage_reference_dataset(args, capi, lib, xf_np, fx, rgba8, scene):
    """the raw files ref_harness reads (same images, same cameras, base.json) + the held-out poses"""
    import ctypes as C
    import shutil
    import tempfile
    import torch
    d = tempfile.mkdtemp(prefix="ngp_refbench_")
    half = torch.empty(rgba8.shape, dtype=torch.int16, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    capi.check(lib.ngp_from_rgba32(s, rgba8.numel() // 4, C.c_void_p(rgba8.data_ptr()), C.c_void_p(half.data_ptr())))
    torch.cuda.synchronize()
    half.cpu().numpy().tofile(os.path.join(d, "images.f16"))
    del half
    np.ascontiguousarray(xf_np, np.float32).tofile(os.path.join(d, "xforms.f32"))
    with open(os.path.join(d, "meta.txt"), "w") as f:
        f.write(f"{args.views} {args.res} {args.res} {float(np.float32(fx))!r} {float(np.float32(fx))!r} 0.33\n")
    shutil.copy(os.path.join(ROOT, PKG, "configs", "base.json"), os.path.join(d, "config.json"))
    np.ascontiguousarray(scene.opengl_to_ngp(held_out_poses(scene)), np.float32).tofile(os.path.join(d, "test_xforms.f32"))
    return d


def harness(*argv, timeout=900):
    r = subprocess.run([REF_HARNESS] + [str(a) for a in argv], capture_output=True, text=True, timeout=timeout)
    if r.returncode != 0:
        return f"ref_harness {argv[0]} failed ({r.returncode}): {r.stderr[-300:]}"
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    return json.loads(lines[-1]) if lines else "ref_harness printed no result line"


def time_reference_cuda(args, d, steps, warmup, extras=True):
    """The UNMODIFIED reference (Testbed::train compiled for sm_100 by oracle/build_ref.sh) on the same scene, same base.json, the
    same train(16, batch) loop for EXACTLY warmup + steps steps (calls of 16 and one remainder call), timed by its own
    training_ms + training_prep_ms; then its render_frame on the held-out poses.  Returns its JSON dict, or a reason string."""
    if not (os.path.exists(REF_HARNESS) and os.access(REF_HARNESS, os.X_OK)):
        return "oracle/_ref/ref_harness not built"
    scene = importlib.import_module(PKG + ".scene")
    if extras:
        return harness("benchsteps", d, steps, warmup, args.batch, args.res, RENDER_W, RENDER_H, scene.CAMERA_ANGLE_X)
    return harness("benchsteps", d, steps, warmup, args.batch)


def c1_inputs():
    g = lambda shape, seed, lo, hi: (np.random.Generator(np.random.PCG64(seed)).random(shape, dtype=np.float32) * np.float32(hi - lo) + np.float32(lo))
    n = 1 << 18
    return n, g((n, 3), 1337, 0, 1), g(12196240, 43, -1e-4, 1e-4).astype(np.float16), (g((n, 32), 7, -1, 1) * np.float32(2 ** -7)).astype(np.float16)


def write_c1_inputs(d):
    n, xyz, table, dy = c1_inputs()
    xyz.tofile(os.path.join(d, "c1_xyz.f32")); table.view(np.uint16).tofile(os.path.join(d, "c1_table.f16")); dy.view(np.uint16).tofile(os.path.join(d, "c1_dy.f16"))


def reference_extras(args, d, r, capi, lib, scene):
    """render / psnr_db / c1 / c5 objects of the reference arm"""
    out = {}
    out["render"] = {"workload": f"configs[3]: {RENDER_W}x{RENDER_H} novel views, spp 1, {HELD_OUT} held-out poses", "mrays_per_s": r.get("render_mrays_per_s"),
                     "ms_per_frame": r.get("render_ms_per_frame"), "frames_timed": r.get("render_frames_timed"),
                     "timing": "wall clock around Testbed::render_frame, device synchronised on both sides"}
    try:
        gt = ground_truth_linear(args, capi, scene, lib, held_out_poses(scene))
        imgs = [np.fromfile(os.path.join(d, f"ref_bench_render_0_{p}.f32"), np.float32).reshape(args.res, args.res, 4) for p in range(HELD_OUT)]
        mean, per = psnr_db(imgs, gt)
        out["psnr_db"] = {"value": mean, "per_view": per, "at_step": int(r.get("i_step", 0)), "views": f"{HELD_OUT} held-out {args.res}x{args.res} poses, spp 1"}
    except Exception as e:
        out["psnr_db"] = {"unavailable": f"{type(e).__name__}: {e}"}
    try:                                        # configs[0]
        c1d = os.path.join(d, "c1")
        os.makedirs(c1d, exist_ok=True)
        write_c1_inputs(c1d)
        c1 = harness("c1", c1d, 100)
        out["c1"] = {"workload": C1_WORKLOAD, "fwd_us": c1["fwd_us"], "bwd_us": c1["bwd_us"]} if isinstance(c1, dict) else {"unavailable": c1}
    except Exception as e:
        out["c1"] = {"unavailable": f"{type(e).__name__}: {e}"}
    try:                                        # configs[4]: same images, the large-table cascaded model
        c5d = os.path.join(d, "c5")
        os.makedirs(c5d, exist_ok=True)
        for f in ("images.f16", "xforms.f32", "meta.txt"):
            os.symlink(os.path.join(d, f), os.path.join(c5d, f))
        write_config(os.path.join(c5d, "config.json"), **C5_EDITS)
        c5 = harness("benchsteps", c5d, args.c5_steps, C5_WARMUP, C5_BATCH)
        out["c5"] = ({"workload": C5_WORKLOAD, "ms_per_step": c5["ms_per_step"], "samples_per_s": c5["samples_per_s"], "steps": args.c5_steps, "warmup": C5_WARMUP,
                      "samples_marched_per_step": c5["marched_per_step"], "loss": c5["loss"]} if isinstance(c5, dict) else {"unavailable": c5})
    except Exception as e:
        out["c5"] = {"unavailable": f"{type(e).__name__}: {e}"}
    return out


def ours_c1(capi, lib):
    """configs[0] through the C ABI, CUDA events on the launching stream"""
    import ctypes as C
    import torch
    n, xyz, table, dy = c1_inputs()
    desc, n_entries = capi.make_grid_desc(device_scales=True)
    d_xyz, d_table, d_dy = torch.from_numpy(xyz).cuda(), torch.from_numpy(table.view(np.int16)).cuda(), torch.from_numpy(dy.view(np.int16)).cuda()
    d_out, d_grad = torch.zeros(n * 32, dtype=torch.int16, device="cuda"), torch.zeros(2 * n_entries, dtype=torch.int16, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    fwd = lambda: capi.check(lib.ngp_hash_encode_fwd(s, p(d_xyz), 3, n, p(d_table), C.byref(desc), p(d_out), 32, None))
    bwd = lambda: capi.check(lib.ngp_hash_encode_bwd(s, p(d_xyz), 3, n, p(d_dy), 32, C.byref(desc), p(d_grad), 1))
    out = {"workload": C1_WORKLOAD}
    for name, fn in (("fwd_us", fwd), ("bwd_us", bwd)):
        for _ in range(10):
            fn()
        ms = []
        for _ in range(100):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); b.synchronize()
            ms.append(a.elapsed_time(b))
        out[name] = float(np.median(ms)) * 1e3
    out["algorithmic_GBs"] = {"fwd": n * 588 / out["fwd_us"] / 1e3, "bwd": (n * 588 + 4 * n_entries) / out["bwd_us"] / 1e3}
    return out


def ours_render_and_psnr(args, tb, pyngp, capi, lib, scene, world, rank, barrier):
    """configs[3] + PSNR on the held-out poses, with the weights as they are after the timed training region"""
    ogl = held_out_poses(scene)
    tb.background_color = [0.0, 0.0, 0.0, 1.0]
    tb.fov = scene.CAMERA_ANGLE_X
    out = {}
    if rank == 0:
        imgs = []
        for m in ogl:
            tb.set_nerf_camera_matrix(m)
            imgs.append(np.array(tb.render(args.res, args.res, 1, False)))
        mean, per = psnr_db(imgs, ground_truth_linear(args, capi, scene, lib, ogl))
        out["psnr_db"] = {"value": mean, "per_view": per, "at_step": int(tb.training_buffer.i_step), "views": f"{HELD_OUT} held-out {args.res}x{args.res} poses, spp 1"}
    # tile-sharded frame: rank r marches the rows [r H / N, (r + 1) H / N) and nothing is exchanged
    if world > 1:
        tb.set_render_rows(round(rank * RENDER_H / world), round((rank + 1) * RENDER_H / world))
    tb.set_nerf_camera_matrix(ogl[0])
    tb.render_frames_ms(RENDER_W, RENDER_H, 1, True)                 # allocation + first touch
    ms, e2e_ms = [], []
    for rep in range(2):
        for m in ogl:
            tb.set_nerf_camera_matrix(m)
            barrier()
            t = tb.render_frames_ms(RENDER_W, RENDER_H, 1, True)
            if world > 1:
                import torch
                import torch.distributed as dist
                tt = torch.tensor([t], dtype=torch.float64, device="cuda")
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                t = float(tt[0])
            ms.append(t)
    for m in ogl:                                                      # the call a user makes: device -> host copy of the frame (band) included
        tb.set_nerf_camera_matrix(m)
        barrier()
        t0 = time.perf_counter()
        band = tb.render_band(RENDER_W, RENDER_H, 1, True) if world > 1 else tb.render(RENDER_W, RENDER_H, 1, True)
        e2e_ms.append((time.perf_counter() - t0) * 1e3)
    tb.set_render_rows(0, -1)
    frame_ms = float(np.mean(ms))
    out["render"] = {"workload": f"configs[3]: {RENDER_W}x{RENDER_H} novel views, spp 1, {HELD_OUT} held-out poses" + (f", {world} row bands (one per GPU, no collective)" if world > 1 else ""),
                     "mrays_per_s": RENDER_W * RENDER_H / frame_ms / 1e3, "ms_per_frame": frame_ms, "frames_timed": len(ms),
                     "timing": "wall clock around render_frame (ray generation -> tonemap), device synchronised on both sides, max over ranks",
                     "e2e_ms_per_frame_with_d2h": float(np.mean(e2e_ms)), "d2h_bytes_per_frame": int(np.asarray(band).nbytes)}
    return out


def ours_c5(args, pyngp, np_rgba8, ogl, scene):
    """configs[4] on the same images: a second Testbed with the large-table cascaded model, bounded to a few train() calls"""
    import tempfile
    import torch
    cfg = write_config(os.path.join(tempfile.mkdtemp(prefix="ngp_c5_"), "config.json"), **C5_EDITS)
    tb = pyngp.Testbed()
    tb.load_model_config(cfg)
    tb.load_training_data_from_memory(np_rgba8, ogl, scene.CAMERA_ANGLE_X, 0.33, [0.5, 0.5, 0.5])
    train_steps(tb, C5_WARMUP, C5_BATCH)
    torch.cuda.synchronize()
    samples, marched, rays, dev_ms = train_steps(tb, args.c5_steps, C5_BATCH)
    return {"workload": C5_WORKLOAD, "ms_per_step": dev_ms / args.c5_steps, "samples_per_s": samples / (dev_ms * 1e-3), "steps": args.c5_steps, "warmup": C5_WARMUP,
            "samples_marched_per_step": marched / args.c5_steps, "n_params": int(tb.n_params), "loss": tb.training_buffer.loss}


def run_reference_cuda(args):
    """--impl reference, CUDA kind: rank 0 times the reference's own CUDA build on GPU 0"""
    import shutil
    import torch
    pkg = importlib.import_module(PKG)
    capi, scene = pkg.capi, importlib.import_module(PKG + ".scene")
    lib = capi.load()
    torch.cuda.set_device(0)
    _, xf_np, fx, rgba8 = make_scene_on_device(args, capi, scene, lib)
    d = stage_reference_dataset(args, capi, lib, xf_np, fx, rgba8, scene)
    del rgba8
    try:
        clocks = ClockSampler(0)
        clocks.sample_now()
        clocks.start()
        r = time_reference_cuda(args, d, args.steps, args.warmup, extras=not args.no_extras)
        clocks.sample_now()
        clk = clocks.stop()
        if not isinstance(r, dict):
            return None, r
        out = {"impl": "reference", "reference_kind": "cuda: unmodified reference (Testbed::train, tiny-cuda-nn) compiled for sm_100 by oracle/build_ref.sh, run as oracle/_ref/ref_harness",
               "metric": METRIC, "value": r["samples_per_s"], "unit": UNIT, "n_gpus": 1, "steps": int(r["steps"]), "warmup": int(r["warmup_steps"]),
               "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
               "config": config_dict(args, 1), "clocks": clk,
               "e2e": {"value": r["samples_per_s"] * r["ms_per_step"] / r["wall_ms_per_step"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 12,
                       "wall_ms_per_step": r["wall_ms_per_step"]},
               "samples_marched_per_step": r["marched_per_step"], "rays_per_step": r["rays_per_step"], "loss": r["loss"]}
        if not args.no_extras:
            out.update(reference_extras(args, d, r, capi, lib, scene))
        return out, None
    finally:
        shutil.rmtree(d, ignore_errors=True)


def run_reference(args):
    """the reference's algorithm on the host cores: CPU oracle pipeline, all threads, bounded sample per step"""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cuda_line, why = None, "cpu kind requested"
    if args.reference_kind in ("auto", "cuda"):
        try:
            cuda_line, why = run_reference_cuda(args)
        except Exception as e:                                    # no GPU / library missing: fall back to the CPU port below
            cuda_line, why = None, f"{type(e).__name__}: {e}"
    from oracle import pipeline
    cores = os.cpu_count() or 1
    if cuda_line is not None:
        # the CPU port is still reported beside it, on a bounded sample
        cb = pipeline.time_cpu_steps(n_steps=4, n_rays=2048, warmup=1, n_views=8, res=200, batch=args.batch)
        cuda_line["cpu_baseline"] = {"value": cb["samples_per_s"], "unit": UNIT, "cores": cores, "kind": "port", "workload": "reduced (see sample)",
                                     "sample": f"REDUCED WORKLOAD: 4 steps of 2048 rays ({cb['samples_per_step']:.0f} samples/step, {cb['seconds']:.1f} s) on 8 procedural views 200x200, analytic occupancy, same model"}
        print(json.dumps(cuda_line), flush=True)
        return
    budget_s = 150.0
    trainer = pipeline.CpuTrainer(n_views=8, res=200, batch=args.batch)
    t0 = time.perf_counter()
    trainer.step(512)
    per_ray = (time.perf_counter() - t0) / 512
    total_steps = args.steps + args.warmup
    n_rays = int(min(8192, max(64, budget_s / total_steps / per_ray)))
    for _ in range(args.warmup):
        trainer.step(n_rays)
    samples, t0 = 0, time.perf_counter()
    for _ in range(args.steps):
        samples += trainer.step(n_rays)[1]
    dt = time.perf_counter() - t0
    value = samples / dt
    sample = f"{args.steps} steps of {n_rays} rays ({samples / args.steps:.0f} samples/step) on 8 procedural views 200x200, analytic occupancy grid, same model config"
    print(json.dumps({"impl": "reference", "reference_kind": f"cpu port (oracle/); cuda build not used: {why}", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
                      "config": config_dict(args, args.gpus),
                      "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
                      "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)


def train_steps(tb, n_steps, batch):
    """n_steps training steps in calls of CALL_STEPS; returns (post-compaction samples, marched samples, rays, device ms)"""
    global ENQUEUE_MS
    done, samples, marched, rays, dev_ms = 0, 0, 0, 0, 0.0
    while done < n_steps:
        k = min(CALL_STEPS, n_steps - done)
        r = tb.training_buffer.n_rays_per_batch
        tb.train(k, batch)
        b = tb.training_buffer
        samples += sum(min(c, batch) for c in b.n_compacted_coords_counter_cpu)
        marched += sum(b.n_coords_counter_cpu)
        rays += r * k
        dev_ms += b.training_ms + b.training_prep_ms
        if os.environ.get("NGP_PRINT_CALLS"):           # per-call trace: marched samples, rays and device ms per step
            print(f"call i_step={b.i_step} marched/step={sum(b.n_coords_counter_cpu) / k:.0f} rays={r} ms/step={(b.training_ms + b.training_prep_ms) / k:.4f} prep_ms={b.training_prep_ms:.3f}",
                  file=sys.stderr, flush=True)
        ENQUEUE_MS += b.enqueue_ms
        done += k
    return samples, marched, rays, dev_ms


def run_ours(args):
    import ctypes as C
    import torch
    import torch.distributed as dist
    pkg = importlib.import_module(PKG)
    capi, scene = pkg.capi, importlib.import_module(PKG + ".scene")
    lib = capi.load()                       # raises if the CUDA library is missing: there is no fallback path
    pyngp = pkg.load_pyngp()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    # ---- scene: rendered once on the device, then parked in PINNED HOST memory like a decoded dataset ----
    ogl, xf_np, fx, rgba8 = make_scene_on_device(args, capi, scene, lib)
    host_rgba8 = torch.empty(rgba8.shape, dtype=torch.uint8, pin_memory=True)
    host_rgba8.copy_(rgba8)
    del rgba8
    np_rgba8 = host_rgba8.numpy()

    tb = pyngp.Testbed()
    tb.load_model_config(os.path.join(ROOT, PKG, "configs", "base.json"))
    if os.environ.get("NGP_SERIAL_MARCH"):        # experiment switch: no look-ahead march
        tb.pipeline_march = False
    comm = 0
    if world > 1:
        box = [pyngp.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        comm = pyngp.nccl_comm_init(box[0], rank, world)
        tb.set_data_parallel(rank, world, comm, args.dp_mode == "peer")
    # host -> device upload of the dataset through the public API, timed on its second (warm) execution -- on a scratch Testbed, so
    # that the one that trains loads its data exactly once, like the reference arm (a second load re-seeds the sampler from an
    # advanced generator: another ray stream, and PSNR after a few hundred steps moves by tenths of a dB with the stream)
    scratch = pyngp.Testbed()
    scratch.load_model_config(os.path.join(ROOT, PKG, "configs", "base.json"))
    scratch.load_training_data_from_memory(np_rgba8, ogl, scene.CAMERA_ANGLE_X, 0.33, [0.5, 0.5, 0.5])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    scratch.load_training_data_from_memory(np_rgba8, ogl, scene.CAMERA_ANGLE_X, 0.33, [0.5, 0.5, 0.5])
    torch.cuda.synchronize()
    t_upload = time.perf_counter() - t0
    del scratch
    h2d_bytes = np_rgba8.nbytes + ogl.nbytes
    tb.load_training_data_from_memory(np_rgba8, ogl, scene.CAMERA_ANGLE_X, 0.33, [0.5, 0.5, 0.5])
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    train_steps(tb, args.warmup, args.batch)
    # The timed region.  A clock query normally returns in well under a millisecond, but now and then one blocks for ~200 ms inside
    # the driver and holds up this process's kernel launches with it (seen as host enqueue time 70x its usual value, 7.7 ms/step
    # instead of 0.55): such a region was disturbed by its own instrumentation and is measured again, at most twice, as the
    # timing rules allow; "attempts" in the clocks object says how often that happened.
    attempts = 0
    while True:
        attempts += 1
        clocks = ClockSampler(local_rank)
        barrier()
        launches0 = lib.ngp_launch_count()
        if rank == 0:
            clocks.sample_now()
        clocks.start()
        t0 = time.perf_counter()
        enq0 = ENQUEUE_MS
        samples, marched, rays, dev_ms = train_steps(tb, args.steps, args.batch)
        enqueue_ms = ENQUEUE_MS - enq0
        barrier()
        wall = time.perf_counter() - t0
        if rank == 0:
            clocks.sample_now()                 # the GPU has just finished the last step: still the clocks of the region
        clk = clocks.stop()
        launches = lib.ngp_launch_count() - launches0
        disturbed = (clk.get("max_query_ms") or 0.0) > 5.0 or enqueue_ms > 0.8 * dev_ms
        if world > 1:
            f = torch.tensor([1.0 if disturbed else 0.0], device="cuda")
            dist.all_reduce(f, op=dist.ReduceOp.MAX)
            disturbed = bool(f.item() > 0)
        if not disturbed or attempts >= 3:
            break
    clk["attempts"] = attempts
    st_timed = None
    if world > 1 and args.dp_mode == "peer":
        st_timed = np.zeros((2, 4096), np.uint64)          # read now: the profiling region below leaves stamps of its own
        capi.check(lib.ngp_peer_stamps_read(st_timed.ctypes.data_as(C.c_void_p)))

    # ---- the rest of BASELINE.json's metric, on the weights as the timed region left them: PSNR on held-out poses, configs[3] render ----
    loss_after_timed = tb.training_buffer.loss
    extras = {}
    if not args.no_extras:
        try:
            extras = ours_render_and_psnr(args, tb, pyngp, capi, lib, scene, world, rank, barrier)
        except Exception as e:
            if world > 1:
                raise                           # a rank that dropped out of the banded render would leave the others in a collective
            extras = {"render": {"unavailable": f"{type(e).__name__}: {e}"}}

    # ---- per-kernel durations, live, with CUDA events on the launching stream (same workload, a further timed region) ----
    prof_steps = min(args.steps, 64)
    capi.check(lib.ngp_kernel_timing_enable(1, 64 * prof_steps + 4096))
    tb.pipeline_march = False             # kernels one after another on one stream, so each duration is the kernel's own
    p_samples, p_marched, p_rays, _ = train_steps(tb, prof_steps, args.batch)
    tb.pipeline_march = not os.environ.get("NGP_SERIAL_MARCH")
    kt = capi.kernel_timing_collect()
    capi.check(lib.ngp_kernel_timing_enable(0, 0))

    peer_trace = None
    if world > 1 and args.dp_mode == "peer":
        # free-running trace of the timed region: per step, own work (previous "all shards landed" -> "my gradients published")
        # and exchange (published -> all shards landed), from the two globaltimer stamps the peer kernels leave
        st = st_timed
        e_hi = int(np.argmax(st[1]))                       # newest epoch slot
        idx = [(e_hi - k) % 4096 for k in range(min(args.steps, 2048))][::-1]
        pub, land = st[0][idx].astype(np.int64), st[1][idx].astype(np.int64)
        if os.environ.get("NGP_PEER_TRACE_DUMP"):
            np.save(f"{os.environ['NGP_PEER_TRACE_DUMP']}.rank{rank}.npy", np.stack([pub, land]))
        own = (pub[1:] - land[:-1]) * 1e-3
        exch = (land - pub) * 1e-3
        inner = np.array([k for k in range(len(own)) if own[k] < 5 * np.median(own)])     # drops the call boundaries (refresh + host)
        peer_trace = {"own_us_median": float(np.median(own[inner])), "exchange_us_median": float(np.median(exch)), "exchange_us_p90": float(np.percentile(exch, 90)),
                      "call_boundary_us_mean": float(np.mean(own[own >= 5 * np.median(own)])) if (own >= 5 * np.median(own)).any() else 0.0}
        tt = torch.tensor([peer_trace["own_us_median"], peer_trace["exchange_us_median"], peer_trace["exchange_us_p90"], peer_trace["call_boundary_us_mean"]], dtype=torch.float64, device="cuda")
        gathered = [torch.zeros_like(tt) for _ in range(world)]
        dist.all_gather(gathered, tt)
        peer_trace = {"per_rank_[own_us_median, exchange_us_median, exchange_us_p90, call_boundary_us_mean]": [[round(float(x), 1) for x in g] for g in gathered]}
    if args.timeline:
        capi.check(lib.ngp_kernel_timing_enable(1, 8192))
        train_steps(tb, 2 * CALL_STEPS, args.batch)
        tl = capi.kernel_timing_timeline()
        capi.check(lib.ngp_kernel_timing_enable(0, 0))
        json.dump(tl, open(f"{args.timeline}.rank{rank}.json", "w"))
    if world > 1:
        t = torch.tensor([dev_ms, wall * 1e3, t_upload * 1e3], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, wall, t_upload = float(t[0]), float(t[1]) / 1e3, float(t[2]) / 1e3
        c = torch.tensor([samples, marched, rays], dtype=torch.float64, device="cuda")
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        samples, marched, rays = float(c[0]), float(c[1]), float(c[2])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    hbm_peak, tf_peak, peak_src = measured_peaks()
    # measured ceilings of the L1TEX / L2 path ("L2 peak is not in MEASURED_PEAKS.json -> must be measured on the box"):
    # coalesced 16-byte reads of a 32 MB L2-resident buffer, and scattered gathers / reductions with every lane on its own line
    probe = {}
    try:
        scratch = torch.zeros(32 << 20, dtype=torch.uint8, device="cuda")
        sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        for kind, key in ((0, "l2_stream_read_GBs"), (1, "scattered_gather_4B_G_per_s"), (2, "scattered_gather_16B_G_per_s"),
                          (3, "scattered_red_f16x2_G_per_s"), (4, "scattered_red_v4_f16x2_G_per_s")):
            val = C.c_double(0.0)
            capi.check(lib.ngp_l2_probe(sp, kind, C.c_void_p(scratch.data_ptr()), scratch.numel(), 5, C.byref(val)))
            probe[key] = val.value
        del scratch
    except Exception as e:
        probe = {"unavailable": str(e)}
    total_kernel_ms = sum(v[0] for v in kt.values())
    # dominant kernel of the STEP: the march kernels run one step ahead on the look-ahead stream, underneath the others, so they
    # are listed in kernels_ms_per_step (this profiling region runs them serially) but cannot be the kernel that bounds the step
    on_path = {k: v for k, v in kt.items() if not k.startswith("march_")} or kt
    name, (k_ms, k_n) = max(on_path.items(), key=lambda kv: kv[1][0])
    m, b, r = p_marched / prof_steps, p_samples / prof_steps, p_rays / prof_steps
    first = args.warmup + args.steps                                     # index of the first profiled step
    g = sum((1 << 22) if (first + k) < 256 else (1 << 21) for k in range(0, prof_steps, CALL_STEPS)) / prof_steps
    work = kernel_work(name, m, b, r, tb.n_params, g, world, args.dp_mode == "peer")
    launches_per_step = k_n / prof_steps
    total_kernel_ms = sum(v[0] for v in on_path.values())
    roofline = {"kernel": name, "share_of_kernel_time": k_ms / total_kernel_ms, "avg_launch_us": k_ms / k_n * 1e3, "launches_per_step": launches_per_step}
    if work:
        bound, per_step = work
        per_launch = per_step / launches_per_step
        achieved = per_launch / (k_ms / k_n * 1e-3)
        l2_peak = probe.get("l2_stream_read_GBs")
        if bound == "l2" and not l2_peak:
            bound = "hbm"                       # the probe failed: fall back to the conservative denominator
        peak = hbm_peak if bound == "hbm" else l2_peak if bound == "l2" else tf_peak
        achieved = achieved / 1e12 if bound == "tensor" else achieved / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            # measured DRAM bytes per unit (ncu --set full, profiles/traffic.json) x the units of this run's average launch
            t = json.load(open(tp)).get(name)
            units = {"hash_encode_fwd_kernel": m + g, "hash_encode_bwd_kernel": b, "adam_ema_kernel": tb.n_params}.get(name)
            if t and units:
                traffic = t["dram_bytes_per_unit"] * units / launches_per_step
        roofline.update({"bound": bound, "achieved": achieved, "peak": peak, "unit": "TFLOP/s" if bound == "tensor" else "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src, "algorithmic_per_launch": per_launch})
        if bound == "l2":
            roofline["peak_source"] = "ngp_l2_probe kind 0 measured in this run: coalesced 16-byte reads of a 32 MB L2-resident buffer (MEASURED_PEAKS.json has no L2 figure)"
            roofline["hbm"] = {"peak": hbm_peak, "frac": achieved / hbm_peak, "peak_source": peak_src,
                               "note": "HBM copy peak for comparison; measured DRAM traffic (`traffic`) is a fraction of the algorithmic bytes: the table stays in L2"}
        if name == "adam_ema_kernel" and world > 1 and args.dp_mode == "peer":
            roofline["note"] = "fused exchange kernel (reduce-scatter + Adam + all-gather over NVLink peer memory): its duration includes the wait for the slowest rank's gradient flag"
    roofline["kernels_ms_per_step"] = {k: v[0] / prof_steps for k, v in sorted(kt.items(), key=lambda kv: -kv[1][0])}

    out = {"metric": METRIC, "value": samples / (dev_ms * 1e-3), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
           "config": config_dict(args, args.gpus), "clocks": clk, "gpu_launches": int(launches),
           "e2e": {"value": samples / (wall + t_upload), "unit": UNIT, "h2d_bytes_per_step": h2d_bytes / args.steps,
                   "d2h_bytes_per_step": (3 * 4 * CALL_STEPS + 4) / CALL_STEPS, "wall_ms_per_step": wall / args.steps * 1e3, "dataset_upload_ms": t_upload * 1e3,
                   "host_enqueue_ms_per_step": enqueue_ms / args.steps},
           "roofline": roofline,
           "samples_marched_per_step": marched / args.steps / max(world, 1), "rays_per_step": rays / args.steps / max(world, 1),
           "loss": loss_after_timed, "loss_scope": "rank 0's rays, last train() call of the timed region"}
    out.update(extras)
    # ---- the encoder kernels against the measured ceilings, access by access ----
    try:
        if "unavailable" in probe:
            raise RuntimeError(probe["unavailable"])
        k = roofline["kernels_ms_per_step"]
        # accesses the kernels issue per point and level: 4 x-pairs, 3 of 4 served by one 16-byte access, the 4th by two 4-byte ones
        per_point = 16 * 5
        fwd = (m + g) * per_point / (k["hash_encode_fwd_kernel"] * 1e-3) / 1e9 if k.get("hash_encode_fwd_kernel") else None
        bwd = b * per_point / (k["hash_encode_bwd_kernel"] * 1e-3) / 1e9 if k.get("hash_encode_bwd_kernel") else None
        out["l2_roofline"] = {"probes": probe, "table": "32 MB scratch, L2-resident; every lane of a warp hits its own line (worst case of the fine levels)",
                              "accesses_per_point": per_point,
                              "reading": "a frac above 1 means the kernel beats the all-lanes-scattered ceiling because neighbouring samples of a ray share lines on the coarse levels; "
                                         "sector_GBs = accesses x one 32-byte L2 sector each, against the coalesced L2 read rate",
                              "sector_bytes": 32,
                              "hash_encode_fwd": {"achieved_G_accesses_per_s": fwd, "frac_of_scattered_gather_16B": fwd / probe["scattered_gather_16B_G_per_s"] if fwd else None,
                                                  "sector_GBs": fwd * 32 if fwd else None, "frac_of_l2_stream_read": fwd * 32 / probe["l2_stream_read_GBs"] if fwd else None},
                              "hash_encode_bwd": {"achieved_G_accesses_per_s": bwd, "frac_of_scattered_red_v4": bwd / probe["scattered_red_v4_f16x2_G_per_s"] if bwd else None,
                                                  "sector_GBs": bwd * 32 if bwd else None, "frac_of_l2_stream_read": bwd * 32 / probe["l2_stream_read_GBs"] if bwd else None,
                                                  "note": "before the in-warp merge of lanes that share a cell; the coarse levels issue fewer reductions than counted"}}
    except Exception as e:                      # evidence, not part of the contract line
        out["l2_roofline"] = {"unavailable": str(e)}
    if peer_trace:
        out["peer_trace"] = peer_trace
    if args.gpus == 1 and not args.no_extras:
        for key, fn in (("c1", lambda: ours_c1(capi, lib)), ("c5", lambda: ours_c5(args, pyngp, np_rgba8, ogl, scene))):
            try:
                if key == "c5":
                    del tb                      # the configs[1] model and its buffers make room for the 2^22 / 2^20 configuration
                out[key] = fn()
            except Exception as e:
                out[key] = {"unavailable": f"{type(e).__name__}: {e}"}
    if args.gpus == 1 and not args.no_cpu_baseline:
        from oracle import pipeline                      # the CPU restatement, timed as the baseline only
        cb = pipeline.time_cpu_steps(n_steps=4, n_rays=2048, warmup=1, n_views=8, res=200, batch=args.batch)
        out["cpu_baseline"] = {"value": cb["samples_per_s"], "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port", "workload": "reduced (see sample)",
                               "sample": f"REDUCED WORKLOAD: 4 steps of 2048 rays ({cb['samples_per_step']:.0f} samples/step, {cb['seconds']:.1f} s) on 8 procedural views 200x200, analytic occupancy, same model"}
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
