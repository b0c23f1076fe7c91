#!/usr/bin/env python
"""BASELINE configs[0]: hash-grid encode forward + backward on 2^18 uniformly random points (L=16, F=2, T=2^19), through the C ABI,
timed with CUDA events on the launching stream, next to the CPU oracle on the host cores and to the measured ceilings of the
L1TEX / L2 path (ngp_l2_probe).  Random points are the worst case for the encoder: no two lanes of a warp share a cell above
the coarsest levels, so nothing coalesces and the backward merge finds no runs.
usage: python tests/c1_bench.py [out.json]"""
import ctypes as C
import importlib
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("instant-ngp-nerf-minimal_b200")
capi = pkg.capi
lib = capi.load()
import oracle                                    # the checker and the CPU baseline, never the product
from util import rng_uniform

N, ITERS, WARM = 1 << 18, 100, 10


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "c1_bench.json")
    L = oracle.GridLevels(16, 19, 16)
    desc, n_entries = capi.make_grid_desc(device_scales=True)          # the scales the reference's kernels evaluate (grid.h:157)
    L.scales[:] = np.frombuffer(bytes(desc.scales), np.float32)[:16]    # the oracle restates the same arithmetic with the same inputs
    L.resolutions[:] = np.frombuffer(bytes(desc.resolutions), np.uint32)[:16]
    xyz = rng_uniform((N, 3), 1337)
    table = rng_uniform(L.n_params, 43, -1e-4, 1e-4).astype(np.float16)
    dy = (rng_uniform((N, 32), 7, -1, 1) * np.float32(2 ** -7)).astype(np.float16)
    d_xyz = torch.from_numpy(xyz).cuda()
    d_table = torch.from_numpy(table.view(np.int16)).cuda()
    d_dy = torch.from_numpy(dy.view(np.int16)).cuda()
    d_out = torch.zeros(N * 32, dtype=torch.int16, device="cuda")
    d_grad = torch.zeros(L.n_params, dtype=torch.int16, device="cuda")
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device="cuda")      # larger than L2: written between timed iterations
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())

    def fwd():
        capi.check(lib.ngp_hash_encode_fwd(s, p(d_xyz), 3, N, p(d_table), C.byref(desc), p(d_out), 32, None))

    def bwd():
        capi.check(lib.ngp_hash_encode_bwd(s, p(d_xyz), 3, N, p(d_dy), 32, C.byref(desc), p(d_grad), 1))

    def timed(fn, flush_l2):
        for _ in range(WARM):
            fn()
        ms = []
        for _ in range(ITERS):
            if flush_l2:
                flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); b.synchronize()
            ms.append(a.elapsed_time(b))
        return float(np.median(ms)) * 1e3

    res = {"workload": "configs[0]: hash-grid encode fwd+bwd, 2^18 random points, L=16 F=2 T=2^19 (12 196 240 fp16 parameters)", "n_points": N, "iterations": ITERS}
    for name, fn in (("fwd", fwd), ("bwd", bwd)):
        hot, cold = timed(fn, False), timed(fn, True)
        alg = N * 588 + (2 * L.n_params if name == "bwd" else 0)
        res[name] = {"us_table_in_L2": hot, "us_after_L2_flush": cold, "algorithmic_bytes": alg, "algorithmic_GBs_table_in_L2": alg / hot / 1e3,
                     "G_accesses_per_s_table_in_L2": N * 80 / hot / 1e3}
    # parity of exactly these launches against the oracle (indices are covered by the tests; here encodings and gradient sums)
    ref = oracle.hash_encode_fwd(xyz, 3, N, table.view(np.uint16), L)
    got = d_out.cpu().numpy().view(np.float16).reshape(N, 32)
    res["fwd"]["max_fp16_ulp_vs_oracle"] = int(np.abs(got.view(np.int16).astype(np.int32) - np.asarray(ref).view(np.int16).astype(np.int32)).max())
    def best_of(fn, k=3):
        ts = []
        for _ in range(k):
            t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
        return min(ts)
    t_f = best_of(lambda: oracle.hash_encode_fwd(xyz, 3, N, table.view(np.uint16), L))
    t_b = best_of(lambda: oracle.hash_encode_bwd(xyz, 3, N, dy, L))
    res["cpu_oracle"] = {"cores": os.cpu_count(), "fwd_ms": t_f * 1e3, "bwd_ms": t_b * 1e3, "kind": "port"}
    # ---- the reference's own GridEncoding on the same three files (oracle/_ref/ref_harness c1), outputs compared, times beside ours ----
    harness = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
    if os.path.exists(harness):
        import subprocess, tempfile
        with tempfile.TemporaryDirectory() as d:
            xyz.tofile(os.path.join(d, "c1_xyz.f32")); table.view(np.uint16).tofile(os.path.join(d, "c1_table.f16")); dy.view(np.uint16).tofile(os.path.join(d, "c1_dy.f16"))
            r = subprocess.run([harness, "c1", d, str(ITERS)], capture_output=True, text=True, timeout=600)
            line = [l for l in r.stdout.splitlines() if l.startswith("{")]
            if r.returncode == 0 and line:
                ref_t = json.loads(line[-1])
                ref_enc = np.fromfile(os.path.join(d, "ref_c1_enc.f16"), np.float16).reshape(N, 32)
                ref_grad = np.fromfile(os.path.join(d, "ref_c1_grad.f16"), np.float16).astype(np.float32)
                bwd()
                torch.cuda.synchronize()
                our_grad = d_grad.cpu().numpy().view(np.float16).astype(np.float32)
                ulp = np.abs(got.view(np.int16).astype(np.int32) - ref_enc.view(np.int16).astype(np.int32))
                gscale = np.abs(ref_grad).max()
                gf, rf = got.astype(np.float32), ref_enc.astype(np.float32)
                bad = np.argwhere(ulp > 0)
                examples = [dict(point=int(i), feature=int(f), ours=float(gf[i, f]), reference=float(rf[i, f]), oracle=float(np.asarray(ref).view(np.float16).reshape(N, 32)[i, f]))
                            for i, f in bad[:6]]
                per_level_mismatch = [float((ulp[:, 2 * l:2 * l + 2] > 0).mean()) for l in range(16)]
                res["reference_grid_encoding"] = dict(ref_t, fwd_speedup=ref_t["fwd_us"] / res["fwd"]["us_table_in_L2"], bwd_speedup=ref_t["bwd_us"] / res["bwd"]["us_table_in_L2"],
                                                      encodings_max_fp16_ulp_vs_reference=int(ulp.max()), encodings_bit_identical_fraction=float((ulp == 0).mean()),
                                                      encodings_max_abs_diff_over_max=float(np.abs(gf - rf).max() / np.abs(rf).max()), mismatch_examples=examples,
                                                      mismatch_fraction_per_level=per_level_mismatch,
                                                      grad_max_abs_diff_over_max=float(np.abs(our_grad - ref_grad).max() / gscale),
                                                      grad_same_nonzero_pattern=bool(np.array_equal(our_grad != 0, ref_grad != 0)))
            else:
                res["reference_grid_encoding"] = {"unavailable": (r.stderr or r.stdout)[-300:]}
    scratch = torch.zeros(32 << 20, dtype=torch.uint8, device="cuda")
    probes = {}
    for kind, key in ((0, "l2_stream_read_GBs"), (2, "scattered_gather_16B_G_per_s"), (4, "scattered_red_v4_f16x2_G_per_s")):
        v = C.c_double(0.0)
        capi.check(lib.ngp_l2_probe(s, kind, p(scratch), scratch.numel(), 5, C.byref(v)))
        probes[key] = v.value
    res["ceilings"] = probes
    res["fwd"]["frac_of_scattered_gather_ceiling"] = res["fwd"]["G_accesses_per_s_table_in_L2"] / probes["scattered_gather_16B_G_per_s"]
    res["bwd"]["frac_of_scattered_red_ceiling"] = res["bwd"]["G_accesses_per_s_table_in_L2"] / probes["scattered_red_v4_f16x2_G_per_s"]
    os.makedirs(os.path.dirname(out), exist_ok=True)
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
