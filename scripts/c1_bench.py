#!/usr/bin/env python
"""BASELINE configs[0]: hash-grid encode forward + backward on 2^18 uniformly random points (L=16, F=2, T=2^19), through the C ABI,
timed with CUDA events on the launching stream, next to the CPU oracle on the host cores and to the measured ceilings of the
L1TEX / L2 path (ngp_l2_probe).  Random points are the worst case for the encoder: no two lanes of a warp share a cell above
the coarsest levels, so nothing coalesces and the backward merge finds no runs.
usage: python scripts/c1_bench.py [out.json]"""
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
    desc, n_entries = capi.make_grid_desc()
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
