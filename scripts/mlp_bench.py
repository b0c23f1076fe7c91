"""Times the fused MLP kernels alone through the C ABI (CUDA events on the launching stream, median of 50 after 5 warm-ups):
ngp_mlp_inference on m samples (forward-only kernel) and ngp_mlp_train on b samples.  usage: python scripts/mlp_bench.py [m] [b]"""
import ctypes as C
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
capi = importlib.import_module("instant-ngp-nerf-minimal_b200").capi
lib = capi.load()


def timed(fn, iters=20, burst=20):
    """median over `iters` of (time of `burst` back-to-back launches) / burst: a single launch between two events would mostly
    measure the host's enqueue latency (the queue runs dry), not the kernel"""
    for _ in range(5):
        fn()
    ms = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(burst):
            fn()
        b.record(); b.synchronize()
        ms.append(a.elapsed_time(b) / burst)
    return float(np.median(ms)) * 1e3


def main():
    m = int(sys.argv[1]) if len(sys.argv) > 1 else 360000
    b = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 18
    g = torch.Generator(device="cuda").manual_seed(1)
    n = max(m, b)
    enc = (torch.rand(n * 32, device="cuda", generator=g) - 0.5).half()
    coords = torch.rand(n * 7, device="cuda", generator=g)
    w = ((torch.rand(capi.NGP_MLP_PARAMS, device="cuda", generator=g) - 0.5) * 0.5).half()
    dl = ((torch.rand(n * 4, device="cuda", generator=g) - 0.5) * 0.1).half()
    out = torch.zeros(n * 4, dtype=torch.half, device="cuda")
    denc = torch.zeros(n * 32, dtype=torch.half, device="cuda")
    wg = torch.zeros(capi.NGP_MLP_PARAMS, dtype=torch.float32, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    res = {"m": m, "b": b}
    res["infer_us"] = timed(lambda: capi.check(lib.ngp_mlp_inference(s, p(enc), p(coords), m, p(w), p(out))))
    res["train_us"] = timed(lambda: capi.check(lib.ngp_mlp_train(s, p(enc), p(coords), b, p(w), p(dl), p(out), p(denc), p(wg))))
    res["infer_TFLOPs"] = m * 18816 / res["infer_us"] / 1e6
    res["train_TFLOPs"] = b * 54400 / res["train_us"] / 1e6
    print(json.dumps(res))


if __name__ == "__main__":
    main()
