#!/usr/bin/env python
"""Where does a training step's time go once the workload is stationary?  After `warm` steps: device ms/step with the march
pipelined, serial, and serial with the per-kernel event table -- the difference between the serial step and the sum of its
kernels is what launch boundaries, memsets and event waits cost.
usage: python scripts/step_breakdown.py [warm_calls=40] [calls=8] [out.json]"""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("instant-ngp-nerf-minimal_b200")
capi, pyngp = pkg.capi, pkg.load_pyngp()
lib = capi.load()


def run(tb, calls):
    ms = prep = 0.0
    samples = marched = 0
    for _ in range(calls):
        tb.train(16, 1 << 18)
        b = tb.training_buffer
        ms += b.training_ms
        prep += b.training_prep_ms
        samples += b.measured_batch_size * 16
        marched += b.measured_batch_size_before_compaction * 16
    n = 16 * calls
    return {"train_ms_per_step": ms / n, "prep_ms_per_step": prep / n, "samples_per_step": samples / n, "marched_per_step": marched / n}


def main():
    warm = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    calls = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    out = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "gpurun_out", "step_breakdown.json")
    tb = pyngp.Testbed()
    tb.load_model_config(os.path.join(ROOT, "instant-ngp-nerf-minimal_b200", "configs", "base.json"))
    tb.load_procedural_training_data(100, 800, 800, 2024, 0.33, [0.5, 0.5, 0.5])
    for _ in range(warm):
        tb.train(16, 1 << 18)
    res = {"warm_steps": warm * 16}
    tb.pipeline_march = True
    res["pipelined"] = run(tb, calls)
    tb.pipeline_march = False
    res["serial"] = run(tb, calls)
    l0 = lib.ngp_launch_count()
    capi.check(lib.ngp_kernel_timing_enable(1, 64 * 16 * calls + 4096))
    res["serial_timed"] = run(tb, calls)
    kt = capi.kernel_timing_collect()
    capi.check(lib.ngp_kernel_timing_enable(0, 0))
    res["launches_per_step"] = (lib.ngp_launch_count() - l0) / (16 * calls)
    res["kernels_us_per_step"] = {k: v[0] / (16 * calls) * 1e3 for k, v in sorted(kt.items(), key=lambda kv: -kv[1][0])}
    res["kernel_sum_us_per_step"] = sum(res["kernels_us_per_step"].values())
    tb.pipeline_march = True
    res["pipelined_again"] = run(tb, calls)
    os.makedirs(os.path.dirname(out), exist_ok=True)
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
