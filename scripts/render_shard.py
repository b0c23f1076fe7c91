#!/usr/bin/env python
"""BASELINE configs[3]: 1920x1080 novel-view render, image tiles (row bands) sharded over the GPUs of one box, no collective.

Rank 0 trains briefly and writes a snapshot; every rank loads it (same weights, same occupancy grid), renders its band of rows
through pyngp.Testbed.render and times it; the bands are then gathered to rank 0 ONLY to check them bit for bit against rank 0's
own full-frame render -- the product path needs no exchange, each rank owns its rows.

usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/render_shard.py [train_calls=40] [out.json]
       (also runs as a plain single process: one band = the whole frame)"""
import importlib
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("instant-ngp-nerf-minimal_b200")
pkg.capi.load()
pyngp = pkg.load_pyngp()
W, H, FRAMES = 1920, 1080, 8


def main():
    calls = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "render_shard.json")
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    snap = os.path.join("/tmp", f"render_shard_{os.environ.get('MASTER_PORT', '0')}.msgpack")
    cfg = os.path.join(ROOT, "instant-ngp-nerf-minimal_b200", "configs", "base.json")

    tb = pyngp.Testbed()
    tb.load_model_config(cfg)
    tb.load_procedural_training_data(100, 800, 800, 2024, 0.33, [0.5, 0.5, 0.5])      # same cameras on every rank
    if rank == 0:
        for _ in range(calls):
            tb.train(16, 1 << 18)
        tb.save_snapshot(snap, False)
    if world > 1:
        dist.barrier()
    tb.load_snapshot(snap)                      # rank 0 too: every rank renders from the same bytes
    tb.set_camera_to_training_view(3)
    tb.background_color = [0, 0, 0, 1]

    def timed(frames, band_only=False):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(frames):
            img = tb.render_band(W, H, 1, True) if band_only else tb.render(W, H, 1, True)
        return (time.perf_counter() - t0) / frames, np.array(img)

    r0, r1 = round(rank * H / world), round((rank + 1) * H / world)
    tb.set_render_rows(r0, r1)
    timed(2, True)                              # warm-up: allocations
    if world > 1:
        dist.barrier()
    band_s, band = timed(FRAMES, True)          # render_band: only this rank's rows are cleared, marched, tonemapped and copied out
    t = torch.tensor([band_s], dtype=torch.float64, device="cuda")
    rows = torch.from_numpy(np.ascontiguousarray(band)).cuda()
    gathered = None
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sizes = [round((k + 1) * H / world) - round(k * H / world) for k in range(world)]
        gathered = [torch.empty((s, W, 4), dtype=torch.float32, device="cuda") for s in sizes] if rank == 0 else None
        if len(set(sizes)) == 1:
            dist.gather(rows, gathered, dst=0)
        else:                                   # ragged bands: one broadcast per band
            gathered = []
            for k in range(world):
                buf = rows if k == rank else torch.empty((sizes[k], W, 4), dtype=torch.float32, device="cuda")
                dist.broadcast(buf, src=k)
                gathered.append(buf)
    if rank == 0:
        tb.set_render_rows(0, -1)
        timed(2)
        full_s, full = timed(FRAMES)
        stitched = torch.cat(gathered, 0).cpu().numpy() if gathered is not None else band
        identical = bool(np.array_equal(stitched.view(np.uint32), full.view(np.uint32)))
        res = {"workload": "configs[3]: 1920x1080 novel view, spp 1, occupancy-grid skip + early ray termination, row bands over the GPUs",
               "n_gpus": world, "train_steps": calls * 16, "rows_per_gpu": H / world,
               "full_frame_1gpu_ms": full_s * 1e3, "full_frame_1gpu_mrays_per_s": W * H / full_s / 1e6,
               "sharded_ms_max_over_ranks": float(t.item()) * 1e3, "sharded_mrays_per_s": W * H / float(t.item()) / 1e6,
               "timing": "wall clock around Testbed.render / render_band (ray generation -> tonemap -> copy of the frame / band to the host), mean of %d frames after 2 warm-up frames" % FRAMES,
               "bands_bit_identical_to_full_frame": identical, "alpha_mean": float(full[..., 3].mean())}
        os.makedirs(os.path.dirname(out), exist_ok=True)
        json.dump(res, open(out, "w"), indent=1)
        print(json.dumps(res))
        assert identical, "stitched bands differ from the full frame"
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
