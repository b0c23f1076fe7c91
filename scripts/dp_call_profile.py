#!/usr/bin/env python
"""Where a data-parallel train(16) call spends its time, per rank: wall clock around the call, the two CUDA-event spans the Testbed
keeps (refresh = training_prep_ms, the 16 steps = training_ms) and the host's enqueue time.
usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P scripts/dp_call_profile.py [out.json] [calls=48] [shard_refresh=1]"""
import importlib, json, os, sys, time
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("instant-ngp-nerf-minimal_b200")
pkg.capi.load()
pyngp = pkg.load_pyngp()


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "dp_call_profile.json")
    calls = int(sys.argv[2]) if len(sys.argv) > 2 else 48
    shard = (sys.argv[3] != "0") if len(sys.argv) > 3 else True
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    tb = pyngp.Testbed()
    tb.load_model_config(os.path.join(ROOT, "instant-ngp-nerf-minimal_b200", "configs", "base.json"))
    tb.load_procedural_training_data(100, 800, 800, 2024, 0.33, [0.5, 0.5, 0.5])
    comm = 0
    if world > 1:
        box = [pyngp.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        comm = pyngp.nccl_comm_init(box[0], rank, world)
        tb.set_data_parallel(rank, world, comm, True)
        tb.shard_grid_refresh = shard
    for _ in range(8):
        tb.train(16, 1 << 18)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    rows = []
    t_begin = time.perf_counter()
    for _ in range(calls):
        t0 = time.perf_counter()
        tb.train(16, 1 << 18)
        t1 = time.perf_counter()
        b = tb.training_buffer
        rows.append(((t1 - t0) * 1e3, b.training_prep_ms, b.training_ms, b.enqueue_ms))
    total = (time.perf_counter() - t_begin) * 1e3
    r = np.array(rows)
    mine = {"rank": rank, "ms_per_call_wall": total / calls, "median": dict(zip(["wall", "refresh_gpu", "steps_gpu", "enqueue_host"], np.median(r, axis=0).round(3).tolist())),
            "mean": dict(zip(["wall", "refresh_gpu", "steps_gpu", "enqueue_host"], r.mean(axis=0).round(3).tolist())),
            "p90": dict(zip(["wall", "refresh_gpu", "steps_gpu", "enqueue_host"], np.percentile(r, 90, axis=0).round(3).tolist()))}
    allr = [None] * world
    if world > 1:
        dist.all_gather_object(allr, mine)
    else:
        allr = [mine]
    if rank == 0:
        res = {"world": world, "shard_grid_refresh": shard, "calls": calls, "ranks": allr, "cpu_count": os.cpu_count()}
        json.dump(res, open(out, "w"), indent=1)
        for a in allr:
            print(a["rank"], a["ms_per_call_wall"], a["median"], a["p90"])
        print("cpus", os.cpu_count())
    if world > 1:
        dist.barrier()
        del tb
        pyngp.nccl_comm_destroy(comm)
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
