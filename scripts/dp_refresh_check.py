#!/usr/bin/env python
"""Data-parallel occupancy refresh, sharded over the ranks (host/testbed_train.cpp train_prep_nerf): correctness and timing on N GPUs.

  (a) the FIRST refresh runs on the initial weights, which every process derives from the same seed, so it is deterministic: the
      density grid after train(1) under data parallelism (each rank evaluates 1/N of the samples, one ncclAllReduce(max)) must be
      BIT-IDENTICAL to the grid a single process computes with the replicated refresh;
  (b) 288 more steps (so the refresh has used both the uniform and the thresholded sample set): every rank must hold bit-identical
      density grids and fp16 parameters (sha256 gathered over the ranks), and the loss must have fallen;
  (c) training_prep_ms of the sharded and of the replicated refresh, alternating, on the same process group.

usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P scripts/dp_refresh_check.py [out.json] [nccl|peer]"""
import hashlib
import importlib
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("instant-ngp-nerf-minimal_b200")
pkg.capi.load()
pyngp = pkg.load_pyngp()
BATCH = 1 << 18


def snapshot_fields(tb, path):
    import msgpack
    tb.save_snapshot(path, False)
    snap = msgpack.unpackb(open(path, "rb").read(), raw=False, strict_map_key=False)["snapshot"]
    return np.frombuffer(snap["density_grid"], np.uint32), np.frombuffer(snap["params_binary"], np.uint16)


def new_testbed():
    tb = pyngp.Testbed()
    tb.load_model_config(os.path.join(ROOT, "instant-ngp-nerf-minimal_b200", "configs", "base.json"))
    tb.load_procedural_training_data(20, 400, 400, 2024, 0.33, [0.5, 0.5, 0.5])
    return tb


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "dp_refresh_check.json")
    mode = sys.argv[2] if len(sys.argv) > 2 else "peer"
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    assert world > 1, "run under torch.distributed.run with at least 2 ranks"
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    tmp = f"/tmp/dp_refresh_{os.environ.get('MASTER_PORT', '0')}_{rank}.msgpack"

    # (a) single-process replicated refresh on the initial weights
    single = new_testbed()
    single.train(1, BATCH)
    want_grid, _ = snapshot_fields(single, tmp)
    del single

    tb = new_testbed()
    box = [pyngp.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    comm = pyngp.nccl_comm_init(box[0], rank, world)
    tb.set_data_parallel(rank, world, comm, mode == "peer")
    assert tb.shard_grid_refresh
    tb.train(1, BATCH)
    first_loss = tb.training_buffer.loss
    got_grid, _ = snapshot_fields(tb, tmp)
    n_diff = int(np.count_nonzero(got_grid != want_grid))
    assert n_diff == 0, f"rank {rank}: sharded first refresh differs from the replicated one in {n_diff} cells"
    assert (want_grid.view(np.float32) > 0).any()

    # (b) both sample sets, then agreement over the ranks
    for _ in range(18):
        tb.train(16, BATCH)
    assert tb.training_buffer.i_step == 1 + 18 * 16
    grid, params = snapshot_fields(tb, tmp)
    digest = (hashlib.sha256(grid.tobytes()).hexdigest(), hashlib.sha256(params.tobytes()).hexdigest())
    digests = [None] * world
    dist.all_gather_object(digests, digest)
    assert all(d == digests[0] for d in digests), f"ranks disagree after the sharded refreshes: {digests}"
    loss = tb.training_buffer.loss
    assert np.isfinite(loss) and loss < 0.5 * first_loss, (first_loss, loss)
    occupied = float(np.mean(grid.view(np.float32) > 0.01))

    # (c) timing, alternating the two schedules on the same process group
    prep = {True: [], False: []}
    for k in range(12):
        tb.shard_grid_refresh = (k % 2 == 0)
        tb.train(16, BATCH)
        prep[tb.shard_grid_refresh].append(tb.training_buffer.training_prep_ms)
    tb.shard_grid_refresh = True
    ms = torch.tensor([float(np.median(prep[True])), float(np.median(prep[False]))], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        res = {"world": world, "mode": mode, "first_refresh_bit_identical_to_single_process": True, "ranks_agree_after_steps": tb.training_buffer.i_step,
               "loss_first": first_loss, "loss": loss, "occupied_fraction": occupied,
               "refresh_ms_sharded_max_over_ranks": float(ms[0]), "refresh_ms_replicated_max_over_ranks": float(ms[1])}
        os.makedirs(os.path.dirname(out), exist_ok=True)
        json.dump(res, open(out, "w"), indent=1)
        print(json.dumps(res))
    dist.barrier()
    del tb
    pyngp.nccl_comm_destroy(comm)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
