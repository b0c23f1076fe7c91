"""CPU, world_size 2 over gloo: the host-side logic of the data-parallel training path (host/testbed_train.cpp, testbed.cpp).

What is checked, with the CPU oracle standing in for the kernels:
  * rank r draws its own rays (generator advanced by r * 2^18 * 8) and carries 1/world of the loss scale, so the SUM of the ranks'
    fp16 gradient arenas is the gradient of the mean loss -- and after the all-reduce both ranks hold bit-identical parameters;
  * the sharded optimizer (rank r updates only shard_range(n, r, world) from the reduced gradients, shards are all-gathered)
    produces exactly the parameters the replicated optimizer produces;
  * the distributed result equals a single process replaying both ranks one after the other."""
import os
import sys
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_RAYS, N_VIEWS, RES, BATCH = 192, 4, 64, 1 << 14


def _rank_gradients(world):
    """single process: the gradients every rank would compute in step 0, and the per-rank sample counts"""
    from oracle import pipeline
    t = pipeline.CpuTrainer(n_views=N_VIEWS, res=RES, batch=BATCH)
    out = [t.compute_gradients(N_RAYS, rank=r, world=world) for r in range(world)]
    return t, out


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import importlib
    import torch
    import torch.distributed as dist
    from oracle import pipeline
    import oracle
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pyngp = importlib.import_module("instant-ngp-nerf-minimal_b200").load_pyngp()
        t = pipeline.CpuTrainer(n_views=N_VIEWS, res=RES, batch=BATCH)
        p0 = t.opt.w32.copy()
        grads, marched, n_comp, loss = t.compute_gradients(N_RAYS, rank=rank, world=world)
        # all-reduce(sum) of the fp16 arena (carried in fp32 over gloo, rounded back to fp16 like the NCCL half sum of two addends)
        g = torch.from_numpy(grads.astype(np.float32))
        dist.all_reduce(g, op=dist.ReduceOp.SUM)
        reduced = g.numpy().astype(np.float16)
        # (a) replicated optimizer on every rank
        t.apply_gradients(reduced, N_RAYS)
        replicated = t.opt.w16.copy()
        # (b) sharded optimizer: this rank steps only its shard, then the shards are all-gathered
        begin, end, size = pyngp.shard_range(p0.size, rank, world)
        n_matrix = max(0, min(10240 - begin, end - begin))
        shard = oracle.Optimizer(p0[begin:end], n_matrix=n_matrix)
        shard.step(reduced[begin:end])
        mine = np.zeros(size, np.float16)
        mine[:end - begin] = shard.w16
        parts = [torch.zeros(size, dtype=torch.float16) for _ in range(world)]
        dist.all_gather(parts, torch.from_numpy(mine))
        sharded = np.concatenate([p.numpy() for p in parts])[:p0.size]
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), grads=grads, reduced=reduced, replicated=replicated, sharded=sharded,
                 counts=np.array([marched, n_comp]), loss=np.array([loss]), shard=np.array([begin, end, size]))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_data_parallel_step_world2_gloo():
    import torch.multiprocessing as mp
    world = 2
    port = 29700 + (os.getpid() % 200)
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(world, port, d), nprocs=world, join=True)
        r = [dict(np.load(os.path.join(d, f"rank{k}.npz"))) for k in range(world)]
    # the ranks worked on different rays ...
    assert not np.array_equal(r[0]["grads"], r[1]["grads"])
    assert r[0]["counts"][0] > 0 and r[1]["counts"][0] > 0
    # ... and hold identical reduced gradients and parameters afterwards
    assert np.array_equal(r[0]["reduced"].view(np.uint16), r[1]["reduced"].view(np.uint16))
    assert np.array_equal(r[0]["replicated"].view(np.uint16), r[1]["replicated"].view(np.uint16))
    # shards tile the arena exactly
    assert r[0]["shard"][0] == 0 and r[0]["shard"][1] == r[1]["shard"][0] and r[1]["shard"][1] == r[0]["replicated"].size
    assert r[0]["shard"][2] % 8 == 0
    # the sharded optimizer reproduces the replicated one bit for bit, on both ranks
    for k in range(world):
        assert np.array_equal(r[k]["sharded"].view(np.uint16), r[k]["replicated"].view(np.uint16))
    # single-process replay of both ranks: same gradients, same sum, same parameters
    t, ref = _rank_gradients(world)
    for k in range(world):
        assert np.array_equal(ref[k][0].view(np.uint16), r[k]["grads"].view(np.uint16))
        assert ref[k][1] == r[k]["counts"][0] and ref[k][2] == r[k]["counts"][1]
    total = (ref[0][0].astype(np.float32) + ref[1][0].astype(np.float32)).astype(np.float16)
    assert np.array_equal(total.view(np.uint16), r[0]["reduced"].view(np.uint16))
    t.apply_gradients(total, N_RAYS)
    assert np.array_equal(t.opt.w16.view(np.uint16), r[0]["replicated"].view(np.uint16))
    # the loss scale is split, not the loss: every rank's gradient magnitude is about 1/world of a single-rank step's
    from oracle import pipeline
    single = pipeline.CpuTrainer(n_views=N_VIEWS, res=RES, batch=BATCH).compute_gradients(N_RAYS)[0].astype(np.float32)
    ratio = np.linalg.norm(ref[0][0].astype(np.float32)) / np.linalg.norm(single)
    assert abs(ratio - 1.0 / world) < 0.02, ratio


def test_shard_ranges_tile_the_arena():
    import importlib
    pyngp = importlib.import_module("instant-ngp-nerf-minimal_b200").load_pyngp()
    for n in (12206480, 10240 + 8, 10240 + 123457):
        for world in (1, 2, 3, 4, 8):
            prev_end, size0 = 0, None
            for rank in range(world):
                b, e, s = pyngp.shard_range(n, rank, world)
                assert b == prev_end and b % 8 == 0 and s % 8 == 0 and e <= (n + 7) // 8 * 8
                size0 = s if size0 is None else size0
                assert s == size0 and e - b <= s
                prev_end = e
            assert prev_end == (n + 7) // 8 * 8
            assert world * size0 <= n + 8 + 8 * 8           # what Trainer allocates for the in-place all-gathers
