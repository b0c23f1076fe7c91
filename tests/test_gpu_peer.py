"""GPU test of the data-parallel exchange kernels (csrc/optimizer.cu): ngp_peer_publish_gradients -> ngp_optimizer_step_sharded
(adam_ema_kernel<PEER, WORLD>: reduce-scatter out of all ranks' gradient arenas + Adam + EMA on the own shard + all-gather of the
new fp16 weights into all ranks' parameter arenas) -> ngp_peer_wait_weights.

The `world` ranks live as separate arenas on ONE device and are driven through the same ngp_peer_set the multi-process path fills
from CUDA IPC handles (the kernels only see addresses).  Bar: after a step every rank's fp16 parameter arena is BIT-IDENTICAL to
ngp_optimizer_step run on the fp32-summed (then fp16-rounded) gradients, the owner of each shard holds the same fp32 state, and the
flags carry the epoch.  World 2, 3 and 8 cover the three template instances and a ragged last shard.

All publishes are enqueued and completed before the first sharded step starts, so no kernel ever spins on a flag whose writer is
not yet resident (on separate GPUs the ranks run concurrently; on one GPU that would depend on the block scheduler)."""
import ctypes as C
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
PKG = "instant-ngp-nerf-minimal_b200"
N_MATRIX = 10240
MAX_PEERS = 8


def _shard(n, rank, world):
    padded = (n + 7) // 8 * 8
    size = ((padded + world - 1) // world + 7) // 8 * 8
    return min(rank * size, padded), min((rank + 1) * size, padded)


@pytest.mark.parametrize("world,n", [(2, N_MATRIX + 50000), (3, N_MATRIX + 50003), (8, N_MATRIX + 123457)])
def test_sharded_step_over_peer_memory_matches_the_replicated_step(capi, world, n):
    import gpuutil as G
    import torch
    lib = capi.load()
    pyngp = importlib.import_module(PKG).load_pyngp()
    g = np.random.Generator(np.random.PCG64(world))
    n_alloc = n + 8 + 8 * MAX_PEERS
    p0 = np.zeros(n_alloc, np.float32)
    p0[:n] = (g.random(n, dtype=np.float32) - 0.5)
    cfg = capi.OptimizerCfg(1e-2, 0.9, 0.99, 1e-15, 1e-6, 128.0 / world, 0.95)

    def state():
        return dict(w32=G.dev(p0), w16=G.dev(p0.astype(np.float16).view(np.uint16)), ema=G.dev(p0.astype(np.float16).view(np.uint16)),
                    m1=G.zeros(n_alloc, torch.float32), m2=G.zeros(n_alloc, torch.float32), ps=G.zeros(n_alloc, torch.int32))
    ref = state()
    ranks = [dict(state(), grads=G.zeros(n_alloc, torch.int16), flags=G.zeros(2 * MAX_PEERS, torch.int32), ticket=G.zeros(1, torch.int32),
                  stream=torch.cuda.Stream()) for _ in range(world)]
    peer_sets = []
    for r in range(world):
        ps = capi.PeerSet()
        ps.world, ps.rank = world, r
        for p in range(world):
            ps.grads[p], ps.params[p], ps.flags[p] = ranks[p]["grads"].data_ptr(), ranks[p]["w16"].data_ptr(), ranks[p]["flags"].data_ptr()
        peer_sets.append(ps)
        assert tuple(pyngp.shard_range(n, r, world))[:2] == _shard(n, r, world)
    assert _shard(n, world - 1, world)[1] == (n + 7) // 8 * 8 and all(_shard(n, r, world)[0] < _shard(n, r, world)[1] for r in range(world))

    for epoch in (1, 2, 3):
        # every rank's own gradients: dense on the matrices, sparse on the "grid" part, different per rank
        per_rank = []
        for r in range(world):
            gr = np.zeros(n_alloc, np.float16)
            gr[:n] = (g.normal(size=n) * 0.05).astype(np.float16)
            gr[N_MATRIX:n][g.random(n - N_MATRIX) < 0.6] = 0
            per_rank.append(gr)
            ranks[r]["grads"].copy_(torch.from_numpy(gr.view(np.int16)))
        total = np.zeros(n_alloc, np.float32)
        for gr in per_rank:                                  # fixed rank order, fp32, rounded to fp16 once (adam_ema_kernel<PEER, WORLD>)
            total += gr.astype(np.float32)
        summed = total.astype(np.float16)
        G.sync()
        capi.check(lib.ngp_optimizer_step(G.stream(), C.byref(cfg), n, N_MATRIX, epoch - 1, 1e-2, G.p(ref["w32"]), G.p(ref["w16"]), G.p(G.dev(summed.view(np.uint16))),
                                          G.p(ref["m1"]), G.p(ref["m2"]), G.p(ref["ps"]), G.p(ref["ema"])))
        for r in range(world):
            capi.check(lib.ngp_peer_publish_gradients(C.c_void_p(ranks[r]["stream"].cuda_stream), C.byref(peer_sets[r]), epoch))
        G.sync()
        for r in range(world):
            lo, hi = _shard(n, r, world)
            k = ranks[r]
            capi.check(lib.ngp_optimizer_step_sharded(C.c_void_p(k["stream"].cuda_stream), C.byref(cfg), lo, hi, N_MATRIX, epoch - 1, 1e-2, G.p(k["w32"]), G.p(k["m1"]), G.p(k["m2"]),
                                                      G.p(k["ps"]), G.p(k["ema"]), C.byref(peer_sets[r]), epoch, G.p(k["ticket"])))
        G.sync()
        for r in range(world):
            capi.check(lib.ngp_peer_wait_weights(C.c_void_p(ranks[r]["stream"].cuda_stream), C.byref(peer_sets[r]), epoch))
        G.sync()
        want16 = G.host(ref["w16"], np.uint16)[:n]
        for r in range(world):
            k = ranks[r]
            assert np.array_equal(G.host(k["w16"], np.uint16)[:n], want16), f"epoch {epoch}: rank {r}'s parameters differ from the replicated step"
            flags = G.host(k["flags"], np.uint32)
            assert np.all(flags[:world] == epoch) and np.all(flags[MAX_PEERS:MAX_PEERS + world] == epoch), (epoch, r, flags)
            lo, hi = _shard(n, r, world)
            hi = min(hi, n)
            for key in ("w32", "m1", "m2", "ps", "ema"):
                a, b = G.host(k[key])[lo:hi], G.host(ref[key])[lo:hi]
                assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), f"epoch {epoch}: rank {r} shard state {key}"
            assert int(G.host(k["ticket"], np.uint32)[0]) == 0, "the last CTA resets the ticket"
    assert (G.host(ref["ps"], np.uint32)[N_MATRIX:n] < 3).any() and (G.host(ref["ps"], np.uint32)[N_MATRIX:n] > 0).any(), "the sparse skip path was exercised"


def test_peer_entry_points_reject_bad_sets(capi):
    import gpuutil as G
    import torch
    lib = capi.load()
    ps = capi.PeerSet()
    ps.world, ps.rank = 9, 0
    assert lib.ngp_peer_publish_gradients(G.stream(), C.byref(ps), 1) != 0
    ps.world, ps.rank = 2, 2
    assert lib.ngp_peer_wait_weights(G.stream(), C.byref(ps), 1) != 0
    ps.world, ps.rank = 2, 0                               # null arenas
    assert lib.ngp_peer_publish_gradients(G.stream(), C.byref(ps), 1) != 0
    assert b"peer set" in lib.ngp_last_error()
    G.sync()
