#!/usr/bin/env python
"""Generates tests/golden/ref_cuda_step.npz: outputs of the REFERENCE'S OWN CUDA build (oracle/_ref/ref_harness, the unmodified
reference compiled for sm_100 by oracle/build_ref.sh) for one occupancy refresh + one training step on a tiny procedural
dataset.  Run on a GPU box:

    gpurun -- 'python tests/make_golden_from_ref.py gpurun_out/golden'      then copy the .npz into tests/golden/

The CPU (-m "not gpu") tests check the oracle against these vectors, which pins the oracle to the reference itself."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refcuda  # noqa: E402

PKG = "instant-ngp-nerf-minimal_b200"
N_VIEWS, RES, N_RAYS, BATCH = 4, 64, 256, 1 << 16
N_HEAD = 4096


def main(out_dir):
    pkg = importlib.import_module(PKG)
    scene = importlib.import_module(PKG + ".scene")
    capi = pkg.capi
    _, n_entries = capi.make_grid_desc()
    n_params = 10240 + 2 * n_entries
    d = refcuda.make(capi, scene, params=refcuda.start_params(n_params), batch=BATCH, n_rays=N_RAYS, n_views=N_VIEWS, res=RES)
    m, ds = d.meta, d.dataset
    k, nm = m["n_kept"], m["n_marched"]
    assert nm <= 16 * BATCH and m["n_compacted"] <= BATCH, "the fixture must not hit the sample budget (atomic-order dependent)"
    p_init = d.arr("params_init_f32", np.float32)
    n_tot = m["n_total_elements"]
    g = dict(
        meta=np.array([N_VIEWS, RES, N_RAYS, BATCH, k, nm, m["n_compacted"], n_tot], np.int64),
        fx=np.float32(ds["fx"]), xforms=ds["xforms"], images=ds["images"].view(np.uint16),
        rng=np.array([m["rng_state"], m["rng_inc"], m["sampler_rng_state"], m["sampler_rng_inc"]], np.uint64),
        # Trainer::initialize_params (seed 43): MLP part, head and tail of the grid part, and a checksum of everything
        params_init_head=p_init[:10240 + N_HEAD], params_init_tail=p_init[-N_HEAD:],
        params_init_sum=np.array([p_init.astype(np.float64).sum(), np.abs(p_init.astype(np.float64)).sum()]),
        # occupancy: untrained marks, refresh samples (head), densities fed to the grid update (head), grid after, mean, bits
        grid0_is_unseen=np.packbits(d.arr("grid0_f32", np.float32) < 0),
        grid_positions_head=d.arr("grid_positions_f32", np.float32, (n_tot, 3))[:N_HEAD],
        grid_indices_head=d.arr("grid_indices_u32", np.uint32)[:N_HEAD],
        grid_indices_xor=np.bitwise_xor.reduce(d.arr("grid_indices_u32", np.uint32)),
        grid_density_head=d.arr("grid_density_f16", np.uint16).reshape(n_tot, 16)[:N_HEAD],
        grid_mean=d.arr("grid_mean_f32", np.float32),
        bitfield0=d.arr("bitfield0_u8", np.uint8), bitfield1=d.arr("bitfield1_u8", np.uint8),
        # training march (slots are in atomic arrival order: compare per ray)
        rays=d.arr("rays_f32", np.float32, (k, 6)), rays_rgba=d.arr("rays_rgba_f32", np.float32, (k, 4)),
        numsteps=d.arr("numsteps_u32", np.uint32, (k, 2)), coords=d.arr("coords_f32", np.float32, (nm, 7)),
        # network + compositing
        rgbsigma=d.arr("rgbsigma_f16", np.uint16).reshape(-1, 4)[:nm],
        numsteps_compacted=d.arr("numsteps_compacted_u32", np.uint32, (k, 2)),
        coords_compacted=d.arr("coords_compacted_f32", np.float32, (BATCH, 7))[:m["n_compacted"]],
        dloss=d.arr("dloss_f16", np.uint16).reshape(BATCH, 4)[:m["n_compacted"]],
        loss=d.arr("loss_f32", np.float32)[:k],
    )
    os.makedirs(out_dir, exist_ok=True)
    path = os.path.join(out_dir, "ref_cuda_step.npz")
    np.savez_compressed(path, **g)
    print("wrote", path, os.path.getsize(path), "bytes; kept", k, "marched", nm, "compacted", m["n_compacted"])


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
