"""Two real GPUs, one process each (NCCL): the sharded occupancy refresh of the data-parallel path against the replicated one
(scripts/dp_refresh_check.py does the work: bit-identical first refresh, rank agreement after 289 steps, falling loss).
Skipped on a single-GPU box; the single-GPU halves of the same property are in test_gpu_sampler.py
(test_grid_sample_parts_tile_the_full_call, test_splat_then_ema_equals_update_grid_and_partial_maxima_combine)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("mode", ["peer", "nccl"])
def test_sharded_refresh_on_two_gpus(mode, tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = str(tmp_path / "dp_refresh.json")
    port = 29800 + os.getpid() % 100
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "scripts", "dp_refresh_check.py"), out, mode]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.load(open(out))
    assert res["first_refresh_bit_identical_to_single_process"] and res["world"] == 2 and res["loss"] < res["loss_first"]
