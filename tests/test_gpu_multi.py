"""Two real GPUs, two processes, NCCL: the sharded upload of the replicated tables
(``sharding.load_tables_sharded``) and the device-side exchange of a genome-sharded call
(``sharding.run_slice_device``).  Skipped on a single-GPU box (the same C path is covered
there by ``test_device_exchange_equals_whole`` with a device-copy all-gather)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, ROOT)
import cases
from msprime_b200 import _native, sharding

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
torch.cuda.set_device(rank)
ts = cases.smc_ts(200, 2e5, 3)
L = ts["sequence_length"]
cuts = sharding.slice_cuts(L, world)
mdl = dict(cases.INF_ALLELES, start_allele=500)
ctx = _native.Context(rank)
# the replicated tables: 1 / world of them from this rank's host, the rest over NCCL
sharding.load_tables_sharded(ctx, ts, rank=rank, world=world)
gather = sharding.nccl_device_allgather(ctx, world)
common = dict(map_position=[0, L], map_rate=[2e-6], model=cases.native_model(mdl), discrete_genome=True, keep=False,
              start_time=-cases.FMAX, end_time=cases.FMAX, seed=77)
for _ in range(3):  # repeated calls reuse the exchange buffers
    local, info = sharding.run_slice_device(
        ctx, _native.Params(slice_left=cuts[rank], slice_right=cuts[rank + 1], **common), rank=rank, world=world,
        gather=gather)
merged = sharding.gather_results(local, rank=rank, world=world)
if rank == 0:
    ctx.generate(_native.Params(**common))
    whole = ctx.fetch()
    for tbl in whole:
        for col in whole[tbl]:
            assert np.array_equal(np.asarray(whole[tbl][col]), np.asarray(merged[tbl][col])), (tbl, col)
    assert info["num_mutations"] == ctx.counts["num_mutations"] and info["next_allele"] == ctx.counts["next_allele"]
    print("MULTI_OK", info["num_mutations"])
dist.barrier()
dist.destroy_process_group()
'''


def test_two_gpu_device_exchange(native_built, tmp_path):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "worker.py"
    script.write_text(f"ROOT = {ROOT!r}\n" + WORKER)
    r = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
         "127.0.0.1", "--master-port", "29541", str(script)],
        capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "MULTI_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
