"""Host-side time of each stage of a sharded overlay (device exchange path), per rank 0.

torchrun --nproc-per-node N --master-addr 127.0.0.1 scripts/prof_slice.py
"""
import os
import sys
import time

sys.path.insert(0, os.getcwd())
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from msprime_b200 import _native, sharding, synth  # noqa: E402

rank = int(os.environ["RANK"])
world = int(os.environ["WORLD_SIZE"])
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
torch.cuda.set_device(rank)
ts = synth.native_treeseq(1000000, 2.5e8, 1e-8, 1e4, 1)
L = ts["sequence_length"]
ctx = _native.Context(rank)
ctx.load_tables(ts)
cuts = sharding.slice_cuts(L, world)
gather = sharding.nccl_device_allgather(ctx, world)


def P(step):
    return _native.Params(map_position=[0, L], map_rate=[1e-7], model={"kind": 1, "start_allele": 0},
                          discrete_genome=True, keep=False, start_time=-1e308, end_time=1e308, seed=step,
                          slice_left=cuts[rank], slice_right=cuts[rank + 1], offsets32=True, slice_stream=rank + 1)


T = {}


def tick(name, t0):
    T[name] = T.get(name, 0) + time.perf_counter() - t0


for step in range(13):
    if step == 3:
        T.clear()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        wall0 = time.perf_counter()
    t = time.perf_counter(); p = P(step); tick("params", t)
    t = time.perf_counter(); c1 = ctx.generate_begin(p); tick("begin", t)
    t = time.perf_counter(); ctx.exchange_post_counts(); tick("post", t)
    t = time.perf_counter(); gather(); tick("gather1", t)
    t = time.perf_counter(); ctx.generate_end_device(rank, world); tick("end_device", t)
    t = time.perf_counter(); gather(); tick("gather2", t)
    t = time.perf_counter(); ctx.generate_finish_device(rank, world); tick("finish", t)
torch.cuda.synchronize()
wall = time.perf_counter() - wall0
if rank == 0:
    print({k: round(1e3 * v / 10, 3) for k, v in T.items()}, "ms per step; wall", round(1e3 * wall / 10, 3),
          "begin device_ms", round(c1["device_ms"], 3), [round(x, 3) for x in ctx.phase_ms().values()]
          if isinstance(ctx.phase_ms(), dict) else ctx.phase_ms())
dist.destroy_process_group()
