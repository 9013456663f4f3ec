import sys, time, json, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import bench
from msprime_b200 import _native, build
build.build()
w = bench.WORKLOADS["config5"]
ts = bench.make_input(w)
L = float(ts["sequence_length"])
hold, host_ts = [], {"sequence_length": L, "time_units": ts["time_units"], "nodes": {}, "edges": {}, "sites": ts["sites"], "mutations": ts["mutations"]}
t, host_ts["nodes"]["time"] = bench.pinned_like(ts["nodes"]["time"]); hold.append(t)
for k in ("left", "right", "parent", "child"):
    t, host_ts["edges"][k] = bench.pinned_like(ts["edges"][k]); hold.append(t)
ctx = _native.Context(0)
for i in range(3):
    ctx.load_tables(host_ts)
ctx.kernel_timing(True)
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(5):
    ctx.load_tables(host_ts)
t1 = time.perf_counter()
print("load_tables ms", (t1 - t0) / 5 * 1e3)
kt = ctx.kernel_times()
print({k: (v[0] // 5, round(v[1] / 5, 4)) for k, v in kt.items()}, "sum", sum(v[1] for v in kt.values()) / 5)
# pure H2D rate
a = torch.empty(272_000_000, dtype=torch.uint8).pin_memory(); b = torch.empty_like(a, device="cuda")
torch.cuda.synchronize(); t0 = time.perf_counter()
for i in range(5): b.copy_(a, non_blocking=True)
torch.cuda.synchronize(); print("H2D 272MB ms", (time.perf_counter() - t0) / 5 * 1e3)
