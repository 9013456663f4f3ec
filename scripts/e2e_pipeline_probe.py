"""Where the pipelined end-to-end loop of bench.py loses time against the host link's floor: the same loop with
parts left out (download only, upload only, everything), at depth 1..3, with per-phase wall times per thread.
Usage (under gpurun): python scripts/e2e_pipeline_probe.py > gpurun_out/e2e_probe.json"""
import json
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from msprime_b200 import _native  # noqa: E402

w = bench.WORKLOADS["config5"]
ts = bench.make_input(w)
L = float(ts["sequence_length"])
hold = []
host_ts = {"sequence_length": L, "time_units": ts["time_units"], "nodes": {}, "edges": {}, "sites": ts["sites"],
           "mutations": ts["mutations"]}
t, host_ts["nodes"]["time"] = bench.pinned_like(ts["nodes"]["time"])
hold.append(t)
for k in ("left", "right", "parent", "child"):
    t, host_ts["edges"][k] = bench.pinned_like(ts["edges"][k])
    hold.append(t)


def params(seed):
    return _native.Params(map_position=[0, L], map_rate=[w["rate"]], model=bench.native_model(w["model"]),
                          discrete_genome=True, keep=False, start_time=-bench.FMAX, end_time=bench.FMAX, seed=seed,
                          offsets32=True)


DEPTH_MAX = 3
ctxs = [_native.Context(0) for _ in range(DEPTH_MAX)]
bufs = []
for c in ctxs:
    c.load_tables(host_ts)
    c.generate(params(1))
    sizes = c.result_sizes()
    b = {}
    for name, dtype in _native._COLUMNS:
        if name.endswith("_offset"):
            dtype = np.uint32
        t, b[name] = bench.pinned_empty(int(sizes[name] * 1.3) + 4096, dtype)
        hold.append(t)
    bufs.append(b)
    c.fetch(b)
plist = [params(100 + i) for i in range(64)]


def run(depth, steps, do_load, do_fetch):
    phase = [[0.0, 0.0, 0.0] for _ in range(depth)]

    def worker(k, barrier):
        torch.cuda.set_device(0)
        barrier.wait()
        for i in range(k, steps, depth):
            t0 = time.perf_counter()
            if do_load:
                ctxs[k].load_tables(host_ts)
            t1 = time.perf_counter()
            ctxs[k].generate(plist[i % len(plist)])
            t2 = time.perf_counter()
            if do_fetch:
                ctxs[k].fetch(bufs[k])
            t3 = time.perf_counter()
            phase[k][0] += t1 - t0
            phase[k][1] += t2 - t1
            phase[k][2] += t3 - t2

    barrier = threading.Barrier(depth + 1)
    ths = [threading.Thread(target=worker, args=(k, barrier)) for k in range(depth)]
    for t_ in ths:
        t_.start()
    torch.cuda.synchronize()
    barrier.wait()
    t0 = time.perf_counter()
    for t_ in ths:
        t_.join()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    per = steps / depth
    return {"depth": depth, "load": do_load, "fetch": do_fetch, "ms_per_step": round(1e3 * wall / steps, 3),
            "thread_ms_per_own_step": [[round(1e3 * x / per, 2) for x in p] for p in phase]}


out = []
for depth in (1, 2, 3):
    for do_load, do_fetch in ((False, True), (True, False), (True, True)):
        run(depth, depth, do_load, do_fetch)
        out.append(run(depth, 12, do_load, do_fetch))
        print(json.dumps(out[-1]), flush=True)
