"""Wall-clock split of the end-to-end step of bench.py (pinned host buffers): load_tables
(H2D + validation + K0), generate, fetch (D2H).  One JSON line.  Usage: python scripts/e2e_breakdown.py"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from msprime_b200 import _native, build  # noqa: E402


def main():
    build.build()
    w = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "config5"]
    ts = bench.make_input(w)
    L = float(ts["sequence_length"])
    hold, host_ts = [], {"sequence_length": L, "time_units": ts["time_units"], "nodes": {}, "edges": {},
                         "sites": ts["sites"], "mutations": ts["mutations"]}
    t, host_ts["nodes"]["time"] = bench.pinned_like(ts["nodes"]["time"])
    hold.append(t)
    for k in ("left", "right", "parent", "child"):
        t, host_ts["edges"][k] = bench.pinned_like(ts["edges"][k])
        hold.append(t)
    map_pos, map_rate = bench.rate_map(w, L)
    ctx = _native.Context(0)
    out = None
    res = []
    for i in range(6):
        p = _native.Params(map_position=map_pos, map_rate=map_rate, model=bench.native_model(w["model"]),
                           discrete_genome=w["discrete"], keep=False, start_time=bench.time_window(w)[0],
                           end_time=bench.time_window(w)[1], seed=10 + i, offsets32=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ctx.load_tables(host_ts)
        t1 = time.perf_counter()
        ctx.generate(p)
        t2 = time.perf_counter()
        if out is None:
            sizes = ctx.result_sizes()
            out = {}
            for name, dtype in _native._COLUMNS:
                if name.endswith("_offset"):
                    dtype = np.uint32
                tt, out[name] = bench.pinned_empty(int(sizes[name] * 1.3) + 4096, dtype)
                hold.append(tt)
            t2 = time.perf_counter()
        ctx.fetch(out)
        t3 = time.perf_counter()
        res.append({"load_tables_ms": (t1 - t0) * 1e3, "generate_ms": (t2 - t1) * 1e3, "fetch_ms": (t3 - t2) * 1e3})
    sizes = ctx.result_sizes()
    d2h = sum(int(sizes[n]) * np.dtype(out[n].dtype).itemsize for n in sizes)
    print(json.dumps({"last": res[-1], "all": res, "d2h_bytes": d2h}))


if __name__ == "__main__":
    main()
