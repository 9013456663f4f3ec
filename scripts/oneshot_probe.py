"""Wall time of the one-shot entry point `mspb_sim_mutations` (the call INTEGRATION.md's stub binds: pageable host
tables in, library-owned host columns out) on a bench workload, next to the context calls on pinned buffers."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from msprime_b200 import _native  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="config5")
    ap.add_argument("--calls", type=int, default=6)
    args = ap.parse_args()
    w = dict(bench.WORKLOADS[args.workload])
    ts = bench.make_input(w)
    L = ts["sequence_length"]
    pos, rate = bench.rate_map(w, L)
    lo, hi = bench.time_window(w)
    lib = _native.lib()
    t, hold = _native.Context._tables_struct(ts)
    walls, muts = [], 0
    for i in range(args.calls):
        p = _native.Params(map_position=pos, map_rate=rate, model=bench.native_model(w["model"]),
                           discrete_genome=w["discrete"], keep=bool(w.get("keep_layer")), start_time=lo, end_time=hi,
                           seed=2000 + i, offsets32=True)
        c, cols = _native._Counts(), _native._Columns()
        t0 = time.perf_counter()
        _native.check(lib.mspb_sim_mutations(0, C.byref(t), C.byref(p.struct), C.byref(c), C.byref(cols)))
        t1 = time.perf_counter()
        first_site = float(np.ctypeslib.as_array(C.cast(cols.site_position, C.POINTER(C.c_double)), shape=(1,))[0])
        lib.mspb_columns_free(C.byref(cols))
        t2 = time.perf_counter()
        walls.append({"call_ms": 1e3 * (t1 - t0), "free_ms": 1e3 * (t2 - t1)})
        muts = c.num_new_mutations
        assert 0 <= first_site < L
    del hold
    steady = sorted(x["call_ms"] for x in walls[2:])
    print(json.dumps({"workload": args.workload, "new_mutations_per_call": int(muts), "calls": walls,
                      "call_ms_median_after_two": steady[len(steady) // 2]}))


if __name__ == "__main__":
    main()
