"""File in -> GPU -> file out at benchmark scale.

Writes the workload's input tree sequence as a `.trees` file, then times the stages of
`python -m msprime_b200 mutations` on it without the interpreter start-up: map + schema
check, upload + index build (straight from the mapping), overlay, fetch, `.trees` write.
Prints one JSON line.  Usage: python scripts/file_pipeline.py [--workload config5] [--dir /dev/shm]
"""
import argparse
import json
import os
import sys
import time


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from msprime_b200 import _native, build, mutations, trees_io  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="config5")
    ap.add_argument("--dir", default="/dev/shm")
    ap.add_argument("--repeat", type=int, default=3)
    args = ap.parse_args()
    build.build()
    w = bench.WORKLOADS[args.workload]
    ts = bench.make_input(w)
    src, dst = os.path.join(args.dir, "mspb_in.trees"), os.path.join(args.dir, "mspb_out.trees")
    t0 = time.perf_counter()
    trees_io.dump(ts, src)
    write_in = time.perf_counter() - t0
    L = float(ts["sequence_length"])
    map_pos, map_rate = bench.rate_map(w, L)
    best = None
    with _native.Context(0) as ctx:
        for rep in range(args.repeat):
            st = {}
            t = time.perf_counter()
            store = trees_io.Store(src)
            tables, _ = store.native_tables()
            st["map+check"] = time.perf_counter() - t
            t = time.perf_counter()
            ctx.load_native(tables)
            st["upload+index"] = time.perf_counter() - t
            t = time.perf_counter()
            counts = ctx.generate(_native.Params(
                map_position=map_pos, map_rate=map_rate, model=bench.native_model(w["model"]),
                discrete_genome=w["discrete"], keep=True, start_time=bench.time_window(w)[0],
                end_time=bench.time_window(w)[1], seed=50 + rep, offsets32=True))
            st["overlay"] = time.perf_counter() - t
            t = time.perf_counter()
            res = ctx.fetch()
            st["fetch"] = time.perf_counter() - t
            t = time.perf_counter()
            items = {k: v for k, v in store.items() if not k.startswith(("sites/", "mutations/")) and k != "uuid"}
            d = {"sequence_length": L, "sites": dict(res["sites"]), "mutations": dict(res["mutations"])}
            new = trees_io.file_items(d)
            for k, v in new.items():
                if k.startswith(("sites/", "mutations/")) or k == "uuid":
                    items[k] = v
            trees_io.write_items(dst, items)
            st["write"] = time.perf_counter() - t
            st["total"] = sum(st.values())
            store.close()
            last_count = int(counts["num_new_mutations"])
            if best is None or st["total"] < best["total"]:
                best = dict(st, mutations=int(counts["num_new_mutations"]))
    out = {
        "workload": w["text"], "dir": args.dir, "input_bytes": os.path.getsize(src),
        "output_bytes": os.path.getsize(dst), "write_input_s": round(write_in, 3),
        "stages_s": {k: round(v, 4) for k, v in best.items() if k != "mutations"},
        "new_mutations": best["mutations"],
        "mutations_per_s_file_to_file": best["mutations"] / best["total"],
    }
    # the output must load back (container + schema checks) with the right row count
    with trees_io.Store(dst) as s2:
        t2, _ = s2.native_tables()
        assert t2.num_mutations == last_count, (t2.num_mutations, last_count)
    os.unlink(src)
    os.unlink(dst)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
