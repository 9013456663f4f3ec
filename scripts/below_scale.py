"""Sizes the planned level-synchronous sample-count kernels (scripts/below_levels.py) on the bench
inputs: levels, events sorted, break points.  CPU only (numpy prototype).  Usage:
    python scripts/below_scale.py SAMPLES LENGTH [SEED]   (diploid samples, as bench.py's configs)"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import below_levels  # noqa: E402
from msprime_b200 import synth  # noqa: E402


def main():
    n, L = int(float(sys.argv[1])), float(sys.argv[2])
    seed = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    t = time.time()
    d = synth.native_treeseq(n, L, 1e-8, 1e4, seed)
    gen = time.time() - t
    N, E = len(d["nodes"]["time"]), len(d["edges"]["left"])
    t = time.time()
    _, st = below_levels.build_fast(d)
    per = st.pop("per_level")
    out = dict(samples=2 * n, sequence_length=L, nodes=N, edges=E, trees=d["num_trees"], generate_s=round(gen, 1),
               build_s=round(time.time() - t, 1), **st, break_points_per_edge=round(st["break_points"] / E, 2),
               events_per_edge=round(st["events_sorted"] / E, 2), largest_level_events=max(p[2] for p in per),
               median_level_events=sorted(p[2] for p in per)[len(per) // 2])
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
