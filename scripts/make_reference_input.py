"""Runs the reference's own ancestry simulator (oracle/_ref/libancestry_ref.so: lib/msprime.c compiled unmodified)
for one bench workload and stores the node and edge tables under oracle/_ref/inputs/ (git-ignored, travels to the
GPU box with the snapshot), where `bench.py --input reference` and the tests find them.

    python scripts/make_reference_input.py config3        # minutes
    python scripts/make_reference_input.py config5        # about an hour or more of one core

INPUT PREPARATION ONLY (test / bench infrastructure): nothing in the product path reads these files.
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from oracle import ref_ancestry  # noqa: E402


def main():
    name = sys.argv[1]
    w = bench.WORKLOADS[name]
    t0 = time.time()
    ts = ref_ancestry.sim_ancestry(w["samples"], sequence_length=w["L"], recombination_rate=w["recomb"],
                                   population_size=w["ne"], random_seed=w["ts_seed"])
    seconds = time.time() - t0
    path = ref_ancestry.cached_input_path(w)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    tmp = path + ".tmp.npz"
    np.savez(tmp, sequence_length=np.float64(ts["sequence_length"]), seconds=np.float64(seconds),
             node_time=ts["nodes"]["time"], node_flags=ts["nodes"]["flags"], edge_left=ts["edges"]["left"],
             edge_right=ts["edges"]["right"], edge_parent=ts["edges"]["parent"], edge_child=ts["edges"]["child"])
    os.replace(tmp, path)
    print(f"{name}: N={len(ts['nodes']['time'])} E={len(ts['edges']['left'])} trees={ts['num_trees']} in "
          f"{seconds:.0f} s -> {path} ({os.path.getsize(path) / 1e6:.0f} MB)")


if __name__ == "__main__":
    main()
