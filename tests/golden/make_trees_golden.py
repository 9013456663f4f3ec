"""Writes tests/golden/ref_overlay.trees with the REFERENCE's tsk_table_collection_dump
(through oracle/_ref, only where /root/reference exists) plus the in-memory columns it was
written from (ref_overlay_columns.npz).  Run from the repo root:

    python tests/golden/make_trees_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

import cases  # noqa: E402
from oracle import oracle  # noqa: E402


def main():
    oracle.build(("port", "ref"))
    ts = dict(cases.smc_ts(20, 5e4, 3))
    L = ts["sequence_length"]
    res = oracle.sim_mutations(ts, which="ref", map_position=[0, L], map_rate=[1e-7], model=cases.JC69, seed=11)
    ts = cases.with_kept(ts, res)
    path = os.path.join(HERE, "ref_overlay.trees")
    oracle.trees_dump(ts, path, num_samples=40, time_units="generations")
    cols = {"sequence_length": np.array(L)}
    for tbl in ("nodes", "edges"):
        for col, v in ts[tbl].items():
            cols[f"{tbl}/{col}"] = np.asarray(v)
    cols["nodes/flags"] = np.array([1] * 40 + [0] * (len(ts["nodes"]["time"]) - 40), dtype=np.uint32)
    back = oracle.trees_load(path)
    for tbl in ("sites", "mutations"):
        for col, v in back[tbl].items():
            cols[f"{tbl}/{col}"] = np.asarray(v)
    np.savez_compressed(os.path.join(HERE, "ref_overlay_columns.npz"), **cols)
    print(path, os.path.getsize(path), "bytes;", len(back["mutations"]["site"]), "mutations")


if __name__ == "__main__":
    main()
