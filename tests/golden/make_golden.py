"""Generates tests/golden/mutgen_ref.npz by running the UNMODIFIED reference mutgen
(lib/mutgen.c compiled from /root/reference by oracle/Makefile into oracle/_ref)
on every case of tests/golden_cases.py under a replay stream.

    python tests/golden/make_golden.py

Only runs where /root/reference exists; the fixture it writes is committed.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases  # noqa: E402
import golden_cases  # noqa: E402
from oracle import oracle  # noqa: E402

COLS = {"sites": ("position", "ancestral_state", "ancestral_state_offset", "metadata", "metadata_offset"),
        "mutations": ("site", "node", "parent", "time", "derived_state", "derived_state_offset", "metadata",
                      "metadata_offset")}


def main():
    oracle.build(("ref",))
    assert oracle.have("ref"), "the compiled reference is needed to make golden vectors"
    store = {}
    results = {}
    for c in golden_cases.all_cases():
        ts = golden_cases.build_ts(c["ts"])
        if c["base"] is not None:
            ts = cases.with_kept(ts, results[c["base"]])
        stream = np.random.default_rng(c["seed"]).random(c["n_stream"])
        r = oracle.sim_mutations(ts, which="ref", map_position=c["map_position"], map_rate=c["map_rate"],
                                 model=cases.MODELS[c["model"]], discrete_genome=c["discrete"], keep=c["keep"],
                                 start_time=c["start_time"], end_time=c["end_time"], stream=stream)
        assert not r["stream_exhausted"], c["name"]
        results[c["name"]] = r
        for tbl, cols in COLS.items():
            for col in cols:
                store[f"{c['name']}/{tbl}/{col}"] = np.asarray(r[tbl][col])
        store[f"{c['name']}/draws"] = np.array(r["draws"], dtype=np.uint64)
        store[f"{c['name']}/next_allele"] = np.array(r["next_allele"], dtype=np.uint64)
        print(f"{c['name']:45s} sites={len(r['sites']['position']):6d} mutations={len(r['mutations']['site']):6d} "
              f"draws={r['draws']}")
    path = os.path.join(HERE, "mutgen_ref.npz")
    np.savez_compressed(path, **store)
    print("wrote", path, os.path.getsize(path), "bytes,", len(results), "cases")


if __name__ == "__main__":
    main()
