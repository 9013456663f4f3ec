"""Generates tests/golden/ref_pipeline.npz: the reference's OWN pipeline, both halves, run here.

    python tests/golden/make_ref_ancestry.py

For every case: ``msprime.sim_ancestry`` as the reference's lib/msprime.c computes it (compiled unmodified by
``make -C oracle ancestry``; oracle/ancestry_driver.c makes the calls msprime/ancestry.py makes), then
``msprime.sim_mutations`` as the reference's lib/mutgen.c computes it with a SEEDED MT19937 generator
(``make -C oracle ref``), i.e. what ``sim_mutations(sim_ancestry(..., random_seed=a), rate=, random_seed=b)``
returns up to the GSL boundary (oracle/gsl_shim restates GSL's algorithms; see tests/test_gsl_shim.py).

config1 is BASELINE.json configs[0]: 1000 diploid samples x 1e6 bp, recombination 1e-8, Ne 1e4, seed 1,
JC69 at rate 1e-8, seed 2.  Only runs where /root/reference exists; the fixture it writes is committed.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases  # noqa: E402
from oracle import oracle, ref_ancestry  # noqa: E402

CASES = [
    # name, ancestry arguments, (model, rate, discrete, seed)
    ("config1", dict(samples=1000, sequence_length=1e6, recombination_rate=1e-8, population_size=1e4,
                     random_seed=1), ("jc69", 1e-8, True, 2)),
    ("small-continuous", dict(samples=8, sequence_length=100.0, recombination_rate=1e-3, population_size=100,
                              random_seed=7, discrete_genome=False), ("binary", 1e-3, False, 11)),
    ("haploid-hot", dict(samples=30, sequence_length=2000.0, recombination_rate=1e-4, population_size=500,
                         random_seed=3, ploidy=1), ("infinite", 2e-5, True, 5)),
]
MODELS = {"jc69": cases.JC69, "binary": cases.BINARY, "infinite": {"kind": "infinite_alleles", "start_allele": 0}}
SITE_COLS = ("position", "ancestral_state", "ancestral_state_offset", "metadata", "metadata_offset")
MUT_COLS = ("site", "node", "parent", "time", "derived_state", "derived_state_offset", "metadata",
            "metadata_offset")


def main():
    oracle.build(("ref", "ancestry"))
    assert oracle.have("ref") and ref_ancestry.have(), "the compiled reference is needed"
    store = {"names": np.array([c[0] for c in CASES])}
    for name, anc, (model, rate, discrete, seed) in CASES:
        ts = ref_ancestry.sim_ancestry(**anc)
        L = ts["sequence_length"]
        r = oracle.sim_mutations(ts, which="ref", map_position=[0.0, L], map_rate=[rate], model=MODELS[model],
                                 discrete_genome=discrete, keep=False, seed=seed)
        store[f"{name}/sequence_length"] = np.float64(L)
        store[f"{name}/overlay"] = np.array([model, repr(rate), str(int(discrete)), str(seed)])
        store[f"{name}/draws"] = np.uint64(r["draws"])
        for col in ("time", "flags"):
            store[f"{name}/nodes/{col}"] = ts["nodes"][col]
        for col in ("left", "right", "parent", "child"):
            store[f"{name}/edges/{col}"] = ts["edges"][col]
        for col in SITE_COLS:
            store[f"{name}/sites/{col}"] = np.asarray(r["sites"][col])
        for col in MUT_COLS:
            store[f"{name}/mutations/{col}"] = np.asarray(r["mutations"][col])
        print(name, "nodes", len(ts["nodes"]["time"]), "edges", len(ts["edges"]["left"]), "trees", ts["num_trees"],
              "sites", len(r["sites"]["position"]), "mutations", len(r["mutations"]["site"]), "draws", r["draws"])
    out = os.path.join(HERE, "ref_pipeline.npz")
    np.savez_compressed(out, **store)
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
