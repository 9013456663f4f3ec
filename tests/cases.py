"""Shared inputs for the parity tests: small tree sequences, models and parameter sweeps."""
import functools

import numpy as np

from msprime_b200 import synth

FMAX = np.finfo(np.float64).max


def four_leaf_tree(L=10.0):
    """The hand-built single tree of lib/tests/test_mutations.c:22-94 (4 leaves, 3 internal nodes)."""
    d = {
        "sequence_length": float(L),
        "time_units": "generations",
        "nodes": {"time": np.array([0, 0, 0, 0, 1, 2, 3.0]), "flags": np.array([1, 1, 1, 1, 0, 0, 0], dtype=np.uint32)},
        "edges": {
            "left": np.zeros(6), "right": np.full(6, float(L)),
            "parent": np.array([4, 4, 5, 5, 6, 6], dtype=np.int32),
            "child": np.array([0, 1, 2, 3, 4, 5], dtype=np.int32),
        },
        "sites": synth.empty_sites(),
        "mutations": synth.empty_mutations(),
    }
    return d


def stick_tree(n=5, L=1.0):
    """A chain 0 <- 1 <- ... (tests/test_mutations.py:1134-1165): node k at time k."""
    return {
        "sequence_length": float(L),
        "time_units": "generations",
        "nodes": {"time": np.arange(n, dtype=np.float64), "flags": np.array([1] + [0] * (n - 1), dtype=np.uint32)},
        "edges": {
            "left": np.zeros(n - 1), "right": np.full(n - 1, float(L)),
            "parent": np.arange(1, n, dtype=np.int32), "child": np.arange(0, n - 1, dtype=np.int32),
        },
        "sites": synth.empty_sites(),
        "mutations": synth.empty_mutations(),
    }


@functools.lru_cache(maxsize=None)
def smc_ts(samples, L, seed, recomb=1e-8, ne=1e4):
    return synth.sim_ancestry_smc(samples, L, recomb, ne, seed)


JC69 = {"kind": "matrix", "alleles": ["A", "C", "G", "T"], "root_distribution": [0.25] * 4,
        "transition_matrix": (np.full((4, 4), 1 / 3) - np.eye(4) / 3).tolist()}
JC69_SI = {"kind": "matrix", "alleles": ["A", "C", "G", "T"], "root_distribution": [0.25] * 4,
           "transition_matrix": np.full((4, 4), 0.25).tolist()}
BINARY = {"kind": "matrix", "alleles": ["0", "1"], "root_distribution": [1.0, 0.0],
          "transition_matrix": [[0.0, 1.0], [1.0, 0.0]]}
MULTICHAR = {"kind": "matrix", "alleles": ["", "AA", "\U0001F4A9", "x" * 17], "root_distribution": [0.1, 0.2, 0.3, 0.4],
             "transition_matrix": [[0.1, 0.2, 0.3, 0.4], [0.25, 0.25, 0.25, 0.25], [0.0, 0.5, 0.0, 0.5],
                                   [0.4, 0.3, 0.2, 0.1]]}
INF_ALLELES = {"kind": "infinite_alleles", "start_allele": 0}
# SLiMMutationModel(type=1, next_id=7, slim_generation=50), msprime/mutations.py:109-151
SLIM = {"kind": "slim", "type": 1, "next_id": 7, "slim_generation": 50}
MODELS = {"jc69": JC69, "jc69_si": JC69_SI, "binary": BINARY, "multichar": MULTICHAR, "infinite_alleles": INF_ALLELES,
          "slim": SLIM}


def with_kept(ts, result):
    d = dict(ts)
    d["sites"] = result["sites"]
    d["mutations"] = result["mutations"]
    return d


def native_model(m):
    """oracle-style model dict -> msprime_b200._native model dict"""
    if m["kind"] == "slim":
        return {"kind": 2, "type": m["type"], "next_id": m.get("next_id", 0),
                "slim_generation": m.get("slim_generation", 1)}
    if m["kind"] == "matrix":
        return {"kind": 0, "alleles": m["alleles"], "root_distribution": m["root_distribution"],
                "transition_matrix": m["transition_matrix"]}
    return {"kind": 1, "start_allele": m.get("start_allele", 0)}
