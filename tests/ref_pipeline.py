"""Loader of tests/golden/ref_pipeline.npz (tests/golden/make_ref_ancestry.py): tree sequences from the
reference's own ancestry simulator with the mutation tables the reference's own lib/mutgen.c lays over
them under a seeded MT19937."""
import os

import numpy as np

import cases
from msprime_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
MODELS = {"jc69": cases.JC69, "binary": cases.BINARY, "infinite": {"kind": "infinite_alleles", "start_allele": 0}}
SITE_COLS = ("position", "ancestral_state", "ancestral_state_offset", "metadata", "metadata_offset")
MUT_COLS = ("site", "node", "parent", "time", "derived_state", "derived_state_offset", "metadata",
            "metadata_offset")


def load():
    z = np.load(os.path.join(HERE, "golden", "ref_pipeline.npz"))
    out = {}
    for name in z["names"]:
        name = str(name)
        model, rate, discrete, seed = (str(x) for x in z[f"{name}/overlay"])
        ts = {
            "sequence_length": float(z[f"{name}/sequence_length"]), "time_units": "generations",
            "nodes": {c: z[f"{name}/nodes/{c}"] for c in ("time", "flags")},
            "edges": {c: z[f"{name}/edges/{c}"] for c in ("left", "right", "parent", "child")},
            "sites": synth.empty_sites(), "mutations": synth.empty_mutations(),
        }
        want = {"sites": {c: z[f"{name}/sites/{c}"] for c in SITE_COLS},
                "mutations": {c: z[f"{name}/mutations/{c}"] for c in MUT_COLS}}
        out[name] = dict(ts=ts, want=want, model=MODELS[model], rate=float(rate), discrete=bool(int(discrete)),
                         seed=int(seed), draws=int(z[f"{name}/draws"]))
    return out


def mt19937_doubles(seed, n):
    """The doubles ``gsl_rng_uniform`` returns for ``gsl_rng_set(seed)``: numpy's independent MT19937 with the
    same 2002 initialiser, 32-bit words scaled by 2^-32 (tests/test_gsl_shim.py pins the shim to it).  Fed to the
    replay harness it stands for the seeded generator: every draw of lib/mutgen.c with per-edge means below 10
    is one such double (gsl_ran_flat, gsl_rng_uniform, the product loop of gsl_ran_poisson)."""
    bg = np.random.MT19937()
    bg._legacy_seeding(seed)
    return bg.random_raw(n).astype(np.float64) / 4294967296.0
