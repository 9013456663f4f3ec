"""Writes tests/golden/model_matrices.npz: alleles, root distributions and transition matrices
of the REFERENCE's Python model classes (msprime/mutations.py:179-1400, imported through
scripts/ref_models.py) for the parameter sets in tests/model_cases.py.  Run from the repo root
where /root/reference exists:  python tests/golden/make_model_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "scripts"), os.path.join(ROOT, "tests")]

import model_cases  # noqa: E402
import ref_models  # noqa: E402


def main():
    ref = ref_models.load()
    out = {}
    for key, cls, kwargs in model_cases.CASES:
        m = getattr(ref, cls)(**kwargs)
        out[f"{key}/alleles"] = np.array(list(m.alleles))
        out[f"{key}/root_distribution"] = np.asarray(m.root_distribution, dtype=np.float64)
        out[f"{key}/transition_matrix"] = np.asarray(m.transition_matrix, dtype=np.float64)
    path = os.path.join(HERE, "model_matrices.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes,", len(model_cases.CASES), "models")


if __name__ == "__main__":
    main()
