"""Writes msprime_b200/data/amino_acid_models.npz: alleles, root distribution and transition
matrix of the two empirical amino-acid models (BLOSUM62, Henikoff & Henikoff 1992; PAM, Dayhoff
et al. 1978, both as distributed with Seq-Gen) exactly as the reference's classes build them
(msprime/mutations.py:853-1387).  These are published tables, not code; they are taken from the
reference so that `model="blosum62"` / `model="pam"` give the reference's numbers bit for bit.
Run where /root/reference exists:  python scripts/make_aa_models.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_models  # noqa: E402


def main():
    ref = ref_models.load()
    out = {}
    for name, cls in (("blosum62", ref.BLOSUM62), ("pam", ref.PAM)):
        m = cls()
        out[f"{name}_alleles"] = np.array(list(m.alleles))
        out[f"{name}_root_distribution"] = np.asarray(m.root_distribution, dtype=np.float64)
        out[f"{name}_transition_matrix"] = np.asarray(m.transition_matrix, dtype=np.float64)
    path = os.path.join(os.path.dirname(HERE), "msprime_b200", "data", "amino_acid_models.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
