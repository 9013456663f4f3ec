"""Writes tests/golden/consumers_ref.npz: genotype matrix and polarised site AFS of
tests/golden/ref_overlay.trees as the REFERENCE's tskit computes them (SURVEY.md section 8 f4:
the vectors that pin oracle/consumers.py and, later, the GPU consumers).  Needs an importable
tskit (see tests/test_tskit_interop.py for the half-minute build of the vendored package):

    python tests/golden/make_consumers_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
for p in (os.environ.get("MSPB_TSKIT_PATH"), "/tmp/tskit_build/python"):
    if p and os.path.isdir(os.path.join(p, "tskit")):
        sys.path.append(p)
import tskit  # noqa: E402


def main():
    ts = tskit.load(os.path.join(HERE, "ref_overlay.trees"))
    G = ts.genotype_matrix()
    afs = ts.allele_frequency_spectrum(mode="site", polarised=True, span_normalise=False)
    n_alleles = np.array([len(v.alleles) for v in ts.variants()], dtype=np.int32)
    below = np.zeros(ts.num_mutations, dtype=np.int64)  # samples below each mutation's node at its site
    for tree in ts.trees():
        for site in tree.sites():
            for mut in site.mutations:
                below[mut.id] = tree.num_samples(mut.node)
    path = os.path.join(HERE, "consumers_ref.npz")
    np.savez_compressed(path, genotypes=G.astype(np.int32), afs=afs, num_alleles=n_alleles, samples_below=below)
    print(path, os.path.getsize(path), "bytes;", G.shape, "genotypes, afs sum", afs.sum())


if __name__ == "__main__":
    main()
