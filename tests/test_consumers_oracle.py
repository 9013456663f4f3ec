"""CPU: the numpy restatement of the table consumers (oracle/consumers.py, SURVEY.md section 8
f4: variant decoding and the site allele frequency spectrum) against vectors computed by the
reference's own tskit on the committed fixture (tests/golden/make_consumers_golden.py).  This
pins the checker of the next row; the GPU consumers themselves are not built yet."""
import os

import numpy as np

from msprime_b200 import build, trees_io
from oracle import consumers

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden", "ref_overlay.trees")
WANT = np.load(os.path.join(HERE, "golden", "consumers_ref.npz"))


def tables():
    build.build_treesio()
    return trees_io.load_tables_dict(GOLDEN)


def test_genotypes_match_tskit():
    G, alleles = consumers.genotype_matrix(tables())
    assert G.shape == WANT["genotypes"].shape
    assert np.array_equal(G, WANT["genotypes"])
    assert np.array_equal([len(a) for a in alleles], WANT["num_alleles"])


def test_site_afs_matches_tskit():
    afs = consumers.site_afs(tables())
    assert np.array_equal(afs, WANT["afs"])
    # every derived allele that segregates is counted once; absent and fixed ones are not
    assert afs[0] == 0 and afs[-1] == 0 and afs.sum() <= (WANT["num_alleles"] - 1).sum()


def test_samples_below_a_mutation():
    d = tables()
    samples = consumers.sample_nodes(d)
    pos = np.asarray(d["sites"]["position"])[np.asarray(d["mutations"]["site"])]
    node = np.asarray(d["mutations"]["node"])
    for k in range(0, len(node), 37):
        parent = consumers.parent_array(d, pos[k])
        n = 0
        for u in samples:
            while u != -1 and u != node[k]:
                u = parent[u]
            n += u == node[k]
        assert n == WANT["samples_below"][k], k
