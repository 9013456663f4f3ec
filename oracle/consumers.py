"""TEST INFRASTRUCTURE (oracle).  Not part of the product path.

CPU restatement, in plain numpy / Python loops (small inputs only), of the first consumers of
the overlay's tables: variant decoding and the site allele frequency
spectrum, as tskit computes them.

* genotypes: ``tsk_variant_decode`` (git-submodules/tskit/c/tskit/genotypes.c): the alleles of a
  site are its ancestral state followed by the derived states in order of first appearance; every
  sample starts at allele 0 and the site's mutations are applied in table order, each setting the
  samples below its node (in the tree covering the site) to its allele.
* site AFS, one sample set, polarised, not span-normalised (``tsk_treeseq_allele_frequency_spectrum``
  with TSK_STAT_SITE | TSK_STAT_POLARISED, git-submodules/tskit/c/tskit/trees.c): for every site and
  every allele but the ancestral one that is carried by some but not all samples
  (trees.c:3527-3540), ``afs[number of samples carrying it] += 1``.

Pinned by tests/test_consumers_oracle.py against vectors made with the reference's own tskit
(tests/golden/make_consumers_golden.py).  Input: the ``asdict()``-style dict of numpy columns."""
import numpy as np


def _ragged(data, offset):
    b = np.asarray(data).view(np.uint8).tobytes()
    off = np.asarray(offset).astype(np.int64)
    return [b[off[i]:off[i + 1]] for i in range(len(off) - 1)]


def parent_array(tables, position):
    """parent[u] in the tree covering ``position`` (-1 for roots)."""
    e = tables["edges"]
    parent = np.full(len(tables["nodes"]["time"]), -1, dtype=np.int64)
    cover = (np.asarray(e["left"]) <= position) & (position < np.asarray(e["right"]))
    parent[np.asarray(e["child"])[cover]] = np.asarray(e["parent"])[cover]
    return parent


def sample_nodes(tables):
    return np.flatnonzero(np.asarray(tables["nodes"]["flags"]) & 1)


def genotype_matrix(tables):
    """(num_sites, num_samples) int32 allele indexes and the per-site allele lists."""
    s, m = tables["sites"], tables["mutations"]
    samples = sample_nodes(tables)
    anc = _ragged(s["ancestral_state"], s["ancestral_state_offset"])
    der = _ragged(m["derived_state"], m["derived_state_offset"])
    site_of, node_of = np.asarray(m["site"]), np.asarray(m["node"])
    G = np.zeros((len(anc), len(samples)), dtype=np.int32)
    alleles = []
    first = np.searchsorted(site_of, np.arange(len(anc) + 1))
    for j, x in enumerate(np.asarray(s["position"])):
        parent = parent_array(tables, x)
        names = [anc[j]]
        for k in range(first[j], first[j + 1]):
            if der[k] not in names:
                names.append(der[k])
            a = names.index(der[k])
            target = node_of[k]
            for col, u in enumerate(samples):  # is `target` on the path from the sample to its root?
                while u != -1 and u != target:
                    u = parent[u]
                if u == target:
                    G[j, col] = a
        alleles.append(names)
    return G, alleles


def site_afs(tables):
    """Polarised, un-normalised site frequency spectrum of all samples: length num_samples + 1."""
    G, alleles = genotype_matrix(tables)
    n = G.shape[1]
    afs = np.zeros(n + 1, dtype=np.float64)
    for j, names in enumerate(alleles):
        counts = np.bincount(G[j], minlength=len(names))
        for a in range(1, len(names)):
            if 0 < counts[a] < n:  # absent and fixed alleles are not counted (trees.c:3530)
                afs[counts[a]] += 1
    return afs
