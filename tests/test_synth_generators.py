"""CPU: the synthetic inputs of the bench and the tests (msprime_b200/synth.py: the numpy SMC' generator
and the C one, csrc/treegen.c) are VALID tree sequences to the reference's real tskit and look like
coalescent output (input route 3: "every output is validated with tskit").
`sim_ancestry` itself cannot run here (it needs GSL)."""
import numpy as np
import pytest

import tskit_path

tskit = tskit_path.load()
if tskit is None:
    pytest.skip("the reference's tskit package is not available here", allow_module_level=True)

from msprime_b200 import synth  # noqa: E402


def as_tskit(d):
    t = tskit.TableCollection(d["sequence_length"])
    t.nodes.set_columns(flags=d["nodes"]["flags"].astype(np.uint32), time=d["nodes"]["time"])
    e = d["edges"]
    t.edges.set_columns(left=e["left"], right=e["right"], parent=e["parent"], child=e["child"])
    return t.tree_sequence()  # runs tskit's integrity checks, edge ordering included


CASES = [("numpy", 8, 2e4, 1), ("numpy", 40, 2e5, 2), ("native", 8, 4e4, 5), ("native", 200, 2e5, 3),
         ("native", 1000, 1e6, 1)]


def generate(kind, samples, L, seed, recomb=1e-8):
    if kind == "numpy":
        return synth.sim_ancestry_smc(samples, L, recomb, 1e4, seed)
    return synth.native_treeseq(samples, L, recomb, 1e4, seed)


@pytest.mark.parametrize("kind, samples, L, seed", CASES)
def test_generated_tables_are_a_valid_tree_sequence(kind, samples, L, seed):
    d = generate(kind, samples, L, seed)
    ts = as_tskit(d)
    n = int(np.sum(d["nodes"]["flags"] & 1))
    assert ts.num_samples == n and n in (samples, 2 * samples)
    if "num_trees" in d:
        assert ts.num_trees == d["num_trees"]
    assert ts.num_trees > 1
    # tskit's canonical order is the order the tables are already in
    t = ts.dump_tables()
    t.sort()
    assert t.edges == ts.tables.edges
    # every tree has coalesced, and break points are integers (discrete genome)
    for tree in ts.trees():
        assert tree.num_roots == 1
    assert np.all(np.floor(ts.breakpoints(as_array=True)) == ts.breakpoints(as_array=True))
    # adjacent trees differ (no empty transitions) and edges are maximal (no two abutting copies)
    e = ts.tables.edges
    order = np.lexsort((e.left, e.child, e.parent))
    same = (e.parent[order][1:] == e.parent[order][:-1]) & (e.child[order][1:] == e.child[order][:-1])
    assert not np.any(same & (e.left[order][1:] == e.right[order][:-1]))


def test_no_recombination_gives_one_tree():
    ts = as_tskit(synth.native_treeseq(10, 1e4, 0.0, 1e4, 9))
    assert ts.num_trees == 1 and ts.num_edges == 2 * ts.num_samples - 2


@pytest.mark.parametrize("kind", ["numpy", "native"])
def test_coalescent_scale(kind):
    """Mean total branch length 4 Ne H_(n-1) and about rho H_(n-1) trees (SMC' hides a few events),
    averaged over seeds; loose bounds: this guards against a generator that is off by a factor."""
    n_dip, L, ne, r = 10, 2e5, 1e4, 1e-8
    n = 2 * n_dip if kind == "native" else None
    lengths, trees = [], []
    for seed in range(1, 9):
        d = generate(kind, n_dip, L, seed)
        ts = as_tskit(d)
        n = ts.num_samples
        lengths.append(sum(t.total_branch_length * t.span for t in ts.trees()) / L)
        trees.append(ts.num_trees)
    h = sum(1 / i for i in range(1, n))
    assert 0.7 < np.mean(lengths) / (4 * ne * h) < 1.3
    rho = 4 * ne * r * L
    assert 0.5 < np.mean(trees) / (rho * h) < 1.2
