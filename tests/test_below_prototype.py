"""CPU: the numpy prototype of the level-synchronous sample-count algorithm (scripts/below_levels.py,
the plan for the device consumers) against the reference's real tskit:
`Tree.num_samples(u)` for every node of every tree, and the polarised site AFS of multi-hit overlays.
Groundwork for SURVEY.md section 8 f4; no product code is involved."""
import os
import sys

import numpy as np
import pytest

import tskit_path

tskit = tskit_path.load()
if tskit is None:
    pytest.skip("the reference's tskit package is not available here", allow_module_level=True)

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))
import below_levels  # noqa: E402
import cases  # noqa: E402
from oracle import consumers  # noqa: E402


def as_tskit(d, res=None):
    t = tskit.TableCollection(d["sequence_length"])
    t.nodes.set_columns(flags=d["nodes"]["flags"].astype(np.uint32), time=d["nodes"]["time"])
    e = d["edges"]
    t.edges.set_columns(left=e["left"], right=e["right"], parent=e["parent"], child=e["child"])
    if res is not None:
        s, m = res["sites"], res["mutations"]
        t.sites.set_columns(position=s["position"], ancestral_state=s["ancestral_state"].view(np.int8),
                            ancestral_state_offset=s["ancestral_state_offset"].astype(np.uint64))
        t.mutations.set_columns(site=m["site"], node=m["node"], parent=m["parent"], time=m["time"],
                                derived_state=m["derived_state"].view(np.int8),
                                derived_state_offset=m["derived_state_offset"].astype(np.uint64))
    return t.tree_sequence()


@pytest.mark.parametrize("samples, L, seed", [(2, 2e3, 1), (5, 2e4, 2), (30, 5e4, 3), (60, 2e5, 4)])
def test_sample_counts_match_every_tree(samples, L, seed):
    d = cases.smc_ts(samples, L, seed)
    fns, stats = below_levels.build(d)
    ts = as_tskit(d)
    assert stats["levels"] >= 2 and stats["break_points"] == len(fns.breaks)
    rng = np.random.default_rng(seed)
    samples_mask = (d["nodes"]["flags"] & 1) != 0
    picked = set(rng.choice(ts.num_trees, size=min(ts.num_trees, 25), replace=False).tolist()) | {0, ts.num_trees - 1}
    nodes = np.arange(ts.num_nodes)
    for tree in ts.trees():
        if tree.index not in picked:
            continue
        a, b = tree.interval
        in_tree = (tree.parent_array[:-1] != -1) | (tree.num_children_array[:-1] > 0) | samples_mask
        want = np.where(in_tree, tree.num_samples_array[:-1] if hasattr(tree, "num_samples_array")
                        else np.array([tree.num_samples(u) for u in nodes]), 0)
        for x in (a, (a + b) / 2, np.nextafter(b, a)):
            got = fns.value_at(nodes, np.full(ts.num_nodes, x))
            assert np.array_equal(got, want), (x, np.flatnonzero(got != want)[:5])


def test_four_leaf_tree_and_stick():
    fns, stats = below_levels.build(cases.four_leaf_tree())
    assert fns.value_at(range(7), np.full(7, 3.0)).tolist() == [1, 1, 1, 1, 2, 2, 4]
    assert stats["levels"] == 3
    fns, stats = below_levels.build(cases.stick_tree(5))
    assert fns.value_at(range(5), np.full(5, 0.5)).tolist() == [1, 1, 1, 1, 1] and stats["levels"] == 5


@pytest.mark.parametrize("model, rate, seed", [("jc69", 5e-7, 1), ("jc69", 1e-5, 2), ("binary", 1e-5, 3),
                                               ("multichar", 5e-6, 4), ("infinite_alleles", 5e-6, 5)])
def test_afs_from_sample_counts_matches_tskit(oracle_mod, model, rate, seed):
    d = dict(cases.smc_ts(12, 4e3, seed))
    res = oracle_mod.sim_mutations(d, which="port", map_position=[0, d["sequence_length"]], map_rate=[rate],
                                   model=cases.MODELS[model], seed=seed + 10)
    full = cases.with_kept(d, res)
    ts = as_tskit(d, res)
    want = ts.allele_frequency_spectrum(polarised=True, span_normalise=False)
    got = below_levels.site_afs(full)
    assert (res["mutations"]["parent"] >= 0).any() or rate < 1e-6  # multi-hit sites are in play
    assert np.array_equal(got, want)
    assert np.array_equal(consumers.site_afs(full), want)  # and the genotype-based checker agrees


@pytest.mark.parametrize("samples, L, seed", [(5, 2e4, 2), (30, 5e4, 3), (200, 1e6, 5)])
def test_level_at_a_time_version_is_identical(samples, L, seed):
    """`build_fast` (a whole level per step, per-range bisection, one sort per level: the shape of the
    device version, used by scripts/below_scale.py to size the kernels) builds the same functions."""
    d = cases.smc_ts(samples, L, seed)
    a, sa = below_levels.build(d)
    b, sb = below_levels.build_fast(d)
    assert np.array_equal(a.start, b.start) and np.array_equal(a.breaks, b.breaks)
    assert np.array_equal(a.values, b.values)
    assert (sa["levels"], sa["events_sorted"], sa["break_points"]) == (sb["levels"], sb["events_sorted"],
                                                                      sb["break_points"])
    assert sum(p[2] for p in sb["per_level"]) == sb["events_sorted"]
