"""What mutations MEAN under a model, after the reference's TestFiniteSites / TestInterval
(tests/test_mutations.py:837-1229): parents found on the tree, times on the branch, every transition
possible under the matrix, one-way / flip-flop / silent / circular models, integer sites, windows.
These are invariants of any correct stream, so the same bodies run twice through the public
`sim_mutations`: on the CPU with the device call answered by the oracle stand-in
(tests/oracle_engine.py; `-m "not gpu"`), and on the B200 with the native Philox stream (`-m gpu`)."""
import numpy as np
import pytest

import cases
import msprime_b200 as msp
from msprime_b200 import mutations, tables as mtables

NUC = ["A", "C", "G", "T"]


@pytest.fixture(params=["oracle", pytest.param("device", marks=pytest.mark.gpu)])
def engine(request, monkeypatch):
    if request.param == "oracle":
        import oracle_engine

        eng = oracle_engine.OracleContext()
        monkeypatch.setattr(mutations, "_context", lambda device: eng)
    else:
        request.getfixturevalue("native_built")
    monkeypatch.setattr(mutations, "_tables", mtables)
    return request.param


def small_ts(seed=1, samples=10, L=10.0):
    """Like sim_ancestry(10, sequence_length=10, recombination_rate=1): times of order one, several trees."""
    return cases.smc_ts(samples, L, seed, 0.5, 0.5)


def overlay(d, model, rate, discrete, seed=42, **kw):
    ts = mtables.TableCollection.fromdict(d).tree_sequence()
    return msp.sim_mutations(ts, rate=rate, random_seed=seed, model=model, discrete_genome=discrete, **kw)


def rows(out):
    t = out.tables
    anc, der = out.alleles()
    return (np.asarray(t.sites.position), anc, np.asarray(t.mutations.site).astype(np.int64),
            np.asarray(t.mutations.node).astype(np.int64), np.asarray(t.mutations.parent).astype(np.int64),
            np.asarray(t.mutations.time), der)


def parent_array(d, x):
    e = d["edges"]
    parent = np.full(len(d["nodes"]["time"]), -1, dtype=np.int64)
    cover = (e["left"] <= x) & (x < e["right"])
    parent[e["child"][cover]] = e["parent"][cover]
    return parent


def verify_mutations(d, out, discrete, model=None):
    """TestFiniteSites.verify_mutations (tests/test_mutations.py:844-873) on plain columns."""
    pos, anc, site, node, par, time, der = rows(out)
    node_time = d["nodes"]["time"]
    first = np.searchsorted(site, np.arange(len(pos) + 1))
    for j, x in enumerate(pos):
        tree = parent_array(d, x)
        if not discrete:
            assert first[j + 1] - first[j] == 1
        seen = {}
        for k in range(first[j], first[j + 1]):
            n = node[k]
            assert node_time[n] <= time[k] < node_time[tree[n]]
            while n != -1 and n not in seen:
                n = tree[n]
            if n == -1:
                assert par[k] == -1
                inherited = anc[j]
            else:
                assert par[k] == seen[n] and time[k] < time[seen[n]]
                inherited = der[seen[n]]
            if model is not None:
                a, b = model.alleles.index(inherited), model.alleles.index(der[k])
                assert model.transition_matrix[a][b] > 0
            seen[node[k]] = k  # last: parents are listed before their children


def matrix_model(alleles, matrix, root=None):
    root = [1 / len(alleles)] * len(alleles) if root is None else root
    return msp.MatrixMutationModel(alleles, root, matrix)


def max_per_site(out):
    _, _, site, *_ = rows(out)
    return np.bincount(site).max()


@pytest.mark.parametrize("discrete", [True, False])
def test_default_binary_and_nucleotide_alphabets(engine, discrete):
    d = small_ts()
    out = overlay(d, matrix_model(["0", "1"], [[0, 1], [1, 0]], [1, 0]), 1.0, discrete)
    _, anc, _, _, _, _, der = rows(out)
    assert out.num_sites > 0 and set(anc) == {"0"} and set(der) <= {"0", "1"}
    verify_mutations(d, out, discrete)
    model = matrix_model(NUC, [[0.25] * 4] * 4)
    out = overlay(d, model, 1.0, discrete)
    _, anc, _, _, _, _, der = rows(out)
    assert set(anc) <= set(NUC) and set(der) <= set(NUC) and len(set(anc)) > 1
    verify_mutations(d, out, discrete, model)


@pytest.mark.parametrize("discrete", [True, False])
def test_one_way_mutation(engine, discrete):
    d = small_ts()
    model = matrix_model(["0", "1"], [[0.0, 1.0], [0.0, 1.0]], [1.0, 0.0])
    out = overlay(d, model, 1.0, discrete)
    _, anc, site, _, _, _, der = rows(out)
    assert out.num_sites > 0 and set(anc) == {"0"} and set(der) == {"1"}
    assert np.bincount(site, minlength=out.num_sites).min() >= 1
    if discrete:
        assert max_per_site(out) > 1
    verify_mutations(d, out, discrete, model)


@pytest.mark.parametrize("discrete", [True, False])
def test_flip_flop_mutation(engine, discrete):
    d = small_ts()
    model = matrix_model(NUC, [[0, 0, 0.5, 0.5], [0, 0, 0.5, 0.5], [0.5, 0.5, 0, 0], [0.5, 0.5, 0, 0]])
    out = overlay(d, model, 5.0, discrete)
    _, anc, site, _, par, _, der = rows(out)
    assert out.num_sites > 0
    if discrete:
        assert max_per_site(out) > 1
    for k in range(len(der)):
        inherited = anc[site[k]] if par[k] == -1 else der[par[k]]
        assert der[k] in ("GT" if inherited in "AC" else "AC")
    verify_mutations(d, out, discrete, model)


@pytest.mark.parametrize("discrete", [True, False])
def test_do_nothing_mutations(engine, discrete):
    d = small_ts()
    model = matrix_model(NUC, np.eye(4).tolist())
    out = overlay(d, model, 5.0, discrete)
    _, anc, site, _, _, _, der = rows(out)
    assert out.num_sites > 0 and len(set(anc)) > 1
    assert all(der[k] == anc[site[k]] for k in range(len(der)))  # silent, but recorded
    verify_mutations(d, out, discrete, model)


@pytest.mark.parametrize("discrete", [True, False])
def test_uniform_mutations_away_from_the_root_state(engine, discrete):
    d = small_ts()
    model = matrix_model(NUC, [[0, 1 / 3, 1 / 3, 1 / 3], [0, 0, 0.5, 0.5], [0, 0.5, 0, 0.5], [0, 0.5, 0.5, 0]],
                         [1.0, 0, 0, 0])
    out = overlay(d, model, 20.0, discrete)
    _, anc, _, _, _, _, der = rows(out)
    assert set(anc) == {"A"}
    counts = {a: der.count(a) for a in NUC}
    assert counts["A"] == 0 and min(counts["C"], counts["G"], counts["T"]) > 0
    if discrete:
        assert max_per_site(out) > 1
    verify_mutations(d, out, discrete, model)


@pytest.mark.parametrize("discrete", [True, False])
def test_circular_mutations(engine, discrete):
    d = small_ts()
    model = matrix_model(NUC, [[0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1], [1, 0, 0, 0]], [1.0, 0, 0, 0])
    out = overlay(d, model, 2.0, discrete)
    _, anc, site, _, par, _, der = rows(out)
    assert out.num_sites > 0 and set(anc) == {"A"}
    if discrete:
        assert max_per_site(out) > 1
    for k in range(len(der)):
        inherited = "A" if par[k] == -1 else der[par[k]]
        assert NUC.index(der[k]) == (NUC.index(inherited) + 1) % 4
    verify_mutations(d, out, discrete, model)


def test_integer_sites(engine):
    d = small_ts(seed=5)
    out = overlay(d, matrix_model(["0", "1"], [[0, 1], [1, 0]], [1, 0]), 5.0, True)
    pos = rows(out)[0]
    assert pos[0] == 0.0 and np.all(pos == np.floor(pos)) and pos[-1] < d["sequence_length"]
    assert len(pos) == 10  # at this rate every one of the ten positions is hit


@pytest.mark.parametrize("keep", [True, False])
def test_zero_rate_on_mutated_input(engine, keep):
    d = small_ts()
    first = overlay(d, "binary", 2.0, False, seed=1)
    again = msp.sim_mutations(first, rate=0, keep=keep, model="binary", random_seed=2, discrete_genome=False)
    if keep:
        for a, b in zip(rows(first), rows(again)):
            assert np.array_equal(np.asarray(a), np.asarray(b))
    else:
        assert again.num_sites == 0 and again.num_mutations == 0
    assert again.num_provenances == first.num_provenances + 1


def test_stick_tree_windows_pick_the_branch(engine):
    # TestInterval.test_stick_tree, tests/test_mutations.py:1134-1200
    d = cases.stick_tree(5, 1.0)
    for k in range(4):
        out = overlay(d, "binary", 50.0, False, seed=3 + k, start_time=k, end_time=k + 1)
        _, _, _, node, _, time, _ = rows(out)
        assert len(node) > 0 and set(node.tolist()) == {k}
        assert (time >= k).all() and (time < k + 1).all()
    out = overlay(d, "binary", 50.0, False, seed=9, start_time=1.5, end_time=2.5)
    _, _, _, node, _, time, _ = rows(out)
    assert set(node.tolist()) == {1, 2} and (time >= 1.5).all() and (time < 2.5).all()
    assert (time[node == 1] < 2).all() and (time[node == 2] >= 2).all()
    for empty in ((4.0, 10.0), (-5.0, 0.0)):  # above the root and below the leaves: no branch at all
        assert overlay(d, "binary", 50.0, False, start_time=empty[0], end_time=empty[1]).num_mutations == 0


def test_coalescent_trees_windows(engine):
    # TestInterval.test_coalescent_trees / test_negative_time, :1202-1229
    d = small_ts(seed=2, samples=8, L=20.0)
    tmax = float(d["nodes"]["time"].max())
    for lo, hi in ((0.0, tmax / 4), (tmax / 4, tmax / 2), (tmax / 8, tmax)):
        out = overlay(d, "jc69", 8.0 / tmax, True, seed=4, start_time=lo, end_time=hi)
        time = rows(out)[5]
        assert len(time) > 0 and (time >= lo).all() and (time < hi).all()
        verify_mutations(d, out, True)
    assert overlay(d, "jc69", 1.0, True, start_time=-10.0, end_time=-1.0).num_mutations == 0
