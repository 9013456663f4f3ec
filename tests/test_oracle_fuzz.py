"""CPU: the oracle's C restatement (oracle/mutgen_port.c, the checker that travels to the GPU box)
against the compiled, unmodified reference mutgen (oracle/_ref) on RANDOM inputs: trees, rate maps
with zero stretches, stochastic matrices with structural zeros and alleles of ragged length, time
windows, both genome kinds, two layers with keep=True (sometimes under a model that cannot read
the first layer's alleles), free-running MT19937 as well as replay.
Bit-exact tables, the same number of draws, or the same error code.  Skipped where the compiled
reference is absent (it is only built where /root/reference exists)."""
import numpy as np
import pytest

import cases
import golden_util

N_CASES = 150


def random_model(rng):
    kind = rng.integers(0, 5)
    if kind == 0:
        return dict(cases.INF_ALLELES, start_allele=int(rng.integers(0, 10**6)))
    if kind == 1:
        return dict(cases.SLIM, type=int(rng.integers(0, 5)), next_id=int(rng.integers(0, 1000)))
    n = int(rng.integers(2, 9))
    pool = ["A", "C", "G", "T", "", "AA", "x" * 17, "é", "0", "1", "long-allele-name", "b", "\U0001F333"]
    alleles = [pool[i] for i in rng.permutation(len(pool))[:n]]
    root = rng.random(n) * (rng.random(n) < 0.8)
    if root.sum() == 0:
        root[0] = 1
    root /= root.sum()
    matrix = rng.random((n, n)) * (rng.random((n, n)) < 0.7)
    for i in range(n):
        if matrix[i].sum() == 0:
            matrix[i, (i + 1) % n] = 1
        matrix[i] /= matrix[i].sum()
    # renormalise until numerically tight (the library checks sums with a relative tolerance)
    root[-1] = 1 - root[:-1].sum() if root[-1] > 0.1 else root[-1]
    return {"kind": "matrix", "alleles": alleles, "root_distribution": root.tolist(),
            "transition_matrix": matrix.tolist()}


def random_map(rng, L, scale):
    k = int(rng.integers(1, 8))
    inner = np.sort(rng.choice(np.arange(1, int(L)), size=k - 1, replace=False)).astype(float) if k > 1 else []
    if k > 1 and rng.random() < 0.5:
        inner = inner + rng.random(k - 1) * 0.5  # break points need not be integers
    position = np.concatenate([[0.0], inner, [L]])
    rate = rng.random(k) * scale * (rng.random(k) < 0.75)
    return position.tolist(), rate.tolist()


def run(oracle_mod, which, ts, kw):
    try:
        return oracle_mod.sim_mutations(ts, which=which, **kw)
    except oracle_mod.OracleError as e:
        return e.status


@pytest.mark.parametrize("case", range(N_CASES))
def test_port_equals_compiled_reference(oracle_mod, case):
    if not oracle_mod.have("ref"):
        pytest.skip("compiled reference absent")
    rng = np.random.default_rng(1000 + case)
    shape = rng.integers(0, 3)
    if shape == 0:
        ts, L = cases.four_leaf_tree(), 10.0
    elif shape == 1:
        ts, L = cases.stick_tree(int(rng.integers(2, 8)), 20.0), 20.0
    else:
        L = float(rng.choice([2e3, 2e4, 1e5]))
        ts = cases.smc_ts(int(rng.integers(2, 40)), L, int(rng.integers(1, 5)))
    height = float(np.max(ts["nodes"]["time"]))
    branch = float(np.sum((ts["edges"]["right"] - ts["edges"]["left"])
                          * (ts["nodes"]["time"][ts["edges"]["parent"]] - ts["nodes"]["time"][ts["edges"]["child"]])))
    target = float(rng.choice([5, 50, 400, 3000]))  # expected mutations per layer
    layers = []
    for layer in range(2):
        position, rate = random_map(rng, L, 2 * target / branch)
        kw = dict(map_position=position, map_rate=rate, model=random_model(rng) if layer == 0 else None,
                  discrete_genome=bool(rng.random() < 0.7), keep=bool(layer == 1 or rng.random() < 0.3))
        if rng.random() < 0.4:
            a, b = sorted(rng.random(2) * height * 1.2)
            kw.update(start_time=float(a), end_time=float(b))
        if rng.random() < 0.25:
            kw["seed"] = int(rng.integers(1, 2**32))
        else:
            kw["stream"] = rng.random(int(40 * target) + 10_000)
        layers.append(kw)
    if rng.random() < 0.85:  # else: a foreign model over the first layer's alleles, usually an error (same code)
        layers[1]["model"] = layers[0]["model"]
    else:
        layers[1]["model"] = random_model(rng)
    for kw in layers:
        want, got = run(oracle_mod, "ref", ts, kw), run(oracle_mod, "port", ts, kw)
        if isinstance(want, int) or isinstance(got, int):
            assert want == got, (want, got)
            return
        assert not golden_util.diff_tables(want, got)
        assert (got["draws"], got["next_allele"]) == (want["draws"], want["next_allele"])
        ts = cases.with_kept(ts, want)


@pytest.mark.parametrize("case", range(40))
def test_port_equals_reference_over_hand_made_existing_tables(oracle_mod, case):
    """keep=True over existing rows that no overlay wrote: a mutation on every branch of every tree
    (rows in tree order, which is NOT time order, or oldest first), unique metadata, and new mutations
    landing on the same integer positions.  New rows are merged into a site of existing rows only
    where those are oldest first: otherwise the reference's comparator (existing rows by id, the rest
    by time, lib/mutgen.c:36-49) is not a total order and its own assertion `mut->time <=
    parent_mut->time` (lib/mutgen.c:1200) fires; rows in tree order get a continuous genome."""
    if not oracle_mod.have("ref"):
        pytest.skip("compiled reference absent")
    import test_surface_semantics as sem

    rng = np.random.default_rng(5000 + case)
    d = sem.small_ts(seed=int(rng.integers(1, 6)), samples=int(rng.integers(2, 9)), L=float(rng.choice([5, 10, 30])))
    oldest_first = bool(case % 2)
    existing = sem.branch_mutations(d, oldest_first=oldest_first)
    model = cases.BINARY if rng.random() < 0.7 else cases.JC69  # JC69 cannot read "0" / "1": same error on both sides
    kw = dict(map_position=[0, d["sequence_length"]], map_rate=[float(rng.choice([0.0, 0.2, 1.0, 4.0]))], model=model,
              discrete_genome=oldest_first and bool(rng.random() < 0.8), keep=True, stream=rng.random(200_000))
    want, got = run(oracle_mod, "ref", existing, kw), run(oracle_mod, "port", existing, kw)
    if isinstance(want, int) or isinstance(got, int):
        assert want == got, (want, got)
        return
    assert not golden_util.diff_tables(want, got)
    assert got["draws"] == want["draws"]


@pytest.mark.parametrize("case", range(24))
def test_port_equals_reference_on_trees_with_several_roots(oracle_mod, case):
    """Tree sequences cut at a time (what `sim_ancestry(end_time=...)` leaves): unary stretches, isolated
    samples, several roots per tree; all three model kinds, both genome kinds, a second layer with keep."""
    if not oracle_mod.have("ref"):
        pytest.skip("compiled reference absent")
    import test_surface_semantics as sem

    rng = np.random.default_rng(9000 + case)
    full = sem.small_ts(seed=int(rng.integers(1, 8)), samples=int(rng.integers(2, 12)), L=float(rng.choice([5, 20])))
    ts = sem.cut_above(full, float(np.quantile(full["nodes"]["time"], rng.choice([0.55, 0.7, 0.9]))))
    if len(ts["edges"]["left"]) == 0:  # nothing left below the cut (two samples): the uncut input then
        ts = full
    model = [cases.JC69, cases.BINARY, cases.INF_ALLELES, cases.SLIM, cases.MULTICHAR][case % 5]
    for layer in range(2):
        kw = dict(map_position=[0, ts["sequence_length"]], map_rate=[float(rng.choice([0.3, 2.0]))], model=model,
                  discrete_genome=bool(case % 3), keep=bool(layer), stream=rng.random(100_000))
        want, got = run(oracle_mod, "ref", ts, kw), run(oracle_mod, "port", ts, kw)
        if isinstance(want, int) or isinstance(got, int):
            assert want == got, (want, got)
            return
        assert not golden_util.diff_tables(want, got)
        assert (got["draws"], got["next_allele"]) == (want["draws"], want["next_allele"])
        ts = cases.with_kept(ts, want)


@pytest.mark.parametrize("case", range(60))
def test_port_equals_reference_on_the_reference_simulators_trees(oracle_mod, case):
    """The same comparison on tree sequences from the reference's OWN ancestry simulator (lib/msprime.c compiled
    unmodified, oracle/ref_ancestry.py): random sample sizes, lengths, recombination rates, ploidies and genome kinds;
    two layers, replay and free-running MT19937."""
    from oracle import ref_ancestry

    if not oracle_mod.have("ref") or not ref_ancestry.have():
        pytest.skip("compiled reference absent")
    rng = np.random.default_rng(9000 + case)
    L = float(rng.choice([50, 1e3, 2e4]))
    discrete_trees = bool(rng.random() < 0.7)
    ts = ref_ancestry.sim_ancestry(
        int(rng.integers(1, 25)), sequence_length=L, recombination_rate=float(rng.choice([0.0, 1e-4, 1e-3])) / 4,
        population_size=float(rng.choice([10, 100, 1000])), random_seed=int(rng.integers(1, 2**31)),
        ploidy=int(rng.integers(1, 3)), discrete_genome=discrete_trees)
    branch = float(np.sum((ts["edges"]["right"] - ts["edges"]["left"])
                          * (ts["nodes"]["time"][ts["edges"]["parent"]] - ts["nodes"]["time"][ts["edges"]["child"]])))
    target = float(rng.choice([5, 50, 400, 2000]))
    model = random_model(rng)
    for layer in range(2):
        position, rate = random_map(rng, L, 2 * target / (branch or 1.0))  # (one haploid sample: no edges)
        kw = dict(map_position=position, map_rate=rate, model=model, discrete_genome=bool(rng.random() < 0.7),
                  keep=bool(layer == 1))
        if rng.random() < 0.3:
            kw["seed"] = int(rng.integers(1, 2**32))
        else:
            kw["stream"] = rng.random(int(40 * target) + 10_000)
        want, got = run(oracle_mod, "ref", ts, kw), run(oracle_mod, "port", ts, kw)
        if isinstance(want, int) or isinstance(got, int):
            assert want == got, (want, got)
            return
        assert not golden_util.diff_tables(want, got)
        assert (got["draws"], got["next_allele"]) == (want["draws"], want["next_allele"])
        ts = cases.with_kept(ts, want)
