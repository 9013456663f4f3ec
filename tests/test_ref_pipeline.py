"""CPU: the reference's own pipeline as the source of inputs and golden outputs (route 2).

``tests/golden/ref_pipeline.npz`` holds tree sequences made by the reference's UNMODIFIED ancestry simulator
(lib/msprime.c compiled by ``make -C oracle ancestry``) and the tables the reference's unmodified lib/mutgen.c
lays over them with a seeded MT19937 -- config 1 of BASELINE.json among them.  Here:

* the C restatement, seeded the same way, reproduces those tables bit for bit;
* so does the restatement under a REPLAY stream made of the MT19937 doubles (what the device test feeds the
  CUDA path: tests/test_gpu_ref_pipeline.py);
* where the compiled simulator is present it regenerates the committed tree sequences bit for bit;
* the GSL routines added for the simulator (oracle/gsl_shim/gsl_shim_ancestry.c) against their published
  definitions, numpy's MT19937 words and scipy;
* the simulator's output against coalescent theory (tree length and TMRCA at one locus, number of trees).
"""
import ctypes
import math

import numpy as np
import pytest

import ref_pipeline
import test_gpu_replay as replay
from oracle import oracle, ref_ancestry

CASES = ref_pipeline.load()
needs_simulator = pytest.mark.skipif(not ref_ancestry.have(), reason="oracle/_ref/libancestry_ref.so not built")


@pytest.fixture(scope="module")
def oracle_built():
    oracle.build(("port",))
    return oracle


def overlay_args(c):
    L = c["ts"]["sequence_length"]
    return dict(map_position=[0.0, L], map_rate=[c["rate"]], model=c["model"], discrete_genome=c["discrete"],
                keep=False)


@pytest.mark.parametrize("name", sorted(CASES))
def test_restatement_reproduces_the_seeded_reference(oracle_built, name):
    c = CASES[name]
    got = oracle.sim_mutations(c["ts"], which="port", seed=c["seed"], **overlay_args(c))
    assert not replay.describe_diff(c["want"], got)
    assert got["draws"] == c["draws"]
    if oracle.have("ref"):
        again = oracle.sim_mutations(c["ts"], which="ref", seed=c["seed"], **overlay_args(c))
        assert not replay.describe_diff(c["want"], again)


@pytest.mark.parametrize("name", sorted(CASES))
def test_replay_of_the_mt19937_doubles_is_the_seeded_run(oracle_built, name):
    c = CASES[name]
    stream = ref_pipeline.mt19937_doubles(c["seed"], c["draws"] + 16)
    got = oracle.sim_mutations(c["ts"], which="port", stream=stream, **overlay_args(c))
    assert not got["stream_exhausted"]
    assert not replay.describe_diff(c["want"], got)
    assert got["draws"] == c["draws"]


def test_config1_is_the_configuration_baseline_names():
    c = CASES["config1"]
    ts = c["ts"]
    assert ts["sequence_length"] == 1e6 and int(ts["nodes"]["flags"].sum()) == 2000
    assert np.all(ts["nodes"]["flags"][:2000] == 1) and np.all(ts["nodes"]["time"][:2000] == 0)
    # 1e-8 per bp per generation over the total branch length
    e, t = ts["edges"], ts["nodes"]["time"]
    mass = float(np.sum((e["right"] - e["left"]) * (t[e["parent"]] - t[e["child"]]))) * c["rate"]
    M = len(c["want"]["mutations"]["site"])
    assert abs(M - mass) < 5 * math.sqrt(mass)


@needs_simulator
@pytest.mark.parametrize("name", sorted(CASES))
def test_simulator_regenerates_the_committed_trees(name):
    import importlib.util
    import os

    spec = importlib.util.spec_from_file_location(
        "make_ref_ancestry", os.path.join(ref_pipeline.HERE, "golden", "make_ref_ancestry.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    anc = {n: a for n, a, _ in gen.CASES}[name]
    ts = ref_ancestry.sim_ancestry(**anc)
    for tbl, cols in (("nodes", ("time", "flags")), ("edges", ("left", "right", "parent", "child"))):
        for col in cols:
            a, b = ts[tbl][col], CASES[name]["ts"][tbl][col]
            assert a.dtype == b.dtype and np.array_equal(a.view(np.uint8), b.view(np.uint8)), (tbl, col)


# ---------------------------------------------------------------- the added GSL routines

@pytest.fixture(scope="module")
def shim():
    if not ref_ancestry.have():
        pytest.skip("oracle/_ref/libancestry_ref.so not built")
    lib = ctypes.CDLL(ref_ancestry.ANCESTRY_SO)
    vp, dbl, uint, ulong = ctypes.c_void_p, ctypes.c_double, ctypes.c_uint, ctypes.c_ulong
    for fn_name, res, args in (
            ("gsl_rng_alloc", vp, [vp]), ("gsl_rng_free", None, [vp]), ("gsl_rng_set", None, [vp, ulong]),
            ("gsl_rng_uniform_int", ulong, [vp, ulong]), ("gsl_ran_exponential", dbl, [vp, dbl]),
            ("gsl_ran_geometric", uint, [vp, dbl]),
            ("gsl_ran_multinomial", None, [vp, ctypes.c_size_t, uint, vp, vp]),
            ("gsl_sf_choose", dbl, [uint, uint]), ("gsl_sf_lnchoose", dbl, [uint, uint]),
            ("gsl_sf_lnbeta", dbl, [dbl, dbl]), ("gsl_isnan", ctypes.c_int, [dbl])):
        fn = getattr(lib, fn_name)
        fn.restype, fn.argtypes = res, args
    lib.mt19937 = ctypes.c_void_p.in_dll(lib, "gsl_rng_mt19937")
    return lib


@pytest.fixture
def rng(shim):
    r = shim.gsl_rng_alloc(shim.mt19937)
    yield r
    shim.gsl_rng_free(r)


def numpy_words(seed, n):
    bg = np.random.MT19937()
    bg._legacy_seeding(seed)
    return bg.random_raw(n)


@pytest.mark.parametrize("n", [1, 2, 3, 10, 1000, 2**20 + 7, 2**31, 2**32 - 1])
def test_uniform_int_is_gsl_scaling_of_the_raw_words(shim, rng, n):
    """rng/rng.c: scale = (max - min) / n; k = word / scale, redrawn while k >= n."""
    shim.gsl_rng_set(rng, 99)
    words = [int(w) for w in numpy_words(99, 4000)]
    scale = 0xFFFFFFFF // n
    it = iter(words)
    for _ in range(1000):
        k = next(it) // scale
        while k >= n:
            k = next(it) // scale
        assert shim.gsl_rng_uniform_int(rng, n) == k


def test_exponential_is_minus_mu_log1p_minus_u(shim, rng):
    shim.gsl_rng_set(rng, 5)
    u = numpy_words(5, 500).astype(np.float64) / 4294967296.0
    for x in u:
        assert shim.gsl_ran_exponential(rng, 2.5) == -2.5 * math.log1p(-x)


def test_geometric_counts_trials_to_first_success(shim, rng):
    from scipy import stats

    shim.gsl_rng_set(rng, 6)
    p = 0.3
    draws = np.array([shim.gsl_ran_geometric(rng, p) for _ in range(100_000)])
    assert draws.min() == 1
    top = 15
    obs = np.bincount(np.minimum(draws, top), minlength=top + 1)[1:]
    pmf = stats.geom.pmf(np.arange(1, top + 1), p)
    pmf[-1] = stats.geom.sf(top - 1, p)
    assert stats.chisquare(obs, pmf * len(draws)).pvalue > 1e-4
    assert shim.gsl_ran_geometric(rng, 1.0) == 1


def test_multinomial_adds_up_and_has_the_right_marginals(shim, rng):
    from scipy import stats

    shim.gsl_rng_set(rng, 8)
    p = np.array([0.5, 0.0, 1.5, 3.0])  # unnormalised, as GSL allows
    out = np.zeros(4, dtype=np.uint32)
    tot = np.zeros(4)
    reps, N = 20_000, 40
    for _ in range(reps):
        shim.gsl_ran_multinomial(rng, 4, N, p.ctypes.data, out.ctypes.data)
        assert out.sum() == N and out[1] == 0
        tot += out
    q = p / p.sum()
    z = (tot - reps * N * q)[q > 0] / np.sqrt(reps * N * q * (1 - q))[q > 0]
    assert np.all(np.abs(z) < 5)
    assert stats.norm.sf(np.abs(z)).min() > 1e-7


def test_special_functions(shim):
    from scipy import special

    for n, m in ((2, 2), (10, 3), (52, 5), (2000, 2), (100000, 2), (60, 30)):
        assert shim.gsl_sf_choose(n, m) == pytest.approx(special.comb(n, m, exact=True), rel=1e-13)
        assert shim.gsl_sf_lnchoose(n, m) == pytest.approx(math.log(special.comb(n, m, exact=True)), rel=1e-12,
                                                           abs=1e-12)
    for a, b in ((0.5, 1.5), (2.0, 3.0), (1.2, 0.8)):
        assert shim.gsl_sf_lnbeta(a, b) == pytest.approx(special.betaln(a, b), rel=1e-13)
    assert shim.gsl_isnan(float("nan")) == 1 and shim.gsl_isnan(1.0) == 0 and shim.gsl_isnan(float("inf")) == 0


# ---------------------------------------------------------------- the simulator against theory

@needs_simulator
def test_one_locus_tree_length_and_tmrca_match_the_coalescent():
    """No recombination, n = 10 haploid genomes in a diploid population of size Ne: in generations,
    E[TMRCA] = 4 Ne (1 - 1/n), E[total length] = 4 Ne H(n-1) (lib/msprime.c scales time by 2 Ne ploidy / ... )."""
    n, ne, reps = 10, 100.0, 600
    tmrca, length = [], []
    for seed in range(1, reps + 1):
        ts = ref_ancestry.sim_ancestry(n // 2, sequence_length=1.0, recombination_rate=0.0, population_size=ne,
                                       random_seed=seed)
        t, e = ts["nodes"]["time"], ts["edges"]
        assert len(t) == 2 * n - 1 and len(e["left"]) == 2 * n - 2
        tmrca.append(t.max())
        length.append(float(np.sum(t[e["parent"]] - t[e["child"]])))
    h = np.sum(1.0 / np.arange(1, n))
    for got, want in ((tmrca, 4 * ne * (1 - 1 / n)), (length, 4 * ne * h)):
        got = np.asarray(got)
        assert abs(got.mean() - want) < 5 * got.std(ddof=1) / math.sqrt(reps)


@needs_simulator
def test_number_of_trees_grows_with_the_recombination_rate():
    """E[breakpoints] = 4 Ne r (L - 1) H(n-1) for a discrete genome under Hudson's model; most of them change the tree."""
    n, ne, L, r, reps = 8, 100.0, 1000.0, 1e-4, 200
    trees = []
    for seed in range(1, reps + 1):
        ts = ref_ancestry.sim_ancestry(n // 2, sequence_length=L, recombination_rate=r, population_size=ne,
                                       random_seed=seed)
        e = ts["edges"]
        assert np.all(e["left"] == np.floor(e["left"])) and np.all(e["left"] < e["right"])
        trees.append(ts["num_trees"])
    expected_breaks = 4 * ne * r * (L - 1) * np.sum(1.0 / np.arange(1, n))
    mean = float(np.mean(trees))
    assert 0.6 * expected_breaks < mean - 1 <= 1.05 * expected_breaks


def test_prepared_input_paths_agree():
    """bench.py looks for the simulator's prepared output where scripts/make_reference_input.py (through
    oracle/ref_ancestry.py) leaves it, without importing anything of oracle/."""
    import os

    import bench

    for w in bench.WORKLOADS.values():
        key = "n%d_L%g_r%g_ne%g_seed%d" % (w["samples"], w["L"], w["recomb"], w["ne"], w["ts_seed"])
        want = os.path.join(bench.ROOT, "oracle", "_ref", "inputs", key + ".npz")
        assert os.path.normpath(ref_ancestry.cached_input_path(w)) == os.path.normpath(want)
    src = open(os.path.join(bench.ROOT, "bench.py")).read()
    body = src[src.index("def load_prepared_input"):src.index("def make_input")]
    assert "oracle import" not in body and "import oracle" not in body


@needs_simulator
def test_prepared_input_round_trip(tmp_path, monkeypatch):
    """What the preparation script stores is what the loaders hand back (a small configuration, in a scratch copy
    of the layout)."""
    import bench

    w = dict(samples=6, L=500.0, recomb=1e-4, ne=100.0, ts_seed=4)
    ts = ref_ancestry.sim_ancestry(w["samples"], sequence_length=w["L"], recombination_rate=w["recomb"],
                                   population_size=w["ne"], random_seed=w["ts_seed"])
    monkeypatch.setattr(ref_ancestry, "HERE", str(tmp_path / "oracle"))
    monkeypatch.setattr(bench, "ROOT", str(tmp_path))
    path = ref_ancestry.cached_input_path(w)
    import os

    os.makedirs(os.path.dirname(path))
    np.savez(path, sequence_length=np.float64(ts["sequence_length"]), seconds=np.float64(0.1),
             node_time=ts["nodes"]["time"], node_flags=ts["nodes"]["flags"], edge_left=ts["edges"]["left"],
             edge_right=ts["edges"]["right"], edge_parent=ts["edges"]["parent"], edge_child=ts["edges"]["child"])
    for got in (ref_ancestry.load_cached(w), bench.load_prepared_input(w)):
        assert got is not None and got["num_trees"] == ts["num_trees"]
        for tbl, cols in (("nodes", ("time", "flags")), ("edges", ("left", "right", "parent", "child"))):
            for col in cols:
                assert np.array_equal(got[tbl][col], ts[tbl][col])
    assert bench.load_prepared_input(dict(w, ts_seed=5)) is None
