"""GPU parity, replay harness: the CUDA path and the CPU oracle consume the same
pre-drawn uniform stream and must produce bit-identical site and mutation tables
(BASELINE.json north_star, first correctness gate).  Everything goes through the C ABI."""
import numpy as np
import pytest

import cases
from msprime_b200 import _native

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(native_built):
    c = _native.Context(0)
    yield c
    c.close()


def gpu_run(ctx, ts, *, model, stream, map_position, map_rate, discrete=True, keep=False,
            start_time=-cases.FMAX, end_time=cases.FMAX):
    ctx.load_tables(ts)
    params = _native.Params(
        map_position=map_position, map_rate=map_rate, model=cases.native_model(model),
        discrete_genome=discrete, keep=keep, start_time=start_time, end_time=end_time,
        replay_stream=stream)
    counts = ctx.generate(params)
    res = ctx.fetch()
    res["draws"] = counts["draws"]
    res["next_allele"] = counts["next_allele"]
    return res


def describe_diff(a, b):
    out = []
    for tbl in ("sites", "mutations"):
        for col in a[tbl]:
            x, y = np.asarray(a[tbl][col]), np.asarray(b[tbl][col])
            if x.shape != y.shape:
                out.append(f"{tbl}.{col}: shape {x.shape} vs {y.shape}")
            elif x.size and not np.array_equal(x.view(np.uint8), y.astype(x.dtype).view(np.uint8)):
                bad = np.flatnonzero(x != y.astype(x.dtype))
                i = int(bad[0]) if bad.size else -1
                out.append(f"{tbl}.{col}: {bad.size} differ, first at {i}: {x[i]!r} vs {y[i]!r}")
    return out


def check_case(ctx, oracle_mod, ts, *, model, map_position, map_rate, seed=1, n_stream=None, **kw):
    rng = np.random.default_rng(seed)
    n_stream = n_stream or 200_000
    stream = rng.random(n_stream)
    okw = dict(map_position=map_position, map_rate=map_rate, model=model,
               discrete_genome=kw.get("discrete", True), keep=kw.get("keep", False),
               start_time=kw.get("start_time", -cases.FMAX), end_time=kw.get("end_time", cases.FMAX))
    want = oracle_mod.sim_mutations(ts, which="port", stream=stream, **okw)
    assert not want["stream_exhausted"], "test stream too short"
    got = gpu_run(ctx, ts, model=model, stream=stream, map_position=map_position, map_rate=map_rate, **kw)
    diffs = describe_diff(want, got)
    assert not diffs, diffs
    if model["kind"] == "matrix":
        assert got["draws"] == want["draws"]
    else:
        assert got["next_allele"] == want["next_allele"]
    return want


@pytest.mark.parametrize("model", list(cases.MODELS))
@pytest.mark.parametrize("discrete", [True, False])
@pytest.mark.parametrize("rate", [0.0, 0.05, 0.5, 5.0])
def test_four_leaf_tree(ctx, oracle_mod, model, discrete, rate):
    ts = cases.four_leaf_tree()
    check_case(ctx, oracle_mod, ts, model=cases.MODELS[model], map_position=[0, 10], map_rate=[rate],
               discrete=discrete)


@pytest.mark.parametrize("model", ["jc69", "binary", "infinite_alleles", "multichar", "slim"])
@pytest.mark.parametrize("discrete", [True, False])
@pytest.mark.parametrize("rate", [1e-8, 1e-6, 2e-5])
@pytest.mark.parametrize("window", [None, (1e3, 2e4)])
def test_smc_sweep_and_keep(ctx, oracle_mod, model, discrete, rate, window):
    ts = cases.smc_ts(50, 1e5, 1)
    L = ts["sequence_length"]
    kw = dict(discrete=discrete)
    if window is not None:
        kw.update(start_time=window[0], end_time=window[1])
    first = check_case(ctx, oracle_mod, ts, model=cases.MODELS[model], map_position=[0, L], map_rate=[rate],
                       n_stream=3_000_000, **kw)
    # second layer on top of the first (keep=True): parents of kept mutations are recomputed
    ts2 = cases.with_kept(ts, first)
    check_case(ctx, oracle_mod, ts2, model=cases.MODELS[model], map_position=[0, L], map_rate=[rate], seed=2,
               n_stream=3_000_000, keep=True, **kw)


@pytest.mark.parametrize("discrete", [True, False])
def test_rate_map_intervals(ctx, oracle_mod, discrete):
    ts = cases.smc_ts(50, 1e5, 1)
    L = ts["sequence_length"]
    rng = np.random.default_rng(5)
    pos = np.concatenate([[0], np.sort(rng.uniform(1, L - 1, 40)), [L]])  # non-integer breakpoints
    rate = rng.lognormal(np.log(1e-6), 1.0, 41)
    rate[::7] = 0.0
    check_case(ctx, oracle_mod, ts, model=cases.JC69, map_position=pos, map_rate=rate, discrete=discrete,
               n_stream=2_000_000)


def test_zero_rate_intervals_site_positions(ctx, oracle_mod):
    # tests/test_mutations.py:528-546: rate 1 on [0,3) and [7,10), 0 between
    ts = cases.four_leaf_tree()
    want = check_case(ctx, oracle_mod, ts, model=cases.BINARY, map_position=[0, 3, 7, 10], map_rate=[1, 0, 1],
                      discrete=True)
    assert sorted(set(want["sites"]["position"])) == [0, 1, 2, 7, 8, 9]


def test_stream_exhausted(ctx):
    ts = cases.four_leaf_tree()
    with pytest.raises(_native.LibraryError) as e:
        gpu_run(ctx, ts, model=cases.JC69, stream=np.random.default_rng(1).random(5), map_position=[0, 10],
                map_rate=[1.0])
    assert e.value.status == -1003


@pytest.mark.parametrize("max_span", ["1", "2"])
def test_walk_range_recovery(ctx, oracle_mod, monkeypatch, max_span):
    # K0 stores, per index entry, the parent's entries overlapping it; spans that do not fit are
    # marked unknown and the walk recovers the parent's range from child_start.  A tiny limit
    # makes that path carry most steps of a dense overlay (many multi-record sites).
    monkeypatch.setenv("MSPB_K0_MAX_SPAN", max_span)
    ts = cases.smc_ts(50, 1e5, 1)
    L = ts["sequence_length"]
    first = check_case(ctx, oracle_mod, ts, model=cases.JC69, map_position=[0, L], map_rate=[2e-5],
                       n_stream=3_000_000)
    ts2 = cases.with_kept(ts, first)
    check_case(ctx, oracle_mod, ts2, model=cases.JC69, map_position=[0, L], map_rate=[2e-5], seed=2,
               n_stream=3_000_000, keep=True)


def _class_model(m):
    return {"kind": "matrix", "alleles": list(m.alleles), "root_distribution": np.asarray(m.root_distribution).tolist(),
            "transition_matrix": np.asarray(m.transition_matrix).tolist()}


@pytest.mark.parametrize("name", ["pam", "blosum62", "smm", "tpm", "el2"])
def test_large_alphabets(ctx, oracle_mod, name):
    # 20 amino acids / up to 49 repeat numbers, multi-character alleles ("10", "11", ...):
    # the allele search of kept states and the O(A) running sums of root and transition draws
    import msprime_b200 as msp

    model = {"pam": msp.PAM, "blosum62": msp.BLOSUM62, "smm": lambda: msp.SMM(lo=2, hi=50),
             "tpm": lambda: msp.TPM(p=0.6, m=0.4, lo=8, hi=30),
             "el2": lambda: msp.EL2(m=0.5, u=0.6, v=0.01, lo=1, hi=15)}[name]()
    md = _class_model(model)
    ts = cases.smc_ts(50, 1e5, 1)
    L = ts["sequence_length"]
    first = check_case(ctx, oracle_mod, ts, model=md, map_position=[0, L], map_rate=[1e-5], n_stream=3_000_000)
    ts2 = cases.with_kept(ts, first)
    check_case(ctx, oracle_mod, ts2, model=md, map_position=[0, L], map_rate=[1e-5], seed=2, n_stream=3_000_000,
               keep=True)


def test_long_states_on_full_tiles(ctx, oracle_mod):
    """A tile of K5 holds 1024 records and stages their state bytes in 12 KB of shared memory; states of 13
    bytes or more (long alleles of a matrix model, allele numbers from 10^12 on) do not fit and must go
    straight to the table instead, whatever the model.  Bit-exact against the oracle with tens of thousands
    of records (every tile full), with and without kept rows."""
    ts = cases.smc_ts(50, 1e5, 1)
    L = ts["sequence_length"]
    long_model = {"kind": "matrix", "alleles": ["A" * 40, "C" * 33, "G" * 50, "T" * 21],
                  "root_distribution": [0.25] * 4, "transition_matrix": cases.JC69["transition_matrix"]}
    for model in (long_model, dict(cases.INF_ALLELES, start_allele=10**14), dict(cases.INF_ALLELES, start_allele=2**64 - 5000)):
        first = check_case(ctx, oracle_mod, ts, model=model, map_position=[0, L], map_rate=[4e-6], n_stream=3_000_000)
        assert len(first["mutations"]["site"]) > 50_000
        start2 = first["next_allele"] if model["kind"] == "infinite_alleles" else None
        model2 = dict(model, start_allele=start2) if start2 is not None else model
        check_case(ctx, oracle_mod, cases.with_kept(ts, first), model=model2, map_position=[0, L], map_rate=[4e-6],
                   seed=2, n_stream=3_000_000, keep=True)
