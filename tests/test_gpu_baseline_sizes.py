"""GPU: parity and validity at the sizes BASELINE.json names (the other parity tests stay at or below
50 samples x 1e5 bp, where the CPU oracle answers in milliseconds).

* configs 1 and 2 at FULL size, and config 3's trees and model under a 16-interval map, through the replay harness: the CUDA path and the reference's algorithm
  (the compiled reference where it was built, and the C restatement) consume one pre-drawn uniform stream
  and must produce bit-identical tables, every column (config 2 is DEFINED by this check in BASELINE.json),
  each followed by a second layer with keep=True over the first;
* configs 3, 4 and 5 at FULL size on the native stream (what bench.py times): the structural invariants of
  a valid overlay (msprime_b200/validate.py: lib/mutgen.c:1027-1074 and tskit's table rules), the number of
  new mutations against its Poisson mean within 5 standard deviations, the infinite-alleles numbering
  (tests/test_mutations.py:1704-1711), and that every kept row is still there (config 4).

The code paths that only exist at scale run here: thousands of K1 tiles and the spine scan, look-back
windows over thousands of tiles in P1 / K5 / the radix passes, the chain-free K5 table, the K2 launched
ahead of the K1 read-back, 32-bit offsets near 100 MB."""
import numpy as np
import pytest

import bench
import cases
import test_gpu_replay as replay
from msprime_b200 import _native, synth, validate

pytestmark = pytest.mark.gpu
FMAX = cases.FMAX


def workload_input(name, source="treegen"):
    """``source="reference"``: the tree sequence comes from the reference's own ancestry simulator
    (oracle/ref_ancestry.py: lib/msprime.c compiled unmodified), i.e. what msprime.sim_ancestry returns for
    the configuration; otherwise from this project's SMC' generator."""
    w = bench.WORKLOADS[name]
    if source == "reference":
        from oracle import ref_ancestry

        cached = ref_ancestry.load_cached(w)  # scripts/make_reference_input.py (configs 3-5 take minutes to hours)
        if cached is not None:
            return w, cached
        if w["samples"] > 10_000:
            pytest.skip("no cached tree sequence from the reference's simulator for this size")
        if not ref_ancestry.have():
            pytest.skip("oracle/_ref/libancestry_ref.so not built")
        return w, ref_ancestry.sim_ancestry(w["samples"], sequence_length=w["L"], recombination_rate=w["recomb"],
                                            population_size=w["ne"], random_seed=w["ts_seed"])
    return w, synth.native_treeseq(w["samples"], w["L"], w["recomb"], w["ne"], w["ts_seed"])


@pytest.mark.parametrize("name", ["config1", "config2", "config3-16", "config1-reference", "config2-reference",
                                  "config3-16-reference"])
def test_replay_bit_exact_at_full_size(native_built, oracle_mod, name):
    """configN-reference: the same configuration on the tree sequence the REFERENCE's Hudson simulator makes for it
    (sim_ancestry(samples, sequence_length=, recombination_rate=1e-8, population_size=1e4, random_seed=1));
    config3-16 is config 3's tree sequence and model with a 16-interval map instead of 1000: the replay
    stream holds one uniform or more per (edge, rate interval) pair (lib/mutgen.c:1115-1120), which is 4e8
    pairs for the 1000-interval map (the ~4e5 edges alive at any position each meet every interval) against
    1.5e7 here; the sequential stream walker is what bounds the size of a replay case."""
    w, ts = workload_input(name.split("-")[0], "reference" if name.endswith("-reference") else "treegen")
    if name.startswith("config3-16"):
        w = dict(w, ratemap=16)
    L = ts["sequence_length"]
    E = len(ts["edges"]["left"])
    model = bench.oracle_model(w["model"])
    pos, rate = bench.rate_map(w, L)
    which = ["port"] + (["ref"] if oracle_mod.have("ref") else [])
    # one uniform or more per (edge, rate interval) pair, two per mutation, allele draws
    n_stream = 20_000_000 if name.startswith("config3-16") else 3 * E + 3_000_000
    base = ts
    with _native.Context(0) as ctx:
        for layer, keep in enumerate((False, True)):
            stream = np.random.default_rng(100 + layer).random(n_stream)
            okw = dict(map_position=pos, map_rate=rate, model=model, discrete_genome=w["discrete"], keep=keep)
            want = oracle_mod.sim_mutations(base, which=which[0], stream=stream, **okw)
            assert not want["stream_exhausted"]
            for other in which[1:]:
                again = oracle_mod.sim_mutations(base, which=other, stream=stream, **okw)
                assert not replay.describe_diff(want, again), "restatement and compiled reference disagree"
            got = replay.gpu_run(ctx, base, model=model, stream=stream, map_position=pos, map_rate=rate,
                                 discrete=w["discrete"], keep=keep)
            diffs = replay.describe_diff(want, got)
            assert not diffs, diffs
            assert got["draws"] == want["draws"]
            M = len(want["mutations"]["site"])
            assert M > (2_000 if name.startswith("config1") else 30_000)
            base = cases.with_kept(ts, want)


def native_overlay(ctx, ts, w, seed, keep=False):
    L = ts["sequence_length"]
    pos, rate = bench.rate_map(w, L)
    lo, hi = bench.time_window(w)
    ctx.load_tables(ts)
    counts = ctx.generate(_native.Params(
        map_position=pos, map_rate=rate, model=bench.native_model(w["model"]), discrete_genome=w["discrete"],
        keep=keep, start_time=lo, end_time=hi, seed=seed, offsets32=True))
    return counts, ctx.fetch(), (pos, rate, lo, hi)


def window_args(lo, hi):
    return dict(start_time=None if lo == -FMAX else lo, end_time=None if hi == FMAX else hi)


SOURCES = ["treegen", "reference"]  # this project's generator / the reference's own sim_ancestry (where cached)


@pytest.mark.parametrize("source", SOURCES)
def test_config3_native_stream_is_valid(native_built, source):
    """GTR, 1000-interval lognormal rate map, 1e5 samples x 1e8 bp."""
    w, ts = workload_input("config3", source)
    with _native.Context(0) as ctx:
        counts, res, (pos, rate, lo, hi) = native_overlay(ctx, ts, w, seed=31)
    assert counts["num_new_mutations"] == counts["num_mutations"] == len(res["mutations"]["site"])
    validate.check_result(ts, res, discrete=True)
    validate.check_count(ts, counts["num_new_mutations"], pos, rate, discrete=True)
    # per rate-map interval: sites fall where the rate is positive, in proportion to mass (coarse: 20 bins)
    sp = res["sites"]["position"][res["mutations"]["site"]]
    edges_ = np.linspace(0, ts["sequence_length"], 21)
    got = np.histogram(sp, bins=edges_)[0]
    for b in range(20):
        sub_pos = np.concatenate([[edges_[b]], [p for p in pos if edges_[b] < p < edges_[b + 1]], [edges_[b + 1]]])
        # mean in the bin = total mean restricted to [a, b): edges clipped to the bin
        e = ts["edges"]
        clipped = {"sequence_length": ts["sequence_length"], "nodes": ts["nodes"],
                   "edges": {"left": np.clip(e["left"], edges_[b], edges_[b + 1]),
                             "right": np.clip(e["right"], edges_[b], edges_[b + 1]),
                             "parent": e["parent"], "child": e["child"]}}
        mean = validate.expected_new_mutations(clipped, pos, rate, discrete=True)
        assert abs(got[b] - mean) < 6 * np.sqrt(mean + 1), (b, got[b], mean, len(sub_pos))
    # alleles are nucleotides, one byte each
    assert counts["derived_state_bytes"] == counts["num_mutations"]
    assert set(np.unique(res["mutations"]["derived_state"]).tolist()) <= {65, 67, 71, 84}


@pytest.mark.parametrize("source", SOURCES)
def test_config4_native_stream_keeps_and_adds(native_built, oracle_mod, source):
    """HKY, keep=True over an existing HKY layer, time window [1e3, 2e4), 1e5 samples x 1e8 bp."""
    w, ts0 = workload_input("config4", source)
    kl = w["keep_layer"]
    L = ts0["sequence_length"]
    which = "ref" if oracle_mod.have("ref") else "port"
    first = oracle_mod.sim_mutations(ts0, which=which, map_position=[0, L], map_rate=[kl["rate"]],
                                     model=bench.oracle_model(kl["model"]), discrete_genome=True, keep=False,
                                     seed=kl["seed"])
    ts = cases.with_kept(ts0, first)
    M0, S0 = len(first["mutations"]["site"]), len(first["sites"]["position"])
    with _native.Context(0) as ctx:
        counts, res, (pos, rate, lo, hi) = native_overlay(ctx, ts, w, seed=41, keep=True)
    M = counts["num_mutations"]
    assert M == M0 + counts["num_new_mutations"] and counts["num_new_mutations"] > 10_000
    # every kept row is still there, with its state: match rows by (position, node, time)
    sp = res["sites"]["position"]
    key = np.rec.fromarrays([sp[res["mutations"]["site"]], res["mutations"]["node"], res["mutations"]["time"]])
    old = np.rec.fromarrays([first["sites"]["position"][first["mutations"]["site"]], first["mutations"]["node"],
                             first["mutations"]["time"]])
    is_old = np.isin(key, old)
    assert is_old.sum() == M0
    assert np.array_equal(res["mutations"]["derived_state"][is_old], first["mutations"]["derived_state"])
    assert np.isin(first["sites"]["position"], sp).all() and len(sp) <= S0 + counts["num_new_mutations"]
    validate.check_result(ts, res, discrete=True, kept_mutations=M0, new_rows=~is_old, **window_args(lo, hi))
    validate.check_count(ts, counts["num_new_mutations"], pos, rate, discrete=True, **window_args(lo, hi))
    # parents point at earlier rows of the same site whose time is not younger (kept and new alike here,
    # because the kept layer came out of an overlay of the same trees)
    par = res["mutations"]["parent"].astype(np.int64)
    has = par >= 0
    assert np.all(res["mutations"]["time"][par[has]] >= res["mutations"]["time"][has])


@pytest.mark.parametrize("source", SOURCES)
def test_config5_native_stream_is_valid(native_built, source):
    """The workload bench.py times: InfiniteAlleles, rate 1e-7, 1e6 samples x 2.5e8 bp, 1.45e7 mutations."""
    w, ts = workload_input("config5", source)
    with _native.Context(0) as ctx:
        ctx.load_tables(ts)
        L = ts["sequence_length"]
        results = []
        for seed in (51, 52):  # the second call runs on warm buffers with K2 launched ahead of the K1 read-back
            counts = ctx.generate(_native.Params(
                map_position=[0, L], map_rate=[w["rate"]], model=bench.native_model(w["model"]), discrete_genome=True,
                keep=False, start_time=-FMAX, end_time=FMAX, seed=seed, offsets32=True))
            results.append((counts, ctx.fetch()))
    for counts, res in results:
        assert counts["num_new_mutations"] == len(res["mutations"]["site"]) > 10_000_000
        validate.check_count(ts, counts["num_new_mutations"], [0, L], [w["rate"]], discrete=True)
        validate.check_infinite_alleles_numbering(res, 0)
        assert counts["next_allele"] == counts["num_mutations"] + counts["num_sites"]
    validate.check_result(ts, results[1][1], discrete=True)
    assert not np.array_equal(results[0][1]["sites"]["position"][:1000], results[1][1]["sites"]["position"][:1000])
