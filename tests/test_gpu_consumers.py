"""GPU: the consumers of the resident tables (include/msprime_b200.h): samples below
every mutation and the polarised site allele frequency spectrum, against

* oracle/consumers.py (numpy restatement of tsk_variant_decode / tsk_treeseq_allele_frequency_spectrum, pinned
  to vectors of the reference's own tskit in tests/test_consumers_oracle.py) on inputs small enough for it;
* scripts/below_levels.py (numpy prototype of the level-synchronous step functions, held to tskit's
  Tree.num_samples on every node of every tree in tests/test_below_prototype.py) at mid size;
* independent parent-array walks on a sample of mutations at BASELINE config 3 size.
Integer work: exact equality throughout."""
import os
import sys

import numpy as np
import pytest

import bench
import cases
from msprime_b200 import _native, synth
from oracle import consumers

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))
import below_levels  # noqa: E402

pytestmark = pytest.mark.gpu
FMAX = cases.FMAX


def overlay(ctx, ts, model, rate, seed, discrete=True, keep=False, offsets32=False):
    L = ts["sequence_length"]
    ctx.load_tables(ts)
    ctx.generate(_native.Params(map_position=[0, L], map_rate=[rate], model=cases.native_model(model),
                                discrete_genome=discrete, keep=keep, start_time=-FMAX, end_time=FMAX, seed=seed,
                                offsets32=offsets32))
    return ctx.fetch()


def merged(ts, res):
    d = dict(ts)
    d["sites"], d["mutations"] = res["sites"], res["mutations"]
    return d


@pytest.mark.parametrize("model", ["jc69", "binary", "infinite_alleles", "multichar"])
@pytest.mark.parametrize("discrete", [True, False])
def test_afs_and_samples_below_match_the_tskit_restatement(native_built, model, discrete):
    ts = cases.smc_ts(8, 3e3, 3, 1e-6, 1e3)  # 16 sample nodes, many trees, times ~1e3
    with _native.Context(0) as ctx:
        # rate high enough that most sites of the discrete genome are hit several times (back mutations,
        # repeated alleles, parents inside the site)
        res = overlay(ctx, ts, cases.MODELS[model], 4e-4 if discrete else 1e-4, seed=5, discrete=discrete)
        first = merged(ts, res)
        info = ctx.build_sample_index(ts["nodes"]["flags"])
        assert info["num_samples"] == 16 and info["levels"] >= 2
        for layer in range(2):
            d = merged(ts, res)
            M = len(res["mutations"]["site"])
            assert M > 200
            want_afs = consumers.site_afs(d)
            got_afs = ctx.site_afs()
            assert np.array_equal(got_afs, want_afs.astype(np.uint64)), (model, discrete, layer)
            # variant decoding, every site: allele indexes of all samples (tsk_variant_decode)
            want_G, _ = consumers.genotype_matrix(d)
            G = ctx.genotypes()
            assert G.shape == want_G.shape and np.array_equal(G, want_G), (model, discrete, layer)
            part = ctx.genotypes(5, 17)
            assert np.array_equal(part, want_G[5:17])
            # samples below = carriers of a mutation with nothing underneath it: the genotype matrix knows
            below = ctx.samples_below()
            samples = consumers.sample_nodes(d)
            pos = np.asarray(d["sites"]["position"])[np.asarray(d["mutations"]["site"])]
            for k in range(0, M, max(1, M // 40)):
                parent = consumers.parent_array(d, pos[k])
                n = 0
                for u in samples:
                    while u != -1 and u != d["mutations"]["node"][k]:
                        u = parent[u]
                    n += u == d["mutations"]["node"][k]
                assert below[k] == n, k
            # second layer: keep=True over the first (kept rows, kept sites without new mutations), 32-bit offsets
            res = overlay(ctx, first, cases.MODELS[model], 2e-4 if discrete else 1e-4, seed=6, discrete=discrete,
                          keep=True, offsets32=True)
            ctx.build_sample_index(ts["nodes"]["flags"])  # load_tables dropped the index


@pytest.mark.parametrize("source", ["smc", "hudson"])
@pytest.mark.parametrize("samples, L, seed", [(30, 5e4, 3), (60, 2e5, 4), (500, 1e6, 5)])
def test_step_functions_match_the_numpy_prototype(native_built, samples, L, seed, source):
    """source="hudson": the same on tree sequences from the reference's own ancestry simulator."""
    if source == "hudson":
        from oracle import ref_ancestry

        if not ref_ancestry.have():
            pytest.skip("oracle/_ref/libancestry_ref.so not built")
        ts = ref_ancestry.sim_ancestry(samples, sequence_length=L, recombination_rate=1e-8, population_size=1e4,
                                       random_seed=seed)
    else:
        ts = cases.smc_ts(samples, L, seed)
    fns, stats = below_levels.build_fast(ts)
    with _native.Context(0) as ctx:
        res = overlay(ctx, ts, cases.JC69, 3e-6, seed=seed)
        info = ctx.build_sample_index(ts["nodes"]["flags"])
        assert info["num_samples"] == 2 * samples
        assert info["levels"] == stats["levels"]
        below = ctx.samples_below()
        pos = np.asarray(res["sites"]["position"])[np.asarray(res["mutations"]["site"])]
        want = fns.value_at(np.asarray(res["mutations"]["node"]), pos)
        assert np.array_equal(below, want)
        d = merged(ts, res)
        afs = ctx.site_afs()
        assert np.array_equal(afs, below_levels.site_afs(d, fns).astype(np.uint64))
        # the two consumers agree: the spectrum recomputed from decoded genotypes of every site
        S, n = len(res["sites"]["position"]), 2 * samples
        from_G = np.zeros(n + 1, dtype=np.uint64)
        for a in range(0, S, 4096):
            G = ctx.genotypes(a, min(S, a + 4096))
            for allele in range(1, int(G.max()) + 1):
                c = (G == allele).sum(axis=1)
                from_G += np.bincount(c[(c > 0) & (c < n)], minlength=n + 1).astype(np.uint64)
        assert np.array_equal(from_G, afs)


@pytest.mark.parametrize("source", ["treegen", "reference"])
def test_config3_size(native_built, source):
    """1e5 samples x 1e8 bp (E = 2.3e6): build time, query time, and samples below a mutation checked by
    walking the parent array from every sample for a handful of mutations.  source="reference": on the tree
    sequence the reference's own ancestry simulator makes for config 3 (where it has been prepared), whose
    deepest nodes sit more levels up than the relaxation takes rounds (the index is built twice and must come
    out the same)."""
    w = bench.WORKLOADS["config3"]
    if source == "reference":
        ts = bench.load_prepared_input(w)
        if ts is None:
            pytest.skip("no prepared tree sequence of the reference's simulator for config 3")
    else:
        ts = synth.native_treeseq(w["samples"], w["L"], w["recomb"], w["ne"], w["ts_seed"])
    n = int(np.sum(ts["nodes"]["flags"] & 1))
    with _native.Context(0) as ctx:
        res = overlay(ctx, ts, cases.INF_ALLELES, 1e-8, seed=9, offsets32=True)
        info = ctx.build_sample_index(ts["nodes"]["flags"])
        assert info["num_samples"] == n == 200_000
        afs = ctx.site_afs()
        below = ctx.samples_below()
        again = ctx.build_sample_index(ts["nodes"]["flags"])
        assert (again["levels"], again["steps"]) == (info["levels"], info["steps"])
        assert np.array_equal(ctx.site_afs(), afs) and np.array_equal(ctx.samples_below(), below)
        print(f"config3 sample index: {info}; afs query {ctx.afs_ms:.3f} ms for {len(below)} mutations")
    M = len(below)
    assert M > 400_000 and below.min() >= 1 and below.max() <= n
    d = merged(ts, res)
    samples = consumers.sample_nodes(d)
    pos = np.asarray(res["sites"]["position"])[np.asarray(res["mutations"]["site"])]
    rng = np.random.default_rng(1)
    for k in rng.choice(M, size=6, replace=False):
        parent = consumers.parent_array(d, pos[k])
        # samples under the node = samples whose root path meets it: count by marking descendants level by level
        target = res["mutations"]["node"][k]
        u = samples.copy()
        hit = np.zeros(len(u), dtype=bool)
        while True:
            hit |= u == target
            alive = (u != -1) & ~hit
            if not alive.any():
                break
            u = np.where(alive, parent[np.maximum(u, 0)], u)
            u = np.where(hit, target, u)
        assert below[k] == hit.sum(), k
    # infinite alleles: every derived allele is unique, so a single-mutation site adds one count at below[m];
    # the spectrum of the single-mutation sites must be a lower bound bin by bin
    site = np.asarray(res["mutations"]["site"])
    single = np.ones(M, dtype=bool)
    single[1:] &= site[1:] != site[:-1]
    single[:-1] &= site[:-1] != site[1:]
    lower = np.bincount(below[single & (below < n)], minlength=n + 1)
    assert np.all(afs[: n + 1] >= lower) and afs[0] == 0 and afs[n] == 0
    assert afs.sum() <= M and afs.sum() >= lower.sum()


def test_config5_size(native_built):
    """The size bench.py reports its consumer figures on (1e6 diploid samples x 2.5e8 bp; the reference simulator's
    tree sequence where it has been prepared, else this project's generator's): the deepest nodes sit far more levels
    up than the level relaxation takes rounds, so the index is built twice and must come out the same, and samples
    below a mutation are checked against a walk up the parent array for a handful of mutations."""
    w = bench.WORKLOADS["config5"]
    ts = bench.load_prepared_input(w)
    if ts is None:
        ts = synth.native_treeseq(w["samples"], w["L"], w["recomb"], w["ne"], w["ts_seed"])
    n = int(np.sum(ts["nodes"]["flags"] & 1))
    with _native.Context(0) as ctx:
        res = overlay(ctx, ts, cases.INF_ALLELES, 2e-9, seed=9, offsets32=True)  # ~3e5 mutations
        info = ctx.build_sample_index(ts["nodes"]["flags"])
        afs, below = ctx.site_afs(), ctx.samples_below()
        again = ctx.build_sample_index(ts["nodes"]["flags"])
        assert (again["levels"], again["steps"]) == (info["levels"], info["steps"])
        assert np.array_equal(ctx.site_afs(), afs) and np.array_equal(ctx.samples_below(), below)
    print(f"config5 sample index: {info}")
    M = len(below)
    assert info["num_samples"] == n == 2_000_000 and M > 100_000 and below.min() >= 1 and below.max() <= n
    d = merged(ts, res)
    samples = consumers.sample_nodes(d)
    pos = np.asarray(res["sites"]["position"])[np.asarray(res["mutations"]["site"])]
    rng = np.random.default_rng(2)
    # (mutations high up in the tree are the ones whose counts pass through the deepest levels)
    deep = np.argsort(below)[-2000:]
    for k in list(rng.choice(M, size=2, replace=False)) + list(rng.choice(deep, size=2, replace=False)):
        parent = consumers.parent_array(d, pos[k])
        target = res["mutations"]["node"][k]
        u = samples.copy()
        hit = np.zeros(len(u), dtype=bool)
        while True:
            hit |= u == target
            alive = (u != -1) & ~hit
            if not alive.any():
                break
            u = np.where(alive, parent[np.maximum(u, 0)], u)
        assert below[k] == hit.sum(), (k, below[k], hit.sum())
