"""GPU, native Philox stream: the second correctness gate of BASELINE.json's north star.
Over 1000 seeds the overlay must be statistically indistinguishable from the model the
reference implements (and from the reference itself under MT19937), p > 0.01, on per-edge
mutation counts, site counts, allele-transition frequencies, mutation times, rate-map
bins and the site frequency spectrum -- the numeric form of verification.py:5479-5975."""
import numpy as np
import pytest
from scipy import stats

import cases
import table_checks
from msprime_b200 import _native, synth

pytestmark = pytest.mark.gpu
SEEDS = 1000
P_MIN = 0.01


@pytest.fixture(scope="module")
def ctx(native_built):
    c = _native.Context(0)
    yield c
    c.close()


def overlays(ctx, ts, *, model, seeds, map_position=None, map_rate=None, rate=None, discrete=True,
             start_time=-cases.FMAX, end_time=cases.FMAX):
    ctx.load_tables(ts)
    L = ts["sequence_length"]
    for seed in seeds:
        p = _native.Params(map_position=map_position or [0, L], map_rate=map_rate or [rate],
                           model=cases.native_model(model), discrete_genome=discrete, keep=False,
                           start_time=start_time, end_time=end_time, seed=seed)
        ctx.generate(p)
        yield ctx.fetch()


def branch_lengths(ts):
    e, t = ts["edges"], ts["nodes"]["time"]
    return t[e["parent"]] - t[e["child"]]


def chi2_p(observed, expected, min_expected=5.0):
    """Pearson chi-square with small-expectation bins pooled into one."""
    observed, expected = np.asarray(observed, float), np.asarray(expected, float)
    small = expected < min_expected
    if small.any():
        observed = np.concatenate([observed[~small], [observed[small].sum()]])
        expected = np.concatenate([expected[~small], [expected[small].sum()]])
    keep = expected > 0
    stat = ((observed[keep] - expected[keep]) ** 2 / expected[keep]).sum()
    return stats.chi2.sf(stat, keep.sum() - 1)


@pytest.fixture(scope="module")
def ts_rec():
    return synth.native_treeseq(8, 4e4, 1e-8, 1e4, 5)   # 16 sample nodes, a few trees


def test_per_edge_counts_are_poisson(ctx, ts_rec):
    ts = ts_rec
    rate = 8e-8
    e = ts["edges"]
    mu = branch_lengths(ts) * (e["right"] - e["left"]) * rate  # discrete genome, integer breakpoints
    E = len(mu)
    counts = np.zeros((SEEDS, E), dtype=np.int64)
    for k, res in enumerate(overlays(ctx, ts, model=cases.JC69, rate=rate, seeds=range(1, SEEDS + 1))):
        m = res["mutations"]
        idx = table_checks.edge_of(ts, m["node"].astype(np.int64), res["sites"]["position"][m["site"]])
        assert (idx >= 0).all()
        counts[k] = np.bincount(idx, minlength=E)
    assert mu.max() > 12 and mu.min() < 1  # both samplers (inversion, transformed rejection) are exercised
    # (a) totals per edge over all seeds: standardised residuals are N(0, 1)
    z = (counts.sum(axis=0) - SEEDS * mu) / np.sqrt(SEEDS * mu)
    assert stats.kstest(z, "norm").pvalue > P_MIN
    assert stats.chi2.sf((z ** 2).sum(), E) > P_MIN
    # (b) the whole distribution on a small-mean, a mid and the largest-mean edge
    order = np.argsort(mu)
    for j in (order[len(order) // 4], order[len(order) // 2], order[-1]):
        hi = int(counts[:, j].max()) + 1
        pmf = stats.poisson.pmf(np.arange(hi + 1), mu[j])
        pmf[-1] += stats.poisson.sf(hi, mu[j])
        obs = np.bincount(counts[:, j], minlength=hi + 1)
        assert chi2_p(obs, SEEDS * pmf) > P_MIN, (j, mu[j])
    # (c) index of dispersion of the total count per call
    tot = counts.sum(axis=1)
    d = ((tot - tot.mean()) ** 2).sum() / tot.mean()
    assert P_MIN / 2 < stats.chi2.cdf(d, SEEDS - 1) < 1 - P_MIN / 2


def test_site_counts(ctx, ts_rec):
    ts = ts_rec
    rate = 1.6e-6
    e = ts["edges"]
    bl = branch_lengths(ts)
    # total branch length of the tree covering each integer position
    L = int(ts["sequence_length"])
    diff = np.zeros(L + 1)
    np.add.at(diff, e["left"].astype(int), bl)
    np.add.at(diff, e["right"].astype(int), -bl)
    lam = np.cumsum(diff)[:L] * rate
    expect_sites = (1 - np.exp(-lam)).sum()
    var_sites = (np.exp(-lam) * (1 - np.exp(-lam))).sum()
    sites, muts = [], []
    for res in overlays(ctx, ts, model=cases.BINARY, rate=rate, seeds=range(1, SEEDS + 1)):
        sites.append(len(res["sites"]["position"]))
        muts.append(len(res["mutations"]["site"]))
    sites, muts = np.array(sites), np.array(muts)
    assert muts.mean() > sites.mean() * 1.05  # the genome is dense enough for multiple hits
    z = (sites.mean() - expect_sites) / np.sqrt(var_sites / SEEDS)
    assert 2 * stats.norm.sf(abs(z)) > P_MIN
    zm = (muts.mean() - lam.sum()) / np.sqrt(lam.sum() / SEEDS)
    assert 2 * stats.norm.sf(abs(zm)) > P_MIN
    assert stats.kstest((sites - expect_sites) / np.sqrt(var_sites), "norm").pvalue > P_MIN


def test_allele_transitions_and_roots(ctx, ts_rec):
    # verification.py:5564-5728: transition counts vs row sums x P, root counts vs root distribution
    import msprime_b200 as msp

    hky = msp.HKY(kappa=3.0, equilibrium_frequencies=[0.1, 0.2, 0.3, 0.4])
    model = {"kind": "matrix", "alleles": hky.alleles, "root_distribution": hky.root_distribution.tolist(),
             "transition_matrix": hky.transition_matrix.tolist()}
    P, root = hky.transition_matrix, hky.root_distribution
    trans = np.zeros((4, 4))
    roots = np.zeros(4)
    code = {ord(a): i for i, a in enumerate("ACGT")}
    lut = np.full(256, -1)
    for k, v in code.items():
        lut[k] = v
    for res in overlays(ctx, ts_rec, model=model, rate=3e-7, seeds=range(1, 301)):
        s, m = res["sites"], res["mutations"]
        anc = lut[s["ancestral_state"]]
        der = lut[m["derived_state"]]
        par = m["parent"]
        from_state = np.where(par >= 0, der[np.maximum(par, 0)], anc[m["site"]])
        np.add.at(trans, (from_state, der), 1)
        roots += np.bincount(anc, minlength=4)
    assert trans.sum() > 20000
    assert chi2_p(roots, roots.sum() * root) > P_MIN
    for i in range(4):
        assert chi2_p(trans[i], trans[i].sum() * P[i]) > P_MIN, i


def test_times_and_positions_are_uniform(ctx, ts_rec):
    ts = ts_rec
    e = ts["edges"]
    nt = ts["nodes"]["time"]
    u_time, u_pos = [], []
    for res in overlays(ctx, ts, model=cases.BINARY, rate=1e-7, discrete=False, seeds=range(1, 201)):
        m = res["mutations"]
        x = res["sites"]["position"][m["site"]]
        idx = table_checks.edge_of(ts, m["node"].astype(np.int64), x)
        lo, hi = nt[e["child"][idx]], nt[e["parent"][idx]]
        u_time.append((m["time"] - lo) / (hi - lo))
        u_pos.append((x - e["left"][idx]) / (e["right"][idx] - e["left"][idx]))
    u_time, u_pos = np.concatenate(u_time), np.concatenate(u_pos)
    assert len(u_time) > 20000
    assert stats.kstest(u_time, "uniform").pvalue > P_MIN   # verification.py:5745-5759
    assert stats.kstest(u_pos, "uniform").pvalue > P_MIN


def test_time_window_mass(ctx, ts_rec):
    ts = ts_rec
    lo_t, hi_t = 2000.0, 15000.0
    e, nt = ts["edges"], ts["nodes"]["time"]
    clipped = np.maximum(0, np.minimum(nt[e["parent"]], hi_t) - np.maximum(nt[e["child"]], lo_t))
    rate = 2e-7
    expect = (clipped * (e["right"] - e["left"])).sum() * rate
    tot = [len(r["mutations"]["site"]) for r in overlays(ctx, ts, model=cases.JC69, rate=rate, start_time=lo_t,
                                                         end_time=hi_t, seeds=range(1, 301))]
    z = (np.mean(tot) - expect) / np.sqrt(expect / len(tot))
    assert 2 * stats.norm.sf(abs(z)) > P_MIN


def test_rate_map_bins(ctx, ts_rec):
    # verification.py:5880-5927: counts per rate-map interval vs Poisson mass (non-integer breakpoints)
    ts = ts_rec
    L = ts["sequence_length"]
    rng = np.random.default_rng(4)
    pos = np.concatenate([[0], np.sort(rng.uniform(1, L - 1, 15)), [L]])
    rate = rng.lognormal(np.log(2e-7), 1.0, 16)
    rate[3] = 0.0
    e = ts["edges"]
    bl = branch_lengths(ts)
    mass = np.zeros(16)
    for j in range(16):
        a, b = np.ceil(pos[j]), np.ceil(pos[j + 1])   # discrete genome: integers in [ceil(a), ceil(b))
        span = np.maximum(0, np.minimum(e["right"], b) - np.maximum(e["left"], a))
        mass[j] = (span * bl).sum() * rate[j]
    counts = np.zeros(16)
    n = 300
    for res in overlays(ctx, ts, model=cases.JC69, seeds=range(1, n + 1), map_position=pos.tolist(),
                        map_rate=rate.tolist()):
        x = res["sites"]["position"][res["mutations"]["site"]]
        counts += np.bincount(np.searchsorted(pos, x, side="right") - 1, minlength=16)
    assert counts[3] == 0
    z = (counts - n * mass)[mass > 0] / np.sqrt(n * mass[mass > 0])
    assert stats.chi2.sf((z ** 2).sum(), len(z)) > P_MIN


def test_site_frequency_spectrum(ctx):
    # infinite sites on one tree: E[#mutations above a clade of k samples] = rate x L x branch length
    ts = synth.native_treeseq(10, 1e4, 0.0, 1e4, 9)   # no recombination: a single tree, 20 leaves
    e, nt = ts["edges"], ts["nodes"]["time"]
    n = int((ts["nodes"]["flags"] & 1).sum())
    leaves = np.zeros(len(nt), dtype=np.int64)
    leaves[:n] = 1
    for p, c in sorted(zip(e["parent"], e["child"]), key=lambda pc: nt[pc[0]]):
        leaves[p] += leaves[c]
    rate = 5e-8
    bl = branch_lengths(ts)
    expected = np.zeros(n + 1)
    np.add.at(expected, leaves[e["child"]], bl * ts["sequence_length"] * rate)
    sfs = np.zeros(n + 1)
    for res in overlays(ctx, ts, model=cases.BINARY, rate=rate, discrete=False, seeds=range(1, SEEDS + 1)):
        sfs += np.bincount(leaves[res["mutations"]["node"]], minlength=n + 1)
    assert sfs.sum() > 20000
    assert chi2_p(sfs[1:n], SEEDS * expected[1:n]) > P_MIN


def test_same_distribution_as_reference(ctx, oracle_mod, ts_rec):
    """Two-sample: CUDA under Philox vs the CPU reference algorithm under MT19937."""
    ts = ts_rec
    rate = 3e-7
    L = ts["sequence_length"]
    n = 300
    g_m, g_s, o_m, o_s = [], [], [], []
    for res in overlays(ctx, ts, model=cases.JC69, rate=rate, seeds=range(1, n + 1)):
        g_m.append(len(res["mutations"]["site"]))
        g_s.append(len(res["sites"]["position"]))
    which = "ref" if oracle_mod.have("ref") else "port"
    for seed in range(1, n + 1):
        r = oracle_mod.sim_mutations(ts, which=which, map_position=[0, L], map_rate=[rate], model=cases.JC69,
                                     keep=False, seed=seed)
        o_m.append(len(r["mutations"]["site"]))
        o_s.append(len(r["sites"]["position"]))
    assert stats.ks_2samp(g_m, o_m).pvalue > P_MIN
    assert stats.ks_2samp(g_s, o_s).pvalue > P_MIN
    assert stats.ttest_ind(g_m, o_m, equal_var=False).pvalue > P_MIN
