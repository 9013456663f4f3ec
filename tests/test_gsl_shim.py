"""The GSL boundary of the oracle (oracle/gsl_shim), pinned as far as this box allows.

GSL is a system library of the reference (not vendored, not version-pinned, absent here), so the
shim restates its published algorithms.  What can be checked without a libgsl:

* MT19937 (raw words, seeding, the [0, 1) scaling) bit for bit against numpy's independent
  implementation of the same generator with the same 2002 initialiser, and against the
  generator's published known answer (10000th word for seed 5489 = 4123659995);
* `gsl_ran_flat`, `gsl_rng_uniform_pos` on top of it;
* the distributions (`gsl_ran_poisson` on both sides of mu = 10, `gsl_ran_gamma_int` on both sides
  of a = 12, `gsl_ran_binomial` in the inverse-CDF and BTPE regimes and with p > 1/2) against
  scipy's exact distributions: chi-square / KS at 2e5 draws, p > 1e-4, fixed seeds.

What stays unpinned is only WHICH uniforms a real libgsl would consume for mu > 10 (bit identity of
the free-running MT19937 stream there); the replay comparisons do not depend on it.
"""
import ctypes
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def shim():
    from oracle import oracle

    oracle.build(("port",))
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_build", "libmutgen_port.so"))
    vp, dbl, uint, ulong = ctypes.c_void_p, ctypes.c_double, ctypes.c_uint, ctypes.c_ulong
    for name, res, args in (
            ("gsl_rng_alloc", vp, [vp]), ("gsl_rng_free", None, [vp]), ("gsl_rng_set", None, [vp, ulong]),
            ("gsl_rng_get", ulong, [vp]), ("gsl_rng_uniform", dbl, [vp]), ("gsl_rng_uniform_pos", dbl, [vp]),
            ("gsl_ran_flat", dbl, [vp, dbl, dbl]), ("gsl_ran_poisson", uint, [vp, dbl]),
            ("gsl_ran_gamma_int", dbl, [vp, uint]), ("gsl_ran_binomial", uint, [vp, dbl, uint]),
            ("gsl_fcmp", ctypes.c_int, [dbl, dbl, dbl])):
        fn = getattr(lib, name)
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


@pytest.mark.parametrize("seed", [1, 42, 4357, 5489, 2**31 - 1, 2**32 - 1])
def test_mt19937_words_match_numpy(shim, rng, seed):
    shim.gsl_rng_set(rng, seed)
    got = np.array([shim.gsl_rng_get(rng) for _ in range(2000)], dtype=np.uint64)
    assert np.array_equal(got, numpy_words(seed, 2000))


def test_mt19937_known_answers(shim, rng):
    # the generator's published check value (also the C++ standard's for std::mt19937)
    shim.gsl_rng_set(rng, 5489)
    assert [shim.gsl_rng_get(rng) for _ in range(10000)][-1] == 4123659995
    # a fresh generator and seed 0 both mean GSL's default seed, 4357
    fresh = shim.gsl_rng_alloc(shim.mt19937)
    shim.gsl_rng_set(rng, 0)
    a = [shim.gsl_rng_get(fresh) for _ in range(1000)]
    b = [shim.gsl_rng_get(rng) for _ in range(1000)]
    shim.gsl_rng_free(fresh)
    assert a == b == list(numpy_words(4357, 1000))
    # GSL's own published values for its default generator: "first value = 4293858116" in the manual's example
    # run, and the 1000th output its rng/test.c checks
    assert a[0] == 4293858116 and a[-1] == 1186927261


def test_uniform_scaling_and_flat(shim, rng):
    words = numpy_words(99, 3000).astype(np.float64)
    shim.gsl_rng_set(rng, 99)
    u = np.array([shim.gsl_rng_uniform(rng) for _ in range(1000)])
    assert np.array_equal(u, words[:1000] / 4294967296.0)
    f = np.array([shim.gsl_ran_flat(rng, -3.5, 12.25) for _ in range(1000)])
    w = words[1000:2000] / 4294967296.0
    assert np.array_equal(f, -3.5 * (1 - w) + 12.25 * w)
    p = np.array([shim.gsl_rng_uniform_pos(rng) for _ in range(1000)])
    assert (p > 0).all() and (p < 1).all()
    assert np.array_equal(p, (words[2000:3000] / 4294967296.0))  # no zero word among these


def chi_square_p(observed_values, pmf, lo, hi):
    """Chi-square p-value of integer draws against `pmf` on [lo, hi] with both tails pooled."""
    from scipy import stats

    n = len(observed_values)
    ks = np.arange(lo, hi + 1)
    expected = pmf(ks) * n
    counts = np.bincount(np.clip(observed_values, lo - 1, hi + 1) - (lo - 1), minlength=hi - lo + 3).astype(float)
    below, above = counts[0], counts[-1]
    body = counts[1:-1]
    tail = n - expected.sum()  # mass outside [lo, hi], both sides pooled
    obs = np.append(body, below + above)
    exp = np.append(expected, max(tail, 1e-12))
    keep = exp >= 5
    obs = np.append(obs[keep], obs[~keep].sum())
    exp = np.append(exp[keep], exp[~keep].sum())
    if exp[-1] < 5:  # pooled remainder too thin: fold it into its neighbour
        obs[-2] += obs[-1]
        exp[-2] += exp[-1]
        obs, exp = obs[:-1], exp[:-1]
    stat = ((obs - exp) ** 2 / exp).sum()
    return stats.chi2.sf(stat, len(obs) - 1)


N_DRAWS = 200_000


@pytest.mark.parametrize("mu", [0.05, 1.0, 9.99, 10.01, 13.7, 37.0, 250.0, 5000.0])
def test_poisson_distribution(shim, rng, mu):
    from scipy import stats

    shim.gsl_rng_set(rng, 1234)
    x = np.array([shim.gsl_ran_poisson(rng, mu) for _ in range(N_DRAWS)], dtype=np.int64)
    sd = np.sqrt(mu)
    lo, hi = max(0, int(mu - 5 * sd)), int(mu + 5 * sd) + 2
    assert chi_square_p(x, lambda k: stats.poisson.pmf(k, mu), lo, hi) > 1e-4
    assert abs(x.mean() - mu) < 5 * sd / np.sqrt(N_DRAWS)


def test_poisson_consumes_a_draw_at_zero_mean(shim, rng):
    # lib/mutgen.c:1120 calls it for every edge, zero-mass ones included: one uniform each
    shim.gsl_rng_set(rng, 7)
    assert [shim.gsl_ran_poisson(rng, 0.0) for _ in range(5)] == [0] * 5
    assert shim.gsl_rng_get(rng) == numpy_words(7, 6)[5]


@pytest.mark.parametrize("a", [1, 5, 11, 12, 40, 4375])
def test_gamma_int_distribution(shim, rng, a):
    from scipy import stats

    shim.gsl_rng_set(rng, 4321)
    x = np.array([shim.gsl_ran_gamma_int(rng, a) for _ in range(N_DRAWS // 4)])
    assert stats.kstest(x, stats.gamma(a).cdf).pvalue > 1e-4


@pytest.mark.parametrize("p, n", [(0.1, 20), (0.3, 30), (0.02, 600), (0.4, 100), (0.5, 1000), (0.9, 8),
                                  (0.85, 2000), (0.999, 4374)])
def test_binomial_distribution(shim, rng, p, n):
    from scipy import stats

    shim.gsl_rng_set(rng, 999)
    x = np.array([shim.gsl_ran_binomial(rng, p, n) for _ in range(N_DRAWS)], dtype=np.int64)
    assert x.min() >= 0 and x.max() <= n
    sd = np.sqrt(n * p * (1 - p))
    lo, hi = max(0, int(n * p - 5 * sd)), min(n, int(n * p + 5 * sd) + 2)
    assert chi_square_p(x, lambda k: stats.binom.pmf(k, n, p), lo, hi) > 1e-4


def test_fcmp_is_knuths_relative_compare(shim):
    eps = 1e-9
    assert shim.gsl_fcmp(1.0, 1.0 + 1e-12, eps) == 0
    assert shim.gsl_fcmp(1.0, 1.0 + 1e-6, eps) == -1
    assert shim.gsl_fcmp(1.0 + 1e-6, 1.0, eps) == 1
    assert shim.gsl_fcmp(1e6, 1e6 + 1e-4, eps) == 0  # tolerance scales with the larger exponent
    assert shim.gsl_fcmp(1e-6, 1e-6 + 1e-12, eps) == -1
    assert shim.gsl_fcmp(0.0, 0.0, eps) == 0
