"""Case list shared by ``tests/golden/make_golden.py`` (which runs the compiled,
unmodified reference mutgen in the container that has /root/reference) and the
tests that replay the same cases on the C restatement (CPU) and the CUDA path (GPU).

Every case is fully determined by this file: the input tree sequence comes from a
deterministic builder, the uniform stream from ``numpy.random.default_rng(seed)``.
"""
import numpy as np

import cases

WINDOW = (1e3, 2e4)


def _ratemap(L):
    rng = np.random.default_rng(5)
    pos = np.concatenate([[0], np.sort(rng.uniform(1, L - 1, 12)), [L]])  # non-integer breakpoints
    rate = rng.lognormal(np.log(7e-7), 1.0, 13)
    rate[::4] = 0.0
    return pos.tolist(), rate.tolist()


def build_ts(spec):
    kind = spec[0]
    if kind == "four_leaf":
        return cases.four_leaf_tree()
    if kind == "stick":
        return cases.stick_tree(5, 1.0)
    if kind == "smc":
        return cases.smc_ts(*spec[1:])
    raise ValueError(kind)


def all_cases():
    """Yields dicts: name, ts (spec), model, map_position, map_rate, discrete, keep, base
    (name of the case whose output is the kept input), start_time, end_time, seed, n_stream."""
    out = []

    def add(name, ts, model, pos, rate, discrete=True, keep=False, base=None, window=None, seed=1, n=60_000):
        out.append(dict(name=name, ts=ts, model=model, map_position=pos, map_rate=rate, discrete=discrete,
                        keep=keep, base=base, start_time=-cases.FMAX if window is None else window[0],
                        end_time=cases.FMAX if window is None else window[1], seed=seed, n_stream=n))

    # the hand-built tree of lib/tests/test_mutations.c:22-94
    for model in cases.MODELS:
        for discrete in (True, False):
            for rate in (0.0, 0.3, 4.0):
                add(f"four_leaf-{model}-{'d' if discrete else 'c'}-{rate}", ("four_leaf",), model, [0, 10], [rate],
                    discrete=discrete)
    # zero-rate intervals (tests/test_mutations.py:528-546)
    add("four_leaf-zero_rate_intervals", ("four_leaf",), "binary", [0, 3, 7, 10], [1, 0, 1])
    # stick tree with time windows (tests/test_mutations.py:1134-1165)
    for k in range(4):
        add(f"stick-window-{k}", ("stick",), "binary", [0, 1], [5.0], discrete=False, window=(k, k + 1), seed=3 + k)
    # recombining input: models x genome x window, then a second layer with keep
    smc = ("smc", 20, 2e4, 2)
    L = 2e4
    for model in ("jc69", "binary", "infinite_alleles", "multichar", "slim"):
        for discrete in (True, False):
            for window in (None, WINDOW):
                tag = f"smc-{model}-{'d' if discrete else 'c'}-{'w' if window else 'a'}"
                add(tag, smc, model, [0, L], [7e-7], discrete=discrete, window=window, n=100_000)
                add(tag + "-keep", smc, model, [0, L], [7e-7], discrete=discrete, window=window, keep=True, base=tag,
                    seed=2, n=100_000)
    pos, rate = _ratemap(L)
    for discrete in (True, False):
        add(f"smc-ratemap-{'d' if discrete else 'c'}", smc, "jc69", pos, rate, discrete=discrete, n=100_000)
    # continuous genome, second layer drawn from the SAME stream as the first: every first draw of a
    # position lands on a site that is taken and is drawn again from the next uniform
    # (lib/mutgen.c:1127-1134; the reference forces this with equal seeds, tests/test_mutations.py:1340-1354)
    for model in ("jc69", "binary", "infinite_alleles", "slim"):
        add(f"smc-{model}-c-a-sameseed", smc, model, [0, L], [7e-7], discrete=False, keep=True,
            base=f"smc-{model}-c-a", seed=1, n=100_000)
    add("four_leaf-jc69-c-4.0-sameseed", ("four_leaf",), "jc69", [0, 10], [4.0], discrete=False, keep=True,
        base="four_leaf-jc69-c-4.0", seed=1)
    add("four_leaf-multichar-c-0.3-sameseed", ("four_leaf",), "multichar", [0, 10], [0.3], discrete=False, keep=True,
        base="four_leaf-multichar-c-0.3", seed=1)
    # big mu per edge: exercises the gamma / binomial branch of gsl_ran_poisson
    add("four_leaf-bigmu", ("four_leaf",), "jc69", [0, 10], [12.0], n=60_000)
    return out
