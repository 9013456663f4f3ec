"""Host-only properties of the mutation-model builders, the ones the reference's own suite asserts
(tests/test_mutations.py: TestMatrixMutationModel :173-443, TestMicrosatModels :445-525,
TestMutationModelFactory :2271-2318, TestModelClasses :2320-2333).  The matrices themselves are
held bit for bit by tests/test_model_golden.py; this file covers shapes, stochasticity,
stationarity, the relations between the named models, argument errors and the factory."""
import numpy as np
import pytest

import msprime_b200 as msp

EQ = [0.1, 0.2, 0.3, 0.4]

NAMED = {
    "binary": lambda: msp.BinaryMutationModel(),
    "jc69": lambda: msp.JC69(),
    "hky": lambda: msp.HKY(kappa=0.75, equilibrium_frequencies=EQ),
    "f84": lambda: msp.F84(kappa=0.75, equilibrium_frequencies=EQ),
    "gtr": lambda: msp.GTR([1] * 6),
    "pam": lambda: msp.PAM(),
    "blosum62": lambda: msp.BLOSUM62(),
    "smm": lambda: msp.SMM(lo=2, hi=50),
    "tpm": lambda: msp.TPM(p=0.9, m=0.9, lo=2, hi=50),
    "tpm_wide": lambda: msp.TPM(p=0.5, m=0.25, lo=50, hi=500),
    "el2": lambda: msp.EL2(m=0.43, u=0.68, v=0.037),
    "microsat_a": lambda: msp.MicrosatMutationModel(s=0.88, u=0.62, v=0.015, p=0, m=1.0, lo=1, hi=10.0),
    "microsat_b": lambda: msp.MicrosatMutationModel(s=0.5, u=0.68, v=0.037, p=1.0, m=0.43, lo=3.0, hi=10),
    "microsat_c": lambda: msp.MicrosatMutationModel(s=0.88, u=0.58, v=0.37, p=0, m=0.83, lo=1, hi=20),
    "microsat_d": lambda: msp.MicrosatMutationModel(s=1, u=0.18, v=0.5, p=1.0, m=0.03, lo=1, hi=4),
}


@pytest.mark.parametrize("name", sorted(NAMED))
def test_model_is_stochastic_and_printable(name):
    model = NAMED[name]()
    n = len(model.alleles)
    assert len(model.root_distribution) == n
    assert model.transition_matrix.shape == (n, n)
    assert np.isclose(np.sum(model.root_distribution), 1.0)
    assert np.allclose(model.transition_matrix.sum(axis=1), 1.0)
    assert str(model).startswith("Mutation model with alleles")


@pytest.mark.parametrize("name", sorted(set(NAMED) - {"binary"}))
def test_root_distribution_is_stationary(name):
    model = NAMED[name]()
    assert np.allclose(model.root_distribution @ model.transition_matrix, model.root_distribution)


def same_model(a, b):
    return (len(a.alleles) == len(b.alleles) and np.array_equal(a.root_distribution, b.root_distribution)
            and np.array_equal(a.transition_matrix, b.transition_matrix))


def test_nested_nucleotide_models_collapse():
    jc = msp.JC69()
    assert same_model(msp.HKY(1.0), jc)
    assert same_model(msp.F84(1.0), jc)
    assert same_model(msp.GTR([1] * 6), jc)
    assert same_model(msp.F84(1.0, equilibrium_frequencies=EQ), msp.HKY(1.0, equilibrium_frequencies=EQ))
    assert same_model(msp.GTR([1.0] * 6, equilibrium_frequencies=EQ), msp.HKY(1.0, equilibrium_frequencies=EQ))


@pytest.mark.parametrize("build, kwargs", [
    (msp.BinaryMutationModel, {}),
    (msp.JC69, {}),
    (msp.HKY, {"kappa": 1.0, "equilibrium_frequencies": EQ}),
    (msp.F84, {"kappa": 1.0, "equilibrium_frequencies": EQ}),
    (msp.GTR, {"relative_rates": [1] * 6, "equilibrium_frequencies": EQ}),
])
def test_state_independent_variants(build, kwargs):
    plain = build(**kwargs).transition_matrix
    indep = build(**kwargs, state_independent=True).transition_matrix
    assert np.allclose(indep, indep[0])  # every row is the same distribution
    # off the diagonal the two matrices differ by one constant factor
    off = ~np.eye(len(plain), dtype=bool)
    ratio = indep[off] / plain[off]
    assert np.allclose(ratio, ratio.mean())


def test_user_defined_model():
    alleles = ["Alligator", "Camel", "Goat"]
    root = [0.5, 0.25, 0.25]
    matrix = [[0.0, 0.5, 0.5], [1.0, 0.0, 0.0], [1.0, 0.0, 0.0]]
    model = msp.MatrixMutationModel(alleles, root, matrix)
    assert model.alleles == alleles
    assert list(model.root_distribution) == root
    assert [list(row) for row in model.transition_matrix] == matrix
    assert np.allclose(model.root_distribution @ model.transition_matrix, root)


def uniform_args(alleles):
    try:
        n = len(alleles)
    except TypeError:
        n = 1
    return dict(alleles=alleles, root_distribution=[1] + [0] * (n - 1), transition_matrix=[[1 / n] * n] * n)


@pytest.mark.parametrize("alleles, err", [
    (0, TypeError), ("xyz", TypeError), ([b"0", "1"], TypeError), ([1, 2, 3], TypeError), (["a"], msp.LibraryError)])
def test_bad_alleles(alleles, err):
    with pytest.raises(err):
        msp.MatrixMutationModel(**uniform_args(alleles))


@pytest.mark.parametrize("root, err", [
    ([1, 0], ValueError), ([1, 0, 0, 0], ValueError), (1.0, ValueError), ("xyz", ValueError), (None, ValueError),
    ([1.5, -0.5, 0], msp.LibraryError), ([0.5, 0, 0], msp.LibraryError)])
def test_bad_root_distribution(root, err):
    with pytest.raises(err):
        msp.MatrixMutationModel(alleles=["a", "b", "c"], root_distribution=root, transition_matrix=[[1 / 3] * 3] * 3)


@pytest.mark.parametrize("matrix, err", [
    ([[1, 0]] * 3, ValueError), ([[1, 0, 0]] * 2, ValueError), (1.0, ValueError), ("xyz", ValueError),
    (None, ValueError), ([[1.5, -0.5, 0]] * 3, msp.LibraryError), ([[0.5, 0, 0]] * 3, msp.LibraryError)])
def test_bad_transition_matrix(matrix, err):
    with pytest.raises(err):
        msp.MatrixMutationModel(alleles=["a", "b", "c"], root_distribution=[0.5, 0.25, 0.25],
                                transition_matrix=matrix)


def test_smm_band_structure():
    t = msp.SMM(lo=2, hi=50).transition_matrix
    assert np.array_equal(t, t.T)
    i, j = np.indices(t.shape)
    assert (t[np.abs(i - j) > 1] == 0).all()
    assert (t[np.abs(i - j) == 1] == t[0, 1]).all()


def geometric_ratio(p, m, lo, hi, i):
    up = p + (m / (1 - (1 - m) ** (hi - i))) * (1 - p)
    down = p + (m / (1 - (1 - m) ** (i - lo))) * (1 - p)
    return up / down


@pytest.mark.parametrize("p, m, lo, hi", [(0.9, 0.9, 2, 50), (0.5, 0.25, 50, 500)])
def test_tpm_step_ratio(p, m, lo, hi):
    t = msp.TPM(p=p, m=m, lo=lo, hi=hi).transition_matrix
    for i in range(lo + 1, hi):
        k = i - lo
        assert np.isclose(t[k, k + 1] / t[k, k - 1], geometric_ratio(p, m, lo, hi, i))


def test_el2_step_ratio_and_direction_bias():
    lo, hi, m, u, v = 1, 10, 0.43, 0.68, 0.037
    t = msp.EL2(lo=lo, hi=hi, m=m, u=u, v=v).transition_matrix
    for i in range(lo + 1, hi):
        k = i - lo
        a = u - v * (i - lo)
        bias = a / (1 - a)
        assert abs(t[k, k + 1] / t[k, k - 1] - geometric_ratio(0, m, lo, hi, i) * bias) < 1e-6
        assert abs(t[k, k:].sum() / t[k, :k].sum() - bias) < 1e-6


@pytest.mark.parametrize("m, u, v, p, s", [
    (1.0, 0.62, 0.015, 0, 0.88), (0.43, 0.68, 0.037, 0, 0.88), (0.43, 0.68, 0.037, 0.5, 0.88),
    (0.43, 0.68, 0.037, 1.0, 0.88)])
def test_microsat_total_rates(m, u, v, p, s):
    lo, hi = 1, 10
    t = msp.MicrosatMutationModel(s=s, u=u, v=v, p=p, m=m, lo=lo, hi=hi).transition_matrix
    leave = t.sum(axis=1) - t.diagonal()
    beta = np.array([1 + (i - lo) * s for i in range(lo, hi + 1)])
    alpha = np.array([u - v * (i - lo) for i in range(lo, hi + 1)])
    # the reference states these as one-sided bounds (tests/test_mutations.py:466-473)
    assert (leave[0] / leave - beta[0] / beta * alpha[0] / (1 - alpha) < 1e-6).all()
    assert (leave[0] / leave[2:-1] - beta[1] / beta[2:-1] < 1e-6).all()


@pytest.mark.parametrize("m, u, v, p, s, lo, hi", [
    (1.0, 0.62, 0.015, 0, 0.88, 1, np.inf), (0.43, 0.68, 0.037, 0, 0.88, np.inf, 10),
    (0.43, 0.68, 0.037, 0, 0.88, 10, 1), (0.43, 0.68, 0.037, 0, 0.88, -1, 4), (0.43, 0.68, 0.037, 1.0, -10, 1, 4),
    (0.43, 10, 0.037, 0, 0.88, 1, 5)])
def test_microsat_bad_parameters(m, u, v, p, s, lo, hi):
    with pytest.raises(ValueError):
        msp.MicrosatMutationModel(s=s, u=u, v=v, p=p, m=m, lo=lo, hi=hi)


@pytest.mark.parametrize("build, kwargs, err, match", [
    (msp.TPM, dict(m=0.9, p=1.1), ValueError, "p must be"),
    (msp.TPM, dict(m=1.1, p=0.9), ValueError, "m must be"),
    (msp.TPM, dict(p=0.1), TypeError, None),
    (msp.TPM, dict(m=0.1), TypeError, None),
    (msp.EL2, dict(m=1.1, u=0.6, v=0.3), ValueError, "m must be"),
    (msp.EL2, dict(m=0.9, u=1.1, v=0.6), ValueError, "u must be"),
    (msp.EL2, dict(u=0.6, v=0.5), TypeError, None),
    (msp.EL2, dict(m=0.6, v=0.5), TypeError, None),
    (msp.EL2, dict(m=0.6, u=0.5), TypeError, None),
    (msp.EL2, dict(v=np.inf, m=0.9, u=0.6), ValueError, "finite"),
    (msp.EL2, dict(v=np.nan, m=0.9, u=0.6), ValueError, "finite"),
])
def test_tpm_el2_argument_errors(build, kwargs, err, match):
    with pytest.raises(err, match=match):
        build(**kwargs)


def test_factory_names_and_types():
    factory = msp.mutations.mutation_model_factory
    for bad in ["NOT", "", "MODEL", "gtr", "slim"]:
        with pytest.raises(ValueError):
            factory(bad)
    for bad in [1234, {}]:
        with pytest.raises(TypeError):
            factory(model=bad)
    classes = {"infinite_alleles": msp.InfiniteAlleles, "binary": msp.BinaryMutationModel, "jc69": msp.JC69,
               "blosum62": msp.BLOSUM62, "pam": msp.PAM}
    for name, cls in classes.items():
        for spelling in (name, name.upper(), name.title()):
            assert isinstance(factory(model=spelling), cls)


def test_factory_passes_instances_through():
    for model in [msp.SLiMMutationModel(0, 0), msp.InfiniteAlleles(), msp.BinaryMutationModel(), msp.JC69(),
                  msp.HKY(0.75), msp.F84(0.75), msp.GTR([1 / 6] * 6), msp.BLOSUM62(), msp.PAM()]:
        before = dict(model.__dict__)
        assert msp.mutations.mutation_model_factory(model=model) is model
        assert model.__dict__.keys() == before.keys()


def test_slim_and_infinite_alleles_descriptions():
    m = msp.SLiMMutationModel(type=1, next_id=2, slim_generation=9)
    assert (m.next_id, m.slim_generation) == (2, 9)
    assert str(m) == "Mutation model for SLiM mutations of type m1\n  next ID: 2\n"
    m = msp.InfiniteAlleles(start_allele=1)
    assert str(m) == "Infinite alleles mutation model, beginning with allele 1\n    next allele: 1\n"
