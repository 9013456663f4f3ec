"""CPU: every model class builds the alleles, root distribution and transition matrix of the
reference's Python class of the same name and arguments (msprime/mutations.py:179-1400).
Golden vectors: tests/golden/model_matrices.npz, made from the reference by
tests/golden/make_model_golden.py."""
import os

import numpy as np
import pytest

import model_cases
import msprime_b200 as msp

GOLDEN = np.load(os.path.join(os.path.dirname(__file__), "golden", "model_matrices.npz"))


@pytest.mark.parametrize("key,cls,kwargs", model_cases.CASES, ids=[c[0] for c in model_cases.CASES])
def test_model_matches_reference(key, cls, kwargs):
    m = getattr(msp, cls)(**kwargs)
    assert list(m.alleles) == [str(a) for a in GOLDEN[f"{key}/alleles"]]
    want = GOLDEN[f"{key}/transition_matrix"]
    got = np.asarray(m.transition_matrix)
    assert got.shape == want.shape
    # the entries feed running sums compared with uniforms (lib/util.c:400-424): identical doubles
    assert np.array_equal(got.view(np.uint64), want.view(np.uint64)), np.max(np.abs(got - want))
    root_want = GOLDEN[f"{key}/root_distribution"]
    root = np.asarray(m.root_distribution)
    if isinstance(m, msp.MicrosatMutationModel) and "root_distribution" not in kwargs:
        # stationary distribution from LAPACK's eigen-solver: equal up to the library's rounding
        np.testing.assert_allclose(root, root_want, rtol=0, atol=1e-12)
        np.testing.assert_allclose(root @ got, root, atol=1e-12)
    else:
        assert np.array_equal(root.view(np.uint64), root_want.view(np.uint64))


def test_model_map_names():
    # msprime/mutations.py:1509-1519
    assert sorted(msp.mutations.MODEL_MAP) == ["binary", "blosum62", "infinite_alleles", "jc69", "pam"]
    assert isinstance(msp.mutation_model_factory("PAM"), msp.PAM)
    assert isinstance(msp.mutation_model_factory("blosum62"), msp.BLOSUM62)


def test_microsat_argument_errors():
    # msprime/mutations.py:368-385, 456-460, 506-510; tests/test_mutations.py microsat error cases
    with pytest.raises(ValueError, match="lo > hi"):
        msp.SMM(lo=5, hi=4)
    with pytest.raises(ValueError, match="lo must be >= 0"):
        msp.SMM(lo=-1, hi=4)
    with pytest.raises(ValueError, match="finite"):
        msp.MicrosatMutationModel(s=float("inf"))
    with pytest.raises(ValueError, match="s must be"):
        msp.MicrosatMutationModel(s=-1.0, lo=1, hi=10)
    with pytest.raises(ValueError, match="u must be"):
        msp.MicrosatMutationModel(u=1.0)
    with pytest.raises(ValueError, match="p must be"):
        msp.TPM(p=1.0, m=0.5)
    with pytest.raises(ValueError, match="m must be"):
        msp.TPM(p=0.5, m=0.0)
    with pytest.raises(ValueError, match="m must be"):
        msp.EL2(m=1.0, u=0.5, v=0.0)
    with pytest.raises(ValueError, match="u must be"):
        msp.EL2(m=0.5, u=0.0, v=0.0)


def test_smm_is_a_reflecting_walk():
    m = msp.SMM(lo=2, hi=6)
    assert m.alleles == ["2", "3", "4", "5", "6"]
    t = m.transition_matrix
    for k in range(5):
        assert t[k, k] == (0.5 if k in (0, 4) else 0.0)
        if k > 0:
            assert t[k, k - 1] == 0.5
        if k < 4:
            assert t[k, k + 1] == 0.5
    np.testing.assert_allclose(m.root_distribution, 0.2)
