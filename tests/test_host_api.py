"""CPU: the Python surface (models, RateMap, argument handling) mirrors the reference's
behaviour (msprime/mutations.py, msprime/intervals.py; tests/test_mutations.py:271-335,
686-835).  Nothing here needs a GPU: every check fails before the device call."""
import numpy as np
import pytest

import cases
import msprime_b200 as msp
from msprime_b200 import tables as mtables


def light_ts(d):
    return mtables.TableCollection.fromdict(d).tree_sequence()


def test_jc69_matrix():
    m = msp.JC69()
    assert m.alleles == ["A", "C", "G", "T"]
    assert np.array_equal(m.root_distribution, [0.25] * 4)
    want = np.full((4, 4), 1 / 3)
    np.fill_diagonal(want, 0)
    assert np.array_equal(m.transition_matrix, want)
    assert np.array_equal(msp.JC69(state_independent=True).transition_matrix, np.full((4, 4), 0.25))


def test_binary_model():
    m = msp.BinaryMutationModel()
    assert m.alleles == ["0", "1"]
    assert np.array_equal(m.root_distribution, [1, 0])
    assert np.array_equal(m.transition_matrix, [[0, 1], [1, 0]])


def test_hky_rows_and_kappa():
    # tests/test_mutations.py:271-300: rows sum to one, transitions kappa times transversions
    freqs = [0.3, 0.2, 0.2, 0.3]
    m = msp.HKY(kappa=2.0, equilibrium_frequencies=freqs)
    P = m.transition_matrix
    assert np.allclose(P.sum(axis=1), 1.0, atol=1e-15)
    assert np.isclose(P[0, 2] / freqs[2], 2 * P[0, 1] / freqs[1])
    assert np.array_equal(m.root_distribution, freqs)


def test_gtr_reversible():
    freqs = np.array([0.3, 0.2, 0.2, 0.3])
    m = msp.GTR(relative_rates=[1, 2, 1, 1, 2, 1], equilibrium_frequencies=freqs)
    P = m.transition_matrix
    assert np.allclose(P.sum(axis=1), 1.0, atol=1e-15)
    flow = freqs[:, None] * P
    off = ~np.eye(4, dtype=bool)
    assert np.allclose(flow[off], flow.T[off])


def test_f84_is_stochastic():
    P = msp.F84(kappa=1.5, equilibrium_frequencies=[0.1, 0.2, 0.3, 0.4]).transition_matrix
    assert np.allclose(P.sum(axis=1), 1.0) and (P >= 0).all()


@pytest.mark.parametrize("alleles,root,matrix,status", [
    (["A"], [1.0], [[1.0]], -52),
    (["A", "B"], [0.5, 0.6], [[0.5, 0.5], [0.5, 0.5]], -53),
    (["A", "B"], [-0.5, 1.5], [[0.5, 0.5], [0.5, 0.5]], -53),
    (["A", "B"], [0.5, 0.5], [[0.5, 0.6], [0.5, 0.5]], -54),
])
def test_matrix_model_validation(alleles, root, matrix, status):
    # lib/mutgen.c:501-524, lib/tests/test_mutations.c:1152-1265
    with pytest.raises(msp.LibraryError) as e:
        msp.MatrixMutationModel(alleles, root, matrix)
    assert e.value.status == status


def test_matrix_model_shape_errors():
    with pytest.raises(ValueError):
        msp.MatrixMutationModel(["A", "B"], [1.0], [[0.5, 0.5], [0.5, 0.5]])
    with pytest.raises(ValueError):
        msp.MatrixMutationModel(["A", "B"], [0.5, 0.5], [[0.5, 0.5]])
    with pytest.raises(TypeError):
        msp.MatrixMutationModel("AB", [0.5, 0.5], [[0.5, 0.5], [0.5, 0.5]])


def test_model_factory():
    assert isinstance(msp.mutation_model_factory(None), msp.JC69)
    assert isinstance(msp.mutation_model_factory("binary"), msp.BinaryMutationModel)
    assert isinstance(msp.mutation_model_factory("JC69"), msp.JC69)
    assert isinstance(msp.mutation_model_factory("infinite_alleles"), msp.InfiniteAlleles)
    m = msp.JC69()
    assert msp.mutation_model_factory(m) is m
    with pytest.raises(ValueError):
        msp.mutation_model_factory("nope")
    with pytest.raises(TypeError):
        msp.mutation_model_factory(3)


def test_rate_map_validation():
    # msprime/intervals.py:61-87
    r = msp.RateMap(position=[0, 5, 10], rate=[1, 2])
    assert r.sequence_length == 10 and r.num_intervals == 2
    assert np.array_equal(r.asdict()["position"], [0, 5, 10])
    for bad in (dict(position=[1, 10], rate=[1]), dict(position=[0, 5, 5], rate=[1, 1]),
                dict(position=[0, 10], rate=[-1]), dict(position=[0], rate=[]),
                dict(position=[0, 10], rate=[1, 2]), dict(position=[0, 10], rate=[np.nan])):
        with pytest.raises(ValueError):
            msp.RateMap(**bad)
    u = msp.RateMap.uniform(100, 0.5)
    assert np.array_equal(u.position, [0, 100]) and np.array_equal(u.rate, [0.5])


def test_sim_mutations_argument_errors():
    # msprime/mutations.py:1637-1684; tests/test_mutations.py:686-835
    ts = light_ts(cases.four_leaf_tree())
    with pytest.raises(ValueError):
        msp.sim_mutations("not a tree sequence", rate=1)
    with pytest.raises(TypeError):
        msp.sim_mutations(ts, rate=[1.0, 2.0])
    with pytest.raises(ValueError):
        msp.sim_mutations(ts, rate=1, start_time=2, end_time=1)
    with pytest.raises(TypeError):
        msp.sim_mutations(ts, rate=1, keep=1)
    with pytest.raises(TypeError):
        msp.sim_mutations(ts, rate=1, discrete_genome="yes")
    with pytest.raises(ValueError):
        msp.sim_mutations(ts, rate=1, model="nope")
    with pytest.raises(ValueError):
        msp.sim_mutations(ts, rate=1, random_seed=0)
    with pytest.raises(ValueError):
        msp.sim_mutations(ts, rate=1, random_seed=2**32)
    d = cases.four_leaf_tree()
    d["time_units"] = "uncalibrated"
    with pytest.raises(ValueError):
        msp.sim_mutations(light_ts(d), rate=1)


def test_mutate_rejects_general_models():
    ts = light_ts(cases.four_leaf_tree())
    with pytest.raises(ValueError):
        msp.mutate(ts, rate=1, model=msp.JC69())


def test_infinite_alleles_start():
    m = msp.InfiniteAlleles(start_allele=7)
    assert m.start_allele == 7 and m.next_allele == 7
    with pytest.raises(OverflowError):
        msp.InfiniteAlleles(start_allele=-1)


def test_slim_model_parameters():
    # msprime/mutations.py:109-151, tests/test_mutations.py:1554-1600
    m = msp.SLiMMutationModel(type=3, next_id=11, slim_generation=40)
    assert (m.type, m.next_id, m.slim_generation) == (3, 11, 40)
    assert "m3" in str(m) and "next ID: 11" in str(m)
    assert m._native() == {"kind": 2, "type": 3, "next_id": 11, "slim_generation": 40}
    assert msp.SLiMMutationModel(type=0).next_id == 0
    for bad in (dict(type=-1), dict(type=1, next_id=-5)):
        with pytest.raises(msp.LibraryError, match="nonnegative"):
            msp.SLiMMutationModel(**bad)
    with pytest.raises(TypeError):
        msp.SLiMMutationModel(type="m1")
    with pytest.raises(OverflowError):
        msp.SLiMMutationModel(type=1, next_id=2**63)


# The argument errors the reference's suite asserts of sim_mutations (tests/test_mutations.py:756-835,
# TestSimMutations.test_bad_rates / test_bad_models / test_bad_tree_sequence / test_boolean_inputs):
# each must be raised on the host, before any device is touched (there is none here).
@pytest.mark.parametrize("kwargs, err", [
    (dict(rate={}), TypeError), (dict(rate=[]), TypeError), (dict(rate="self"), TypeError),
    (dict(rate="abc"), ValueError), (dict(rate="xxx"), ValueError),
    (dict(rate=-1), ValueError), (dict(rate=-1e-6), ValueError), (dict(rate=-1e7), ValueError),
    (dict(rate=0, model={}), TypeError), (dict(rate=0, model=True), TypeError), (dict(rate=0, model=123), TypeError),
    (dict(rate=0, model=""), ValueError), (dict(rate=0, model="coffee"), ValueError),
    (dict(rate=10, discrete_genome=[]), TypeError), (dict(rate=10, keep=[]), TypeError),
])
def test_sim_mutations_rejects_what_the_reference_rejects(kwargs, err):
    ts = light_ts(cases.four_leaf_tree())
    if kwargs.get("rate") == "self":
        kwargs = dict(kwargs, rate=ts)
    with pytest.raises(err):
        msp.sim_mutations(ts, **kwargs)


@pytest.mark.parametrize("bad", [None, {}, "sdrf"])
def test_sim_mutations_wants_a_tree_sequence(bad):
    with pytest.raises(ValueError, match="TreeSequence"):
        msp.sim_mutations(bad)


@pytest.mark.parametrize("rate", [0, 1])
def test_uncalibrated_time_units_both_interfaces(rate):
    # tests/test_mutations.py:703-713
    d = cases.four_leaf_tree()
    d["time_units"] = "uncalibrated"
    for call in (msp.sim_mutations, msp.mutate):
        with pytest.raises(ValueError, match="uncalibrated"):
            call(light_ts(d), rate=rate, random_seed=1)


def test_provenance_record_contents():
    # msprime/provenance.py:52-220; what tests/test_mutations.py:114-171 and tests/test_provenance.py:170-245
    # read back from the record
    import json

    from msprime_b200 import provenance

    ts = light_ts(cases.four_leaf_tree())
    rate_map = msp.RateMap(position=[0, 0.25, 1.0], rate=[1e-3, 0])
    params = dict(command="sim_mutations", tree_sequence=ts, rate=rate_map, model=msp.HKY(kappa=2.0),
                  start_time=np.int64(-1), end_time=np.array(100), discrete_genome=None, keep=False, random_seed=42)
    text = provenance.json_encode_provenance(provenance.get_provenance_dict(params))
    record = json.loads(text)
    assert record["schema_version"] == "1.0.0"
    assert record["software"]["name"] == "msprime_b200"
    assert {"os", "python", "libraries"} <= set(record["environment"])
    p = record["parameters"]
    assert p["command"] == "sim_mutations"
    assert p["tree_sequence"] == {"__constant__": "__current_ts__"}  # the input is referenced, not embedded
    assert p["random_seed"] == 42 and p["keep"] is False and p["discrete_genome"] is None
    assert p["start_time"]["__npgeneric__"] == "-1" and p["end_time"]["__ndarray__"] == 100
    assert p["rate"]["__class__"] == "msprime_b200.intervals.RateMap"
    assert p["rate"]["position"]["__ndarray__"] == [0, 0.25, 1.0] and p["rate"]["rate"]["__ndarray__"] == [1e-3, 0]
    assert p["model"]["__class__"] == "msprime_b200.mutations.HKY" and p["model"]["kappa"] == 2.0
    # and back: the call can be made again from the record
    command, again = provenance.parse_provenance(text, ts)
    assert command == "sim_mutations" and again["tree_sequence"] is ts
    assert isinstance(again["rate"], msp.RateMap) and again["rate"] == rate_map
    assert isinstance(again["model"], msp.HKY)
    assert np.array_equal(again["model"].transition_matrix, msp.HKY(kappa=2.0).transition_matrix)
    assert again["start_time"] == -1 and again["end_time"] == 100 and again["random_seed"] == 42
    foreign = json.dumps(dict(record, software={"name": "skynet", "version": "1"}))
    with pytest.raises(ValueError, match="Only msprime_b200"):
        provenance.parse_provenance(foreign, None)
    with pytest.raises(TypeError, match="asdict"):
        provenance.json_encode_provenance({"x": object()})


def test_star_import_gives_the_paths_names():
    # tests/test_imports_star.py: `from msprime import *` must work and export the documented names
    ns = {}
    exec("from msprime_b200 import *", ns)
    wanted = {"BinaryMutationModel", "BLOSUM62", "F84", "GTR", "HKY", "InfiniteAlleles", "JC69", "MatrixMutationModel",
              "MicrosatMutationModel", "MutationModel", "SMM", "TPM", "EL2", "NUCLEOTIDES", "PAM", "SLiMMutationModel",
              "sim_mutations", "BINARY", "mutate", "InfiniteSites", "RateMap"}  # msprime/__init__.py, this path's share
    assert wanted <= set(ns)
    assert set(msp.__all__) == wanted | {"LibraryError", "InputError"}
    assert "np" not in ns and "core" not in ns
