"""CPU: the Python surface (`sim_mutations`, `mutate`) with REAL `tskit.TreeSequence` objects in and
out, after the reference's own suite (tests/test_mutations.py: TestDefaults :44, TestMutateProvenance
:114, TestSimMutations :686, TestKeep :1231, MutateMixin :593).  There is no GPU here and no tskit on
the GPU box, so this is the one place the tskit side of the host layer meets the real library: the
device call is answered by a stand-in on the CPU oracle (tests/oracle_engine.py; the pattern of
tests/test_sharding_gloo.py) that receives the very parameter struct the device would.  What is under
test is everything around that call: table hand-off in both directions, what is kept and what is
replaced, provenance, argument handling, the model's allele counter."""
import importlib
import json

import numpy as np
import pytest

import tskit_path

tskit = tskit_path.load()
if tskit is None:
    pytest.skip("the reference's tskit package is not available here", allow_module_level=True)

import cases  # noqa: E402
import oracle_engine  # noqa: E402
import msprime_b200 as msp  # noqa: E402
from msprime_b200 import mutations, tables  # noqa: E402


@pytest.fixture(autouse=True)
def engine(monkeypatch):
    importlib.reload(tables)  # pick tskit up if the module was imported before it became importable
    monkeypatch.setattr(mutations, "_tables", tables)
    eng = oracle_engine.OracleContext()
    monkeypatch.setattr(mutations, "_context", lambda device: eng)
    return eng


def base_ts(samples=10, L=2e4, seed=1, extras=True):
    """A recombining tree sequence as a real tskit object, with the tables the overlay must not touch."""
    d = cases.smc_ts(samples, L, seed)
    t = tskit.TableCollection(d["sequence_length"])
    t.time_units = "generations"
    n = len(d["nodes"]["time"])
    pop = np.arange(n, dtype=np.int32) % 2 if extras else np.full(n, -1, dtype=np.int32)
    ind = np.full(n, -1, dtype=np.int32)
    if extras:
        for _ in range(2):
            t.populations.add_row(metadata=b"pop")
        t.individuals.add_row(flags=3, location=[1.5, 2.5], metadata=b"someone")
        ind[0] = 0
        t.migrations.add_row(left=0, right=L, node=0, source=0, dest=1, time=0.5)
    t.nodes.set_columns(flags=d["nodes"]["flags"].astype(np.uint32), time=d["nodes"]["time"], population=pop,
                        individual=ind)
    e = d["edges"]
    t.edges.set_columns(left=e["left"], right=e["right"], parent=e["parent"], child=e["child"])
    t.provenances.add_row(record=json.dumps({"software": "test", "parameters": {"command": "sim_ancestry"}}))
    return t.tree_sequence()


def last_record(ts):
    return json.loads(ts.provenance(ts.num_provenances - 1).record)


def assert_rest_untouched(before, after):
    """MutateMixin.verify_topology / verify_provenance, tests/test_mutations.py:593-616."""
    a, b = before.dump_tables(), after.dump_tables()
    for name in ("nodes", "edges", "populations", "migrations", "individuals"):
        assert getattr(a, name) == getattr(b, name), name
    assert a.sequence_length == b.sequence_length and a.time_units == b.time_units
    assert len(b.provenances) == len(a.provenances) + 1
    for j in range(len(a.provenances)):
        assert a.provenances[j] == b.provenances[j]


def test_provenance_record_passes_tskits_schema():
    # tests/test_provenance.py:55-64: tskit.validate_provenance on what the call writes
    out = msp.sim_mutations(base_ts(), rate=msp.RateMap(position=[0, 1e4, 2e4], rate=[1e-6, 0]), random_seed=3,
                            model=msp.GTR([1, 2, 1, 1, 2, 1]))
    tskit.validate_provenance(last_record(out))


def test_defaults_add_nothing_but_a_provenance_record():
    ts = base_ts()
    out = msp.sim_mutations(ts)
    assert isinstance(out, tskit.TreeSequence)
    assert out.num_sites == 0 and out.num_mutations == 0
    assert_rest_untouched(ts, out)
    assert ts.num_provenances == 1  # the input is never modified
    assert last_record(out)["parameters"]["command"] == "sim_mutations"


def test_result_is_a_valid_tree_sequence_with_consistent_parents_and_times():
    ts = base_ts(samples=20)
    out = msp.sim_mutations(ts, rate=2e-5, random_seed=3, model=msp.HKY(kappa=2.0))
    assert out.num_mutations > 100 and out.num_sites <= out.num_mutations
    assert_rest_untouched(ts, out)
    t = out.dump_tables()
    parent = t.mutations.parent.copy()
    t.compute_mutation_parents()  # tskit's own derivation from the trees
    assert np.array_equal(parent, t.mutations.parent)
    assert (parent >= 0).any()  # some sites were hit more than once
    node_time = t.nodes.time
    assert (t.mutations.time >= node_time[t.mutations.node]).all()
    for tree in out.trees():
        for site in tree.sites():
            assert site.ancestral_state in "ACGT"
            for m in site.mutations:
                assert m.derived_state in "ACGT"
                assert m.time < tree.time(tree.parent(m.node))
    assert np.all(np.floor(t.sites.position) == t.sites.position)  # discrete genome by default


@pytest.mark.parametrize("rate", [0, 1e-6, 1e-5])
def test_provenance_rate_and_seed(rate):
    out = msp.sim_mutations(base_ts(), rate)
    p = last_record(out)["parameters"]
    assert p["command"] == "sim_mutations" and p["rate"] == rate and p["random_seed"] > 0
    assert p["tree_sequence"] == {"__constant__": "__current_ts__"}


def test_provenance_times_keep_and_switch():
    ts = base_ts()
    for start, end in ((0, 1), (-1, 100), (1, None)):
        p = last_record(msp.sim_mutations(ts, 1e-6, start_time=start, end_time=end))["parameters"]
        assert p["start_time"] == start and p["end_time"] == end
    for seed in range(1, 6):
        assert last_record(msp.sim_mutations(ts, rate=1e-6, random_seed=seed))["parameters"]["random_seed"] == seed
    for keep in (True, False):
        assert last_record(msp.sim_mutations(ts, rate=1e-6, keep=keep))["parameters"]["keep"] == keep
    assert msp.sim_mutations(ts, rate=1e-6, record_provenance=True).num_provenances == 2
    assert msp.sim_mutations(ts, rate=1e-6, record_provenance=False).num_provenances == 1


def test_default_seeds_differ_and_fixed_seeds_repeat():
    ts = base_ts()
    seeds = [last_record(msp.sim_mutations(ts, 0))["parameters"]["random_seed"] for _ in range(10)]
    assert len(set(seeds)) == 10
    runs = [msp.sim_mutations(ts, rate=1e-5, random_seed=2) for _ in range(3)]
    assert runs[0].num_mutations > 0
    for other in runs[1:]:
        assert other.tables.sites == runs[0].tables.sites and other.tables.mutations == runs[0].tables.mutations
    assert msp.sim_mutations(ts, rate=1e-5, random_seed=3).tables.mutations != runs[0].tables.mutations


def test_numpy_scalars_are_accepted():
    ts = base_ts()
    v = np.array([0, 1, 2.5e-5, 1e6])
    a = msp.sim_mutations(ts, rate=v[2], random_seed=v[1], start_time=v[0], end_time=v[3])
    b = msp.sim_mutations(ts, rate=float(v[2]), random_seed=int(v[1]), start_time=float(v[0]), end_time=float(v[3]))
    assert a.equals(b, ignore_timestamps=True)


def test_zero_rate_keeps_and_keep_false_discards():
    ts = base_ts()
    layer = msp.sim_mutations(ts, rate=1e-5, random_seed=5)
    assert layer.num_mutations > 0
    same = msp.sim_mutations(layer, 0)
    assert same.tables.sites == layer.tables.sites and same.tables.mutations == layer.tables.mutations
    assert_rest_untouched(layer, same)
    wiped = msp.sim_mutations(layer, 0, keep=False)
    assert wiped.num_sites == 0 and wiped.num_mutations == 0
    assert_rest_untouched(layer, wiped)
    assert layer.num_mutations > 0  # still there in the input


def test_layers_accumulate_and_metadata_survives():
    ts = base_ts(samples=12)
    t = msp.sim_mutations(ts, rate=1e-5, random_seed=6, model="binary").dump_tables()
    # give the first layer metadata, as a SLiM-style producer would
    md = [f"m{j}".encode() for j in range(len(t.mutations))]
    t.mutations.packset_metadata(md)
    t.sites.packset_metadata([b"s%d" % j for j in range(len(t.sites))])
    first = t.tree_sequence()
    second = msp.sim_mutations(first, rate=1e-5, random_seed=7, model="binary")
    assert second.num_mutations > first.num_mutations and second.num_sites >= first.num_sites
    kept_md = {m.metadata for m in second.mutations() if m.metadata}
    assert kept_md == set(md)
    by_pos = {s.position: s.metadata for s in second.sites()}
    for s in first.sites():
        assert by_pos[s.position] == s.metadata
    t2 = second.dump_tables()
    parent = t2.mutations.parent.copy()
    t2.compute_mutation_parents()
    assert np.array_equal(parent, t2.mutations.parent)


def test_kept_mutations_need_times():
    # tests/test_mutations.py:1417-1432: existing mutations with unknown times cannot be layered over
    t = msp.sim_mutations(base_ts(), rate=1e-5, random_seed=8).dump_tables()
    t.mutations.time = np.full(len(t.mutations), tskit.UNKNOWN_TIME)
    with pytest.raises(msp.LibraryError, match="known mutation times"):
        msp.sim_mutations(t.tree_sequence(), rate=1e-5, random_seed=9, keep=True)
    assert msp.sim_mutations(t.tree_sequence(), rate=1e-5, random_seed=9, keep=False).num_mutations > 0


def test_unicode_alleles_on_a_continuous_genome():
    alleles = ["\U0001F384\U0001F333", "\U0001F4A9" * 5]
    model = msp.MatrixMutationModel(alleles, [1.0, 0.0], [[0.0, 1.0], [1.0, 0.0]])
    out = msp.sim_mutations(base_ts(samples=8), rate=2e-5, random_seed=1, model=model, discrete_genome=False)
    assert out.num_sites > 0 and out.num_sites == out.num_mutations
    for site in out.sites():
        assert site.ancestral_state == alleles[0]
        assert [m.derived_state for m in site.mutations] == [alleles[1]]


def test_time_window():
    ts = base_ts(samples=15)
    out = msp.sim_mutations(ts, rate=1e-4, random_seed=4, start_time=50, end_time=500)
    t = out.tables.mutations.time
    assert len(t) > 0 and (t >= 50).all() and (t < 500).all()
    with pytest.raises(ValueError, match="start_time"):
        msp.sim_mutations(ts, rate=1, start_time=2, end_time=1)


def test_rate_map_confines_mutations():
    ts = base_ts(samples=15)
    L = ts.sequence_length
    rate_map = msp.RateMap(position=[0, 5000, 7000.5, L], rate=[0, 1e-4, 0])
    out = msp.sim_mutations(ts, rate=rate_map, random_seed=4)
    pos = out.tables.sites.position
    assert len(pos) > 0 and (pos >= 5000).all() and (pos <= 7000).all()
    p = last_record(out)["parameters"]["rate"]
    assert p["position"]["__ndarray__"] == [0, 5000, 7000.5, L] and p["__class__"].endswith("RateMap")
    # the record is enough to make the call again (tests/test_provenance.py:240-245, 340-352)
    from msprime_b200 import provenance

    command, params = provenance.parse_provenance(out.provenance(out.num_provenances - 1), ts)
    assert command == "sim_mutations" and params["tree_sequence"] is ts and params["rate"] == rate_map
    again = getattr(msp, command)(**params)
    assert again.tables.sites == out.tables.sites and again.tables.mutations == out.tables.mutations
    with pytest.raises(msp.InputError, match="rate map"):
        msp.sim_mutations(ts, rate=msp.RateMap(position=[0, L / 2], rate=[1e-5]))


def test_infinite_alleles_counter_is_advanced():
    ts = base_ts()
    model = msp.InfiniteAlleles(start_allele=10)
    out = msp.sim_mutations(ts, rate=1e-5, random_seed=2, model=model)
    assert model.next_allele == 10 + out.num_sites + out.num_mutations
    names = [int(s.ancestral_state) for s in out.sites()] + [int(m.derived_state) for m in out.mutations()]
    assert sorted(names) == list(range(10, model.next_allele))
    more = msp.sim_mutations(out, rate=1e-5, random_seed=3, model=model)  # the counter carries on
    fresh = {int(m.derived_state) for m in more.mutations()} | {int(s.ancestral_state) for s in more.sites()}
    assert max(fresh) == model.next_allele - 1 and len(fresh) == more.num_sites + more.num_mutations


def test_slim_model_ids_and_metadata():
    model = msp.SLiMMutationModel(type=2, next_id=100, slim_generation=7)
    out = msp.sim_mutations(base_ts(), rate=1e-5, random_seed=2, model=model)
    assert out.num_mutations > 0 and model.next_id == 100 + out.num_mutations
    for site in out.sites():
        assert site.ancestral_state == ""
    ids = []
    for m in out.mutations():
        stack = [int(x) for x in m.derived_state.split(",")]
        ids.append(stack[-1])
        assert len(m.metadata) == 17 * len(stack)
    assert sorted(ids) == list(range(100, model.next_id))


def test_legacy_mutate():
    ts = base_ts()
    out = msp.mutate(ts, rate=1e-5, random_seed=2)
    assert out.num_sites == out.num_mutations > 0  # infinite sites
    assert {s.ancestral_state for s in out.sites()} == {"0"} and {m.derived_state for m in out.mutations()} == {"1"}
    pos = out.tables.sites.position
    assert not np.all(np.floor(pos) == pos)
    assert last_record(out)["parameters"]["command"] == "sim_mutations"
    again = msp.mutate(out, rate=1e-5, random_seed=3)  # keep=False by default here
    assert set(again.tables.sites.position).isdisjoint(pos)
    kept = msp.mutate(out, rate=1e-5, random_seed=3, keep=True)
    assert set(pos) <= set(kept.tables.sites.position)
    with pytest.raises(ValueError, match="legacy"):
        msp.mutate(ts, rate=1, model=msp.JC69())


def test_one_device_call_per_overlay(engine):
    ts = base_ts()
    msp.sim_mutations(ts, rate=1e-5, random_seed=1)
    msp.sim_mutations(ts, rate=1e-5, random_seed=2)
    assert engine.calls == 2


def test_command_line_on_tskit_files(tmp_path):
    """tests/test_cli.py:1522-1628 (msp mutations output): a file written by tskit goes through our
    reader, the overlay and our writer; tskit loads the result; keep=True and a discrete genome are
    what the command line means; the time window and the model option are honoured."""
    from msprime_b200 import cli

    first = msp.sim_mutations(base_ts(), rate=1e-5, random_seed=2)
    src, dst = tmp_path / "in.trees", tmp_path / "out.trees"
    first.dump(str(src))
    cli.main(["mutations", "1e-5", str(src), "-o", str(dst), "-s", "3"])
    out = tskit.load(str(dst))
    assert out.num_mutations > first.num_mutations  # the existing ones are kept
    assert set(first.tables.sites.position) <= set(out.tables.sites.position)
    assert_rest_untouched(first, out)
    p = last_record(out)["parameters"]
    assert p["keep"] is True and p["discrete_genome"] is True and p["random_seed"] == 3 and p["model"] == "jc69"
    assert np.all(np.floor(out.tables.sites.position) == out.tables.sites.position)
    assert {m.derived_state for m in out.mutations()} <= set("ACGT")

    bare = tmp_path / "bare.trees"
    base_ts().dump(str(bare))
    cli.main(["mutations", "1e-4", str(bare), "-o", str(dst), "-s", "4", "--start-time", "0", "--end-time", "200",
              "-m", "binary"])
    out = tskit.load(str(dst))
    t = out.tables.mutations.time
    assert len(t) > 0 and (t >= 0).all() and (t < 200).all()
    assert {m.derived_state for m in out.mutations()} <= {"0", "1"}
    # the same seed through the in-memory call gives the same tables
    again = msp.sim_mutations(base_ts(), rate=1e-4, random_seed=4, start_time=0, end_time=200, model="binary")
    assert again.tables.sites == out.tables.sites and again.tables.mutations == out.tables.mutations


def test_legacy_alphabets():
    """TestMutate, tests/test_mutations.py:618-684."""
    ts = base_ts()
    default = msp.mutate(ts, rate=1e-5, random_seed=2)
    binary = msp.mutate(ts, rate=1e-5, random_seed=2, model=msp.InfiniteSites(msp.BINARY))
    nucs = msp.mutate(ts, rate=1e-5, random_seed=2, model=msp.InfiniteSites(msp.NUCLEOTIDES))
    assert default.tables.sites == binary.tables.sites and default.tables.mutations == binary.tables.mutations
    assert {s.ancestral_state for s in binary.sites()} == {"0"}
    assert {m.derived_state for m in binary.mutations()} == {"1"}
    assert {s.ancestral_state for s in nucs.sites()} <= set("ACGT") and len({s.ancestral_state for s in nucs.sites()}) > 1
    assert {m.derived_state for m in nucs.mutations()} <= set("ACGT")
    # the alphabet does not move the mutations: same seed, same places, same branches
    assert binary.num_sites == nucs.num_sites > 0 and binary.num_mutations == nucs.num_mutations
    for a, b in zip(binary.sites(), nucs.sites()):
        assert a.position == b.position and a.mutations[0].node == b.mutations[0].node
    plain = msp.sim_mutations(ts, rate=1e-5, random_seed=2)  # the default here is nucleotides
    assert {m.derived_state for m in plain.mutations()} <= set("ACGT")
    for bad in (-1, 3, "binary"):
        with pytest.raises(ValueError, match="alphabet"):
            msp.InfiniteSites(bad)
    for model in (msp.PAM(), "jc69", msp.BinaryMutationModel()):
        with pytest.raises(ValueError, match="legacy"):
            msp.mutate(ts, rate=1, random_seed=2, model=model)
