"""`.trees` file hand-off: the product's kastore reader / writer
(msprime_b200/csrc/trees_io.c, msprime_b200/trees_io.py) against the reference's own tskit C
(tsk_table_collection_dump / _load / _equals, driven through oracle/_ref) and against a
committed reference-written fixture.  CPU only."""
import os
import struct

import numpy as np
import pytest

import cases
from msprime_b200 import build, trees_io

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "ref_overlay.trees")
GOLDEN_COLUMNS = os.path.join(os.path.dirname(__file__), "golden", "ref_overlay_columns.npz")


@pytest.fixture(scope="module", autouse=True)
def built():
    build.build_treesio()


def have_ref(oracle_mod):
    return oracle_mod.have("ref")


def overlay_tables(oracle_mod, model=cases.MULTICHAR, rate=2e-6):
    ts = dict(cases.smc_ts(20, 5e4, 3))
    res = oracle_mod.sim_mutations(ts, which="port", map_position=[0, ts["sequence_length"]], map_rate=[rate],
                                   model=model, seed=11)
    return cases.with_kept(ts, res)


def without_uuid(data, store):
    """File bytes with the 36 uuid bytes blanked (the only part that differs between two dumps)."""
    idx = [k for k, _ in store.items()].index("uuid")
    (array_start,) = struct.unpack_from("<Q", data, 64 + 64 * idx + 24)
    return data[:array_start] + b"\0" * 36 + data[array_start + 36:]


def test_header_symbols_exported():
    """_treesio.so exports every function include/msprime_b200_io.h declares."""
    import ctypes
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "include", "msprime_b200_io.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = sorted(set(re.findall(r"\b(mspb_[a-z_]+)\s*\(", text)))
    h = ctypes.CDLL(trees_io.SO_PATH)
    for name in names:
        assert hasattr(h, name), name
    assert sorted(trees_io.EXPORTED_SYMBOLS) == names


def test_reads_reference_fixture():
    """A file written by the reference (tests/golden/make_trees_golden.py) reads back with the
    exact columns the reference had in memory."""
    want = np.load(GOLDEN_COLUMNS)
    d = trees_io.load_tables_dict(GOLDEN)
    assert d["sequence_length"] == float(want["sequence_length"])
    assert d["time_units"] == "generations"
    for tbl in ("nodes", "edges", "sites", "mutations"):
        for col in d[tbl]:
            key = f"{tbl}/{col}"
            if key in want.files:
                got = np.asarray(d[tbl][col])
                exp = want[key]
                assert got.shape == exp.shape, key
                assert np.array_equal(got.view(np.uint8), exp.astype(got.dtype).view(np.uint8)), key
    assert len(d["provenances"]) == 1 and d["provenances"][0]["record"] == '{"software": "oracle"}'
    assert d["indexes"]["edge_insertion_order"].dtype == np.int32
    # the native struct points at the same memory (zero copy)
    with trees_io.Store(GOLDEN) as st:
        t, units = st.native_tables()
        assert units == "generations"
        assert t.num_edges == len(want["edges/left"]) and t.num_mutations == len(want["mutations/site"])
        assert t.edge_left == st.get("edges/left").ctypes.data


def test_rewrite_of_fixture_is_byte_identical():
    """load -> dump reproduces the reference's file byte for byte (uuid aside): same key order,
    types, offsets widths and padding."""
    import tempfile

    with tempfile.TemporaryDirectory() as tmp:
        out = os.path.join(tmp, "again.trees")
        ts = trees_io.load(GOLDEN)
        trees_io.dump(ts, out)
        a, b = open(GOLDEN, "rb").read(), open(out, "rb").read()
        assert len(a) == len(b)
        with trees_io.Store(GOLDEN) as sa, trees_io.Store(out) as sb:
            assert [k for k, _ in sa.items()] == [k for k, _ in sb.items()]
            assert without_uuid(a, sa) == without_uuid(b, sb)


def test_reference_loads_our_file_as_equal(oracle_mod, tmp_path):
    if not have_ref(oracle_mod):
        pytest.skip("compiled reference not available")
    for model in (cases.MULTICHAR, cases.INF_ALLELES, cases.SLIM):
        ts = overlay_tables(oracle_mod, model)
        ref_path, our_path = tmp_path / "ref.trees", tmp_path / "ours.trees"
        oracle_mod.trees_dump(ts, ref_path, num_samples=40, time_units="generations")
        # our reader on the reference's file, our writer, then the reference's comparison
        loaded = trees_io.load(ref_path)
        trees_io.dump(loaded, our_path)
        assert oracle_mod.trees_equal(ref_path, our_path)
        got = oracle_mod.trees_load(our_path)
        want = oracle_mod.trees_load(ref_path)
        assert not oracle_mod.tables_equal(want, got)
        # and the reference's re-dump of OUR file is what it makes of its own
        again = tmp_path / "again.trees"
        oracle_mod.trees_copy(our_path, again)
        assert oracle_mod.trees_equal(ref_path, again)


def test_dump_from_dict_without_optional_tables(oracle_mod, tmp_path):
    """A bare dict (nodes, edges only) gets every table the loader requires, with defaults."""
    ts = cases.four_leaf_tree()
    p = tmp_path / "bare.trees"
    trees_io.dump(ts, p)
    d = trees_io.load_tables_dict(p)
    assert len(d["mutations"]["site"]) == 0 and len(d["sites"]["position"]) == 0
    assert np.array_equal(d["nodes"]["population"], np.full(7, -1, dtype=np.int32))
    assert d["populations"]["metadata_offset"].tolist() == [0]
    if have_ref(oracle_mod):
        ref = tmp_path / "ref.trees"
        oracle_mod.trees_dump(ts, ref, num_samples=4)
        got = oracle_mod.trees_load(p)
        assert got["sites"]["position"].size == 0


def test_offsets_are_32_bit_when_they_fit(tmp_path):
    ts = cases.four_leaf_tree()
    p = tmp_path / "x.trees"
    trees_io.dump(ts, p)
    with trees_io.Store(p) as st:
        assert st.get("sites/ancestral_state_offset").dtype == np.uint32
        assert st.get("format/version").tolist() == [12, 7]
        assert bytes(st.get("format/name")) == b"tskit.trees"
        keys = [k for k, _ in st.items()]
        assert keys == sorted(keys)


def _write_raw(path, data):
    with open(path, "wb") as f:
        f.write(data)


def test_container_errors(tmp_path):
    """kastore_read_header / read_descriptors rejections (kastore.c:154-337), reported in the
    numbering the reference stack hands to Python."""
    good = open(GOLDEN, "rb").read()
    p = tmp_path / "bad.trees"

    def status_of(data):
        _write_raw(p, data)
        with pytest.raises(trees_io.FileFormatError) as e:
            trees_io.Store(p)
        return e.value.status

    kas = lambda e: (e ^ (1 << 14)) ^ (1 << 13)  # noqa: E731
    assert status_of(b"") == kas(-14)  # EOF
    assert status_of(good[:40]) == kas(-5)
    assert status_of(b"X" + good[1:]) == kas(-5)  # magic
    assert status_of(good[:8] + struct.pack("<H", 0) + good[10:]) == kas(-6)  # too old
    assert status_of(good[:8] + struct.pack("<H", 2) + good[10:]) == kas(-7)  # too new
    assert status_of(good[:-9]) == kas(-5)  # truncated
    bad_type = bytearray(good)
    bad_type[64] = 10
    assert status_of(bytes(bad_type)) == kas(-8)
    shifted = bytearray(good)
    struct.pack_into("<Q", shifted, 64 + 24, struct.unpack_from("<Q", good, 64 + 24)[0] + 8)
    assert status_of(bytes(shifted)) == kas(-5)  # arrays must be packed
    with pytest.raises(trees_io.FileFormatError):
        trees_io.Store(tmp_path / "missing.trees")


def test_schema_errors(tmp_path):
    """tsk_table_collection_load rejections (tables.c:100-262, 11588-11634)."""
    with trees_io.Store(GOLDEN) as st:
        items = {k: np.array(v) for k, v in st.items()}
    tsk = lambda e: e ^ (1 << 13)  # noqa: E731

    def status_of(changed, drop=()):
        it = {k: v for k, v in items.items() if k not in drop}
        it.update(changed)
        p = tmp_path / "s.trees"
        trees_io.write_items(p, it)
        with pytest.raises(trees_io.FileFormatError) as e:
            trees_io.load_tables_dict(p)
        return e.value.status

    assert status_of({"format/name": np.frombuffer(b"tskit.treez", dtype=np.int8)}) == tsk(-100)
    assert status_of({"format/version": np.array([11, 0], dtype=np.uint32)}) == tsk(-101)
    assert status_of({"format/version": np.array([13, 0], dtype=np.uint32)}) == tsk(-102)
    assert status_of({"sequence_length": np.array([0.0])}) == tsk(-701)
    assert status_of({}, drop=("edges/left",)) == tsk(-103)
    assert status_of({}, drop=("format/name",)) == tsk(-103)
    assert status_of({}, drop=("sites/metadata_offset",)) == tsk(-104)
    assert status_of({"edges/parent": items["edges/parent"].astype(np.int64)}) == tsk(-105)
    assert status_of({"edges/child": items["edges/child"][:-1]}) == tsk(-100)
    off = items["mutations/derived_state_offset"].copy()
    off[-1] += 1
    assert status_of({"mutations/derived_state_offset": off}) == tsk(-200)
    # mutations/time is optional (TSK_COL_OPTIONAL): absent = unknown times
    it = {k: v for k, v in items.items() if k != "mutations/time"}
    p = tmp_path / "notime.trees"
    trees_io.write_items(p, it)
    with trees_io.Store(p) as st:
        t, _ = st.native_tables()
        assert not t.mutation_time


def test_writer_errors(tmp_path):
    kas = lambda e: (e ^ (1 << 14)) ^ (1 << 13)  # noqa: E731
    with pytest.raises(trees_io.FileFormatError) as e:
        trees_io.write_items(tmp_path / "e.trees", {"": np.zeros(1)})
    assert e.value.status == kas(-9)
    assert not os.path.exists(tmp_path / "e.trees")
    # an empty store is a valid kastore file (header only)
    trees_io.write_items(tmp_path / "empty.kas", {})
    with trees_io.Store(tmp_path / "empty.kas") as st:
        assert len(st) == 0
    assert os.path.getsize(tmp_path / "empty.kas") == 64


def test_wide_offsets_round_trip(tmp_path):
    """uint64 offset columns (what tskit writes past 4 GiB, or with TSK_DUMP_FORCE_OFFSET_64) are
    read in place; uint32 ones are widened."""
    with trees_io.Store(GOLDEN) as st:
        items = {k: np.array(v) for k, v in st.items()}
    for k in list(items):
        if k.endswith("_offset"):
            items[k] = items[k].astype(np.uint64)
    p = tmp_path / "wide.trees"
    trees_io.write_items(p, items)
    a, b = trees_io.load_tables_dict(p), trees_io.load_tables_dict(GOLDEN)
    for tbl in ("sites", "mutations"):
        for col in a[tbl]:
            assert np.array_equal(np.asarray(a[tbl][col]).astype(np.int64) if col.endswith("_offset") else a[tbl][col],
                                  np.asarray(b[tbl][col]).astype(np.int64) if col.endswith("_offset") else b[tbl][col]), col
    with trees_io.Store(p) as st:
        t, _ = st.native_tables()
        assert t.site_ancestral_state_offset == st.get("sites/ancestral_state_offset").ctypes.data


def test_cli_parser_mirrors_msp_mutations():
    """Arguments and defaults of `msp mutations` (msprime/cli.py:1132-1173)."""
    from msprime_b200 import cli

    p = cli.get_parser()
    a = p.parse_args(["mutations", "1e-8", "in.trees", "-o", "out.trees"])
    assert a.mutation_rate == 1e-8 and a.input == "in.trees" and a.output == "out.trees"
    assert a.random_seed is None and a.start_time is None and a.end_time is None and a.model == "jc69"
    a = p.parse_args(["mutations", "0.5", "-s", "7", "--start-time", "1", "--end-time", "2.5", "-m", "pam"])
    assert a.input is None and a.output is None and a.random_seed == 7
    assert (a.start_time, a.end_time, a.model) == (1.0, 2.5, "pam")
    with pytest.raises(SystemExit):
        p.parse_args(["mutations", "1e-8", "-m", "hky"])  # needs arguments: not a name (mutations.py:1515-1518)
    with pytest.raises(SystemExit):
        p.parse_args([])


def test_cli_fails_loudly_without_a_gpu(tmp_path):
    """No CPU fallback: without a device the command stops with the library's error."""
    import subprocess
    import sys

    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "msprime_b200", "mutations", "1e-6", GOLDEN, "-o",
                        str(tmp_path / "o.trees"), "-s", "1"], cwd=root, capture_output=True)
    assert r.returncode != 0 and b"no CPU fallback" in r.stderr
    assert not os.path.exists(tmp_path / "o.trees")


@pytest.mark.parametrize("case", range(25))
def test_random_overlays_round_trip_through_both_libraries(oracle_mod, tmp_path, case):
    """Random overlays (random tree, model and rates, from none to thousands of rows): the file our
    writer makes straight from the column dict is, to the reference, EQUAL to its own dump of the same
    columns and passes its integrity check; and our reader returns the columns unchanged."""
    if not have_ref(oracle_mod):
        pytest.skip("compiled reference not available")
    import test_oracle_fuzz as fuzz

    rng = np.random.default_rng(77 + case)
    L = float(rng.choice([2e3, 2e4]))
    n = int(rng.integers(2, 30))
    ts = dict(cases.smc_ts(n, L, int(rng.integers(1, 4))))
    rate = float(rng.choice([0.0, 1e-7, 3e-6, 5e-5]))
    res = oracle_mod.sim_mutations(ts, which="port", map_position=[0, L], map_rate=[rate], model=fuzz.random_model(rng),
                                   discrete_genome=bool(rng.random() < 0.6), seed=int(rng.integers(1, 2**31)))
    ts = cases.with_kept(ts, res)
    ref_path, our_path = tmp_path / "ref.trees", tmp_path / "ours.trees"
    oracle_mod.trees_dump(ts, ref_path, num_samples=2 * n, time_units="generations")
    want = oracle_mod.trees_load(ref_path)
    trees_io.dump(trees_io.load(ref_path), our_path)
    assert oracle_mod.trees_equal(ref_path, our_path)  # includes TSK_CHECK_TREES on both
    got = trees_io.load_tables_dict(our_path)
    for tbl in ("sites", "mutations"):
        for col in res[tbl]:
            a, b = np.asarray(want[tbl][col]), np.asarray(got[tbl][col])
            assert a.shape == b.shape and np.array_equal(a.view(np.uint8), b.astype(a.dtype).view(np.uint8)), (tbl, col)
