"""CPU, optional: the file hand-off against the reference's REAL tskit Python package, where it can be
had.  tskit is not installed in this image, but the reference vendors it: `make -C oracle tskit`
(run by `__graft_entry__.build()` where /root/reference exists) compiles its C extension from the
sources where they lie into oracle/_ref/tskit_py/, and the pure-Python part is imported from the
reference tree itself (nothing is copied into this repository).  MSPB_TSKIT_PATH may point at another
tskit checkout.  Skipped wherever neither is there (the GPU box: the reference tree does not travel)."""
import os

import numpy as np
import pytest

import tskit_path

tskit = tskit_path.load()
if tskit is None:
    pytest.skip("the reference's tskit package is not available here", allow_module_level=True)

import cases  # noqa: E402
from msprime_b200 import build, tables, trees_io  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "ref_overlay.trees")


@pytest.fixture(scope="module", autouse=True)
def built():
    build.build_treesio()
    import importlib

    importlib.reload(tables)  # pick tskit up if msprime_b200.tables was imported before the path was set
    importlib.reload(trees_io)


def test_tskit_loads_our_files(tmp_path):
    ts_ref = tskit.load(GOLDEN)
    ours = tmp_path / "ours.trees"
    trees_io.dump(trees_io.load_tables_dict(GOLDEN), ours)
    ts = tskit.load(str(ours))
    assert ts.tables.equals(ts_ref.tables)
    assert ts.num_trees == ts_ref.num_trees and ts.num_mutations == ts_ref.num_mutations
    # the trees are usable: genotypes decode identically
    assert np.array_equal(ts.genotype_matrix(), ts_ref.genotype_matrix())


def test_we_load_tskit_files(tmp_path):
    ts_ref = tskit.load(GOLDEN)
    theirs = tmp_path / "theirs.trees"
    ts_ref.dump(str(theirs))
    d = trees_io.load_tables_dict(theirs)
    want = ts_ref.tables.asdict()
    for tbl in ("nodes", "edges", "sites", "mutations"):
        for col, v in want[tbl].items():
            if isinstance(v, np.ndarray):
                got = np.asarray(d[tbl][col])
                if col.endswith("_offset"):
                    got, v = got.astype(np.uint64), v.astype(np.uint64)
                assert np.array_equal(got.view(np.uint8), v.view(np.uint8)), (tbl, col)
    assert d["sequence_length"] == ts_ref.sequence_length and d["time_units"] == ts_ref.time_units


def test_load_returns_a_real_tree_sequence_and_dump_takes_one(tmp_path):
    ts = trees_io.load(GOLDEN)
    assert isinstance(ts, tskit.TreeSequence)
    ref = tskit.load(GOLDEN)
    assert ts.tables.equals(ref.tables)
    out = tmp_path / "again.trees"
    trees_io.dump(ts, out)
    assert tskit.load(str(out)).tables.equals(ref.tables)


def test_dict_protocol_round_trip():
    """What sim_mutations does around the device call: asdict -> replace sites / mutations ->
    TableCollection.fromdict (msprime/mutations.py:1686-1702)."""
    ts = tskit.load(GOLDEN)
    d, kind = tables.from_tree_sequence(ts)
    assert kind == "tskit"
    from msprime_b200 import mutations

    s, m = d["sites"], d["mutations"]
    result = {"sites": {k: np.asarray(s[k]).view(np.uint8) if k in ("ancestral_state", "metadata") else s[k]
                        for k in ("position", "ancestral_state", "ancestral_state_offset", "metadata", "metadata_offset")},
              "mutations": {k: np.asarray(m[k]).view(np.uint8) if k in ("derived_state", "metadata") else m[k]
                            for k in ("site", "node", "parent", "time", "derived_state", "derived_state_offset",
                                      "metadata", "metadata_offset")}}
    out = mutations._merge_result(d, result)
    ts2 = tables.to_tree_sequence(out, kind, '{"software": "test"}')
    assert isinstance(ts2, tskit.TreeSequence)
    assert ts2.tables.equals(ts.tables, ignore_provenance=True)
    assert ts2.num_provenances == ts.num_provenances + 1
