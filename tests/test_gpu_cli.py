"""File in, file out on the GPU: the `mutations` command line
(msprime/cli.py:1043-1055) reads a reference-written `.trees` file, overlays on the device
and writes a `.trees` file whose tables equal the in-memory call with the same seed."""
import os
import subprocess
import sys

import numpy as np
import pytest

import table_checks
from msprime_b200 import mutations, trees_io

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "ref_overlay.trees")


def run_cli(*args, stdin=None):
    r = subprocess.run([sys.executable, "-m", "msprime_b200", *args], cwd=ROOT, input=stdin, capture_output=True)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    return r


@pytest.mark.parametrize("model", ["jc69", "infinite_alleles"])
def test_cli_matches_in_memory_call(native_built, tmp_path, model):
    out = tmp_path / "out.trees"
    run_cli("mutations", "1e-6", GOLDEN, "-o", str(out), "-s", "17", "-m", model,
            "--start-time", "10", "--end-time", "50000")
    got = trees_io.load_tables_dict(out)
    ts_in = trees_io.load(GOLDEN)
    want_ts = mutations.sim_mutations(ts_in, rate=1e-6, random_seed=17, model=model, keep=True,
                                      start_time=10, end_time=50000, discrete_genome=True)
    want = want_ts.tables.asdict()
    for tbl in ("sites", "mutations"):
        for col, v in want[tbl].items():
            if not isinstance(v, np.ndarray):
                assert v == got[tbl][col], (tbl, col)
                continue
            a, b = np.asarray(v), np.asarray(got[tbl][col])
            if col.endswith("_offset"):
                a, b = a.astype(np.uint64), b.astype(np.uint64)
            assert a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8)), (tbl, col)
    # kept rows of the input survive, new ones were added, nothing else changed
    src = trees_io.load_tables_dict(GOLDEN)
    assert len(got["mutations"]["site"]) > len(src["mutations"]["site"])
    for tbl in ("nodes", "edges", "indexes"):
        for col, v in src[tbl].items():
            if isinstance(v, np.ndarray):
                assert np.array_equal(v, got[tbl][col]), (tbl, col)
    assert len(got["provenances"]) == len(src["provenances"]) + 1
    table_checks.check_result(src, {"sites": got["sites"], "mutations": got["mutations"]}, discrete=True,
                              kept_mutations=len(src["mutations"]["site"]))


def test_cli_streams(native_built, tmp_path):
    """'-' / omitted arguments mean standard input and output, as in the reference."""
    data = open(GOLDEN, "rb").read()
    r = run_cli("mutations", "1e-6", "-s", "3", stdin=data)
    p = tmp_path / "piped.trees"
    p.write_bytes(r.stdout)
    d = trees_io.load_tables_dict(p)
    assert len(d["mutations"]["site"]) > len(trees_io.load_tables_dict(GOLDEN)["mutations"]["site"])


def test_reference_accepts_cli_output(native_built, oracle_mod, tmp_path):
    if not oracle_mod.have("ref"):
        pytest.skip("compiled reference not available")
    out, again = tmp_path / "out.trees", tmp_path / "again.trees"
    run_cli("mutations", "1e-6", GOLDEN, "-o", str(out), "-s", "5")
    # the reference loads it, passes its integrity check with trees, and re-dumps it equal
    oracle_mod.trees_copy(out, again)
    assert oracle_mod.trees_equal(out, again)
    got, ours = oracle_mod.trees_load(out), trees_io.load_tables_dict(out)
    assert np.array_equal(got["mutations"]["parent"], ours["mutations"]["parent"])


def test_cli_reports_library_errors(native_built, tmp_path):
    """Kept nucleotide alleles under the binary model: MSP_ERR_UNKNOWN_ALLELE, as the reference."""
    r = subprocess.run([sys.executable, "-m", "msprime_b200", "mutations", "1e-5", GOLDEN, "-o",
                        str(tmp_path / "x.trees"), "-s", "2", "-m", "binary"], cwd=ROOT, capture_output=True)
    assert r.returncode != 0
    assert b"incompatible with mutation model alphabet" in r.stderr
    assert not os.path.exists(tmp_path / "x.trees")
