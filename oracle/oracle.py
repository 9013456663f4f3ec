"""TEST INFRASTRUCTURE (oracle).  Not part of the product path.

ctypes front end for the two CPU checkers that implement ``oracle_api.h``:

* ``ref``  -- ``oracle/_ref/libmutgen_ref.so``: the unmodified reference
  ``lib/mutgen.c`` (+ util.c, rate_map.c, avl.c, tskit C) built by
  ``oracle/Makefile`` against the GSL shim;
* ``port`` -- ``oracle/_build/libmutgen_port.so``: this project's plain-C
  restatement (``oracle/mutgen_port.c``).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` arm may import this module.  Tables travel as the tskit
``asdict()`` layout (dict of numpy columns).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libmutgen_ref.so")
PORT_SO = os.path.join(HERE, "_build", "libmutgen_port.so")

MODEL_MATRIX = 0
MODEL_INFINITE_ALLELES = 1
MODEL_SLIM = 2

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_up = C.POINTER(C.c_uint64)
_cp = C.c_char_p


class _Input(C.Structure):
    _fields_ = [
        ("sequence_length", C.c_double),
        ("num_nodes", C.c_uint64),
        ("node_time", C.c_void_p),
        ("num_edges", C.c_uint64),
        ("edge_left", C.c_void_p),
        ("edge_right", C.c_void_p),
        ("edge_parent", C.c_void_p),
        ("edge_child", C.c_void_p),
        ("num_sites", C.c_uint64),
        ("site_position", C.c_void_p),
        ("site_ancestral_state", C.c_void_p),
        ("site_ancestral_state_offset", C.c_void_p),
        ("site_metadata", C.c_void_p),
        ("site_metadata_offset", C.c_void_p),
        ("num_mutations", C.c_uint64),
        ("mutation_site", C.c_void_p),
        ("mutation_node", C.c_void_p),
        ("mutation_time", C.c_void_p),
        ("mutation_derived_state", C.c_void_p),
        ("mutation_derived_state_offset", C.c_void_p),
        ("mutation_metadata", C.c_void_p),
        ("mutation_metadata_offset", C.c_void_p),
        ("map_size", C.c_uint64),
        ("map_position", C.c_void_p),
        ("map_rate", C.c_void_p),
        ("model_kind", C.c_int32),
        ("num_alleles", C.c_uint64),
        ("alleles", C.c_void_p),
        ("allele_offset", C.c_void_p),
        ("root_distribution", C.c_void_p),
        ("transition_matrix", C.c_void_p),
        ("start_allele", C.c_uint64),
        ("discrete_genome", C.c_int32),
        ("keep", C.c_int32),
        ("start_time", C.c_double),
        ("end_time", C.c_double),
        ("rng_kind", C.c_int32),
        ("seed", C.c_uint32),
        ("stream", C.c_void_p),
        ("stream_len", C.c_uint64),
        ("slim_mutation_type_id", C.c_int32),
        ("slim_generation", C.c_int32),
    ]


class _Output(C.Structure):
    _fields_ = [
        ("status", C.c_int32),
        ("stream_exhausted", C.c_int32),
        ("draws", C.c_uint64),
        ("next_allele", C.c_uint64),
        ("seconds", C.c_double),
        ("num_sites", C.c_uint64),
        ("site_position", C.c_void_p),
        ("site_ancestral_state", C.c_void_p),
        ("site_ancestral_state_offset", C.c_void_p),
        ("site_metadata", C.c_void_p),
        ("site_metadata_offset", C.c_void_p),
        ("num_mutations", C.c_uint64),
        ("mutation_site", C.c_void_p),
        ("mutation_node", C.c_void_p),
        ("mutation_parent", C.c_void_p),
        ("mutation_time", C.c_void_p),
        ("mutation_derived_state", C.c_void_p),
        ("mutation_derived_state_offset", C.c_void_p),
        ("mutation_metadata", C.c_void_p),
        ("mutation_metadata_offset", C.c_void_p),
    ]


class OracleError(RuntimeError):
    def __init__(self, status):
        super().__init__(f"oracle returned status {status}")
        self.status = status


_libs = {}


def build(which=("port", "ref")):
    """Runs ``make`` for the requested targets (``ref``, ``ancestry`` and ``tskit``, the extension of the
    reference's tskit Python package, are no-ops without /root/reference)."""
    for target in which:
        subprocess.run(["make", "-s", "-C", HERE, target], check=True)


def have(which):
    return os.path.exists(REF_SO if which == "ref" else PORT_SO)


def _load(which):
    if which not in _libs:
        path = REF_SO if which == "ref" else PORT_SO
        lib = C.CDLL(path)
        lib.orc_sim_mutations.argtypes = [C.POINTER(_Input), C.POINTER(_Output)]
        lib.orc_sim_mutations.restype = C.c_int
        lib.orc_free_output.argtypes = [C.POINTER(_Output)]
        lib.orc_name.restype = C.c_char_p
        _libs[which] = lib
    return _libs[which]


def _arr(x, dtype):
    a = np.ascontiguousarray(np.asarray(x, dtype=dtype))
    return a


def _ptr(a):
    return a.ctypes.data if a.size > 0 else None


def _copy(ptr, n, dtype):
    if n == 0 or not ptr:
        return np.zeros(0, dtype=dtype)
    nbytes = n * np.dtype(dtype).itemsize
    buf = C.string_at(ptr, nbytes)
    return np.frombuffer(buf, dtype=dtype).copy()


def _ragged(tbl, name, n):
    """Returns (bytes u8 array, u64 offsets) for column ``name``; missing -> empty."""
    if tbl is not None and name in tbl and name + "_offset" in tbl:
        data = _arr(np.asarray(tbl[name]).view(np.uint8) if np.asarray(tbl[name]).dtype.itemsize == 1
                    else tbl[name], np.uint8)
        off = _arr(tbl[name + "_offset"], np.uint64)
    else:
        data = np.zeros(0, dtype=np.uint8)
        off = np.zeros(n + 1, dtype=np.uint64)
    return data, off


def encode_alleles(alleles):
    """List of str/bytes -> (u8 data, u64 offsets)."""
    bs = [a.encode("utf8") if isinstance(a, str) else bytes(a) for a in alleles]
    off = np.zeros(len(bs) + 1, dtype=np.uint64)
    off[1:] = np.cumsum([len(b) for b in bs])
    data = np.frombuffer(b"".join(bs), dtype=np.uint8).copy()
    return data, off


def _fill_tables(inp, tables, hold):
    """Points the tree-sequence fields of ``inp`` at the columns of the dict ``tables``."""
    inp.sequence_length = float(tables["sequence_length"])
    nt = hold(_arr(tables["nodes"]["time"], np.float64))
    inp.num_nodes = nt.size
    inp.node_time = _ptr(nt)
    e = tables["edges"]
    el, er = hold(_arr(e["left"], np.float64)), hold(_arr(e["right"], np.float64))
    ep, ec = hold(_arr(e["parent"], np.int32)), hold(_arr(e["child"], np.int32))
    inp.num_edges = el.size
    inp.edge_left, inp.edge_right, inp.edge_parent, inp.edge_child = map(_ptr, (el, er, ep, ec))

    s = tables.get("sites")
    ns = 0 if s is None else len(s["position"])
    inp.num_sites = ns
    if ns > 0:
        sp = hold(_arr(s["position"], np.float64))
        sa, sao = map(hold, _ragged(s, "ancestral_state", ns))
        sm, smo = map(hold, _ragged(s, "metadata", ns))
        inp.site_position = _ptr(sp)
        inp.site_ancestral_state, inp.site_ancestral_state_offset = _ptr(sa), _ptr(sao)
        inp.site_metadata, inp.site_metadata_offset = _ptr(sm), _ptr(smo)
    m = tables.get("mutations")
    nm = 0 if m is None else len(m["site"])
    inp.num_mutations = nm
    if nm > 0:
        ms, mn = hold(_arr(m["site"], np.int32)), hold(_arr(m["node"], np.int32))
        inp.mutation_site, inp.mutation_node = _ptr(ms), _ptr(mn)
        if "time" in m and m["time"] is not None:
            mt = hold(_arr(m["time"], np.float64))
            inp.mutation_time = _ptr(mt)
        md, mdo = map(hold, _ragged(m, "derived_state", nm))
        mm, mmo = map(hold, _ragged(m, "metadata", nm))
        inp.mutation_derived_state, inp.mutation_derived_state_offset = _ptr(md), _ptr(mdo)
        inp.mutation_metadata, inp.mutation_metadata_offset = _ptr(mm), _ptr(mmo)



def _export(out):
    """Copies the site / mutation columns of an ``orc_output_t`` into numpy arrays."""
    S, M = out.num_sites, out.num_mutations
    sao = _copy(out.site_ancestral_state_offset, S + 1, np.uint64)
    smo = _copy(out.site_metadata_offset, S + 1, np.uint64)
    mdo = _copy(out.mutation_derived_state_offset, M + 1, np.uint64)
    mmo = _copy(out.mutation_metadata_offset, M + 1, np.uint64)
    res = {
        "sites": {
            "position": _copy(out.site_position, S, np.float64),
            "ancestral_state": _copy(out.site_ancestral_state, int(sao[-1]), np.uint8),
            "ancestral_state_offset": sao,
            "metadata": _copy(out.site_metadata, int(smo[-1]), np.uint8),
            "metadata_offset": smo,
        },
        "mutations": {
            "site": _copy(out.mutation_site, M, np.int32),
            "node": _copy(out.mutation_node, M, np.int32),
            "parent": _copy(out.mutation_parent, M, np.int32),
            "time": _copy(out.mutation_time, M, np.float64),
            "derived_state": _copy(out.mutation_derived_state, int(mdo[-1]), np.uint8),
            "derived_state_offset": mdo,
            "metadata": _copy(out.mutation_metadata, int(mmo[-1]), np.uint8),
            "metadata_offset": mmo,
        }
    }
    return res


def sim_mutations(
    tables,
    *,
    which="ref",
    map_position,
    map_rate,
    model,
    discrete_genome=True,
    keep=True,
    start_time=-np.finfo(np.float64).max,
    end_time=np.finfo(np.float64).max,
    seed=None,
    stream=None,
):
    """Runs one overlay on the CPU checker ``which`` ("ref" or "port").

    ``model`` is a dict: ``{"kind": "matrix", "alleles": [...], "root_distribution": [...],
    "transition_matrix": [[...]]}`` or ``{"kind": "infinite_alleles", "start_allele": int}``.
    Exactly one of ``seed`` (MT19937) and ``stream`` (replay) must be given.
    Returns a dict with ``sites`` and ``mutations`` column dicts plus bookkeeping.
    """
    lib = _load(which)
    keep_refs = []

    def hold(a):
        keep_refs.append(a)
        return a

    inp = _Input()
    _fill_tables(inp, tables, hold)

    mpos = hold(_arr(map_position, np.float64))
    mrate = hold(_arr(map_rate, np.float64))
    inp.map_size = mrate.size
    inp.map_position, inp.map_rate = _ptr(mpos), _ptr(mrate)

    if model["kind"] == "matrix":
        inp.model_kind = MODEL_MATRIX
        ad, ao = map(hold, encode_alleles(model["alleles"]))
        rd = hold(_arr(model["root_distribution"], np.float64))
        tm = hold(_arr(model["transition_matrix"], np.float64))
        inp.num_alleles = len(model["alleles"])
        inp.alleles = ad.ctypes.data  # may be empty but must be non-NULL for pointer maths
        inp.allele_offset = _ptr(ao)
        inp.root_distribution, inp.transition_matrix = _ptr(rd), _ptr(tm)
    elif model["kind"] == "infinite_alleles":
        inp.model_kind = MODEL_INFINITE_ALLELES
        inp.start_allele = int(model.get("start_allele", 0)) & 0xFFFFFFFFFFFFFFFF
    elif model["kind"] == "slim":
        # {"kind": "slim", "type": mutation type id, "next_id": ..., "slim_generation": ...}
        inp.model_kind = MODEL_SLIM
        inp.start_allele = int(model.get("next_id", 0)) & 0xFFFFFFFFFFFFFFFF
        inp.slim_mutation_type_id = int(model["type"])
        inp.slim_generation = int(model.get("slim_generation", 1))
    else:
        raise ValueError("unknown model kind")

    inp.discrete_genome = int(bool(discrete_genome))
    inp.keep = int(bool(keep))
    inp.start_time = float(start_time)
    inp.end_time = float(end_time)
    if (seed is None) == (stream is None):
        raise ValueError("give exactly one of seed / stream")
    if stream is not None:
        st = hold(_arr(stream, np.float64))
        inp.rng_kind = 1
        inp.stream, inp.stream_len = _ptr(st), st.size
    else:
        inp.rng_kind = 0
        inp.seed = int(seed)

    out = _Output()
    status = lib.orc_sim_mutations(C.byref(inp), C.byref(out))
    try:
        if status != 0:
            raise OracleError(status)
        res = _export(out)
        res.update({
            "draws": int(out.draws),
            "stream_exhausted": bool(out.stream_exhausted),
            "next_allele": int(out.next_allele),
            "seconds": float(out.seconds),
        })
        return res
    finally:
        lib.orc_free_output(C.byref(out))


def tables_equal(a, b):
    """Bit-exact comparison of the sites/mutations columns of two results; returns list of diffs."""
    diffs = []
    for tbl in ("sites", "mutations"):
        for col in a[tbl]:
            x, y = np.asarray(a[tbl][col]), np.asarray(b[tbl][col])
            if x.dtype.kind == "f":
                same = x.shape == y.shape and np.array_equal(
                    x.view(np.uint64), np.asarray(y, dtype=np.float64).view(np.uint64))
            else:
                same = x.shape == y.shape and np.array_equal(
                    x.astype(np.int64, copy=False) if x.dtype.kind in "iu" else x,
                    y.astype(np.int64, copy=False) if y.dtype.kind in "iu" else y)
            if not same:
                diffs.append(f"{tbl}.{col}")
    return diffs


# ---- `.trees` files through the reference's own tskit C (reference build only): the checker
# of the product's file reader / writer

def _trees_lib():
    lib = _load("ref")
    if not hasattr(lib, "_trees_ready"):
        lib.orc_trees_dump.argtypes = [C.POINTER(_Input), C.c_uint64, C.c_char_p, C.c_char_p]
        lib.orc_trees_load.argtypes = [C.c_char_p, C.POINTER(_Output)]
        lib.orc_trees_equal.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(C.c_int32)]
        lib.orc_trees_copy.argtypes = [C.c_char_p, C.c_char_p]
        lib._trees_ready = True
    return lib


def trees_dump(tables, path, num_samples=0, time_units="generations"):
    """Writes ``tables`` (nodes, edges, sites, mutations) with ``tsk_table_collection_dump``."""
    keep_refs = []

    def hold(a):
        keep_refs.append(a)
        return a

    inp = _Input()
    _fill_tables(inp, tables, hold)
    status = _trees_lib().orc_trees_dump(C.byref(inp), int(num_samples), time_units.encode(), str(path).encode())
    if status != 0:
        raise OracleError(status)


def trees_load(path):
    """Site and mutation tables of a file as ``tsk_table_collection_load`` reads them."""
    lib = _trees_lib()
    out = _Output()
    status = lib.orc_trees_load(str(path).encode(), C.byref(out))
    try:
        if status != 0:
            raise OracleError(status)
        return _export(out)
    finally:
        lib.orc_free_output(C.byref(out))


def trees_equal(path_a, path_b):
    """True when the reference loads both files and finds the table collections equal."""
    which = C.c_int32()
    r = _trees_lib().orc_trees_equal(str(path_a).encode(), str(path_b).encode(), C.byref(which))
    if r < 0:
        raise OracleError(r)
    return r == 1


def trees_copy(path_in, path_out):
    status = _trees_lib().orc_trees_copy(str(path_in).encode(), str(path_out).encode())
    if status != 0:
        raise OracleError(status)
