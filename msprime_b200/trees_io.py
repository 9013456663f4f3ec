"""``.trees`` files on either side of the overlay.

The reference's command line reads its input with ``tskit.load`` and writes the result with
``TreeSequence.dump`` (``msprime/cli.py:1043-1055``); both are kastore containers written
by ``tsk_table_collection_dump`` (``git-submodules/tskit/c/tskit/tables.c:12049-12140``).
This module is the ctypes binding of ``include/msprime_b200_io.h`` (``_treesio.so``,
``csrc/trees_io.c``) plus the tskit key schema:

* :func:`load` maps a file and returns a tree sequence whose columns are numpy views INTO the
  mapping (no copy: the overlay uploads them to the GPU straight from the page cache);
* :func:`dump` writes a table collection with the key set, array types, ordering and padding of
  ``tsk_table_collection_dump``, so that the reference loads it back as an equal collection.
"""
from __future__ import annotations

import ctypes as C
import datetime
import os
import uuid as _uuid

import numpy as np

from . import tables as _tables
from ._native import LibraryError

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "_treesio.so")

EXPORTED_SYMBOLS = [
    "mspb_store_open", "mspb_store_close", "mspb_store_num_items", "mspb_store_item", "mspb_store_get",
    "mspb_store_tables", "mspb_store_write", "mspb_io_strerror",
]

# kastore.h:108-117
_KAS_DTYPES = [np.int8, np.uint8, np.int16, np.uint16, np.int32, np.uint32, np.int64, np.uint64,
               np.float32, np.float64]
_KAS_OF = {np.dtype(t): i for i, t in enumerate(_KAS_DTYPES)}
I8, U8, I32, U32, U64, F64 = 0, 1, 4, 5, 7, 9

FORMAT_NAME = "tskit.trees"
FORMAT_VERSION = (12, 7)  # core.h:206-209
UNKNOWN_TIME = np.array([0x7FF8000000000001], dtype=np.uint64).view(np.float64)[0]  # core.h TSK_UNKNOWN_TIME

# table -> (fixed-width columns (name, type, default or None = required), ragged columns (name, type))
# in the order of tsk_*_table_dump (tables.c:1880, 2551, 3244, 4000, 4803, 5473, 6060, 6690)
_SCHEMA = {
    "individuals": ([("flags", U32, None)], [("location", F64), ("parents", I32), ("metadata", U8)]),
    "nodes": ([("time", F64, None), ("flags", U32, 0), ("population", I32, -1), ("individual", I32, -1)],
              [("metadata", U8)]),
    "edges": ([("left", F64, None), ("right", F64, None), ("parent", I32, None), ("child", I32, None)],
              [("metadata", U8)]),
    "migrations": ([("left", F64, None), ("right", F64, None), ("node", I32, None), ("source", I32, None),
                    ("dest", I32, None), ("time", F64, None)], [("metadata", U8)]),
    "sites": ([("position", F64, None)], [("ancestral_state", U8), ("metadata", U8)]),
    "mutations": ([("site", I32, None), ("node", I32, None), ("parent", I32, -1), ("time", F64, UNKNOWN_TIME)],
                  [("derived_state", U8), ("metadata", U8)]),
    "populations": ([], [("metadata", U8)]),
    "provenances": ([], [("timestamp", U8), ("record", U8)]),
}


class _Item(C.Structure):
    _fields_ = [("key", C.c_void_p), ("key_len", C.c_uint64), ("type", C.c_int32), ("array", C.c_void_p),
                ("array_len", C.c_uint64)]


_lib_handle = None


def lib():
    global _lib_handle
    if _lib_handle is None:
        if not os.path.exists(SO_PATH):
            raise ImportError(f"{SO_PATH} is not built: run `python -m msprime_b200.build`")
        h = C.CDLL(SO_PATH)
        h.mspb_store_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        h.mspb_store_close.argtypes = [C.c_void_p]
        h.mspb_store_close.restype = None
        h.mspb_store_num_items.argtypes = [C.c_void_p]
        h.mspb_store_num_items.restype = C.c_uint64
        h.mspb_store_item.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(_Item)]
        h.mspb_store_get.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(_Item)]
        h.mspb_store_tables.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
        h.mspb_store_write.argtypes = [C.c_char_p, C.POINTER(_Item), C.c_uint64]
        h.mspb_io_strerror.argtypes = [C.c_int]
        h.mspb_io_strerror.restype = C.c_char_p
        _lib_handle = h
    return _lib_handle


class FileFormatError(LibraryError):
    """The file is not a usable ``.trees`` file (mirrors ``tskit.FileFormatError``)."""


def _check(status, what):
    if status != 0:
        msg = lib().mspb_io_strerror(status).decode()
        raise FileFormatError(f"{what}: {msg}", status)


class Store:
    """An open kastore file: read-only numpy views of its arrays, keyed by name."""

    def __init__(self, path):
        self._h = C.c_void_p()
        self.path = os.fspath(path)
        _check(lib().mspb_store_open(self.path.encode(), C.byref(self._h)), f"Cannot load '{self.path}'")

    def close(self):
        if self._h:
            lib().mspb_store_close(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _view(self, item):
        dtype = np.dtype(_KAS_DTYPES[item.type])
        nbytes = int(item.array_len) * dtype.itemsize
        if nbytes == 0:
            return np.zeros(0, dtype=dtype)
        buf = (C.c_char * nbytes).from_address(item.array)
        buf._store = self  # the mapping must outlive every view of it
        a = np.frombuffer(buf, dtype=dtype)
        a.flags.writeable = False
        return a

    def __len__(self):
        return int(lib().mspb_store_num_items(self._h))

    def items(self):
        it = _Item()
        for j in range(len(self)):
            _check(lib().mspb_store_item(self._h, j, C.byref(it)), "item")
            # keys are bytes to kastore; one that is not UTF-8 is simply not a key the schema knows
            yield C.string_at(it.key, it.key_len).decode("utf8", "surrogateescape"), self._view(it)

    def get(self, key):
        it = _Item()
        _check(lib().mspb_store_get(self._h, key.encode(), C.byref(it)), f"key '{key}'")
        return self._view(it)

    def native_tables(self):
        """The overlay's input struct (``mspb_tables_t``) pointing into the mapping, after the
        schema checks of ``tsk_table_collection_load``; returns ``(struct, time_units)``."""
        from ._native import _Tables

        t = _Tables()
        tu, tu_len = C.c_void_p(), C.c_uint64()
        _check(lib().mspb_store_tables(self._h, C.byref(t), C.byref(tu), C.byref(tu_len)),
               f"Cannot load '{self.path}'")
        t._store = self
        units = _text(np.frombuffer(C.string_at(tu, tu_len.value), dtype=np.uint8)) if tu.value else "unknown"
        return t, units


def _text(a) -> str:
    """A text item (time units, schemas, uuid).  The C layers do not look inside; tskit's Python layer
    fails on bytes that are not UTF-8 when the value is read, here that is at load time."""
    try:
        return bytes(np.asarray(a).view(np.uint8)).decode("utf8")
    except UnicodeDecodeError as e:
        raise FileFormatError(f"A text item of the file is not valid UTF-8 ({e})", -100 ^ (1 << 13)) from None


def _fail(path, code, what):
    _check(code ^ (1 << 13), f"Cannot load '{os.fspath(path)}' ({what})")


_OPTIONAL_COLUMNS = {("individuals", "parents"), ("edges", "metadata"), ("migrations", "metadata"), ("mutations", "time")}


def _check_columns(d, path):
    """The column rules of ``read_table_cols`` / ``read_table_ragged_cols`` and ``check_offsets``
    (tables.c:100-262) for EVERY table, as ``tsk_table_collection_load`` applies them (the C entry point
    looks only at what the overlay reads): required columns present (-103), data and offset columns
    together (-104), element types (-105), equal row counts (-100), offsets starting at 0, never
    decreasing and ending at the data length (-200)."""
    if "file_uuid" not in d:  # tables.c: "uuid" is a required item, 36 characters
        _fail(path, -103, "uuid")
    if len(d["file_uuid"]) != 36:
        _fail(path, -100, "uuid")
    idx = d.get("indexes") or {}
    ins, rem = idx.get("edge_insertion_order"), idx.get("edge_removal_order")
    if (ins is None) != (rem is None):  # tables.c:11751-11770: the two index columns come together
        _fail(path, -104, "indexes")
    if ins is not None:
        if ins.dtype != np.int32 or rem.dtype != np.int32:
            _fail(path, -105, "indexes")
        if not len(ins) == len(rem) == len(d.get("edges", {}).get("left", ())):
            _fail(path, -100, "indexes")
    for tbl, (fixed, ragged) in _SCHEMA.items():
        cols = d.get(tbl)
        if not isinstance(cols, dict):
            _fail(path, -103, tbl)
        rows = None
        for name, kas, _ in fixed:
            a = cols.get(name)
            if a is None:
                if (tbl, name) in _OPTIONAL_COLUMNS:
                    continue
                _fail(path, -103, f"{tbl}/{name}")
            if _KAS_OF.get(np.asarray(a).dtype) != kas:
                _fail(path, -105, f"{tbl}/{name}")
            if rows is not None and len(a) != rows:
                _fail(path, -100, f"{tbl}/{name}")
            rows = len(a)
        for name, kas in ragged:
            data, off = cols.get(name), cols.get(name + "_offset")
            if data is None and off is None:
                if (tbl, name) in _OPTIONAL_COLUMNS:
                    continue
                _fail(path, -103, f"{tbl}/{name}")
            if data is None or off is None:
                _fail(path, -104, f"{tbl}/{name}")
            data, off = np.asarray(data), np.asarray(off)
            if _KAS_OF.get(data.dtype) != kas or off.dtype not in (np.uint32, np.uint64):
                _fail(path, -105, f"{tbl}/{name}")
            if len(off) == 0 or (rows is not None and len(off) != rows + 1):
                _fail(path, -100, f"{tbl}/{name}_offset")
            rows = len(off) - 1
            if off[0] != 0 or int(off[-1]) != len(data) or np.any(off[1:] < off[:-1]):
                _fail(path, -200, f"{tbl}/{name}_offset")


def load_tables_dict(path):
    """The file as a tskit ``asdict()``-style dict; columns are views into the mapped file."""
    store = Store(path)
    store.native_tables()  # schema validation as the reference's loader does it
    d = {"encoding_version": (1, 6)}
    for key, arr in store.items():
        if "/" in key:
            tbl, col = key.split("/", 1)
            if tbl == "format":
                continue
            if col == "metadata_schema" or (tbl == "reference_sequence" and col in ("data", "url")):
                if tbl in _SCHEMA and _KAS_OF.get(arr.dtype) != U8:  # a table's schema is a uint8 column to tskit
                    _fail(path, -105, key)
                arr = _text(arr)
            elif tbl == "reference_sequence" and col == "metadata":
                arr = bytes(arr.view(np.uint8))
            d.setdefault(tbl, {})[col] = arr
        elif key == "sequence_length":
            d[key] = float(arr[0])
        elif key in ("time_units", "metadata_schema", "metadata", "uuid"):
            if arr.dtype != np.int8:  # kastore_gets_int8 (tables.c:11636-11700): KAS_ERR_TYPE_MISMATCH
                _check((-13 ^ (1 << 14)) ^ (1 << 13), f"Cannot load '{os.fspath(path)}' ({key})")
            if key == "metadata":
                d[key] = bytes(arr.view(np.uint8))
            else:
                d["file_uuid" if key == "uuid" else key] = _text(arr)
        else:
            d[key] = arr
    _check_columns(d, path)
    prov = d.pop("provenances", None)
    if prov is not None:
        try:
            ts = _tables.decode_ragged(prov["timestamp"], prov["timestamp_offset"])
            rec = _tables.decode_ragged(prov["record"], prov["record_offset"])
        except UnicodeDecodeError as e:
            raise FileFormatError(f"A provenance row is not valid UTF-8 ({e})", -100 ^ (1 << 13)) from None
        d["provenances"] = [{"timestamp": t, "record": r} for t, r in zip(ts, rec)]
    for tbl, (_, ragged) in _SCHEMA.items():
        for name, kas in ragged:
            if kas == U8 and isinstance(d.get(tbl), dict) and isinstance(d[tbl].get(name), np.ndarray):
                d[tbl][name] = d[tbl][name].view(np.int8)  # the dict protocol's dtype for state and metadata bytes
    return d


def load(path):
    """Reads a ``.trees`` file; the counterpart of ``tskit.load`` on this path."""
    d = load_tables_dict(path)
    if _tables.have_tskit():
        import tskit

        d2 = dict(d)
        d2.pop("file_uuid", None)
        prov = _provenance_columns(d2.pop("provenances", []))
        d2["provenances"] = {k: (v if k.endswith("_offset") else v.view(np.int8)) for k, v in prov.items()}
        return tskit.TableCollection.fromdict(d2).tree_sequence()
    return _tables.TableCollection.fromdict(d).tree_sequence()


def _offsets(off, n_rows):
    """The offset array as the reference writes it: uint32 unless the last entry needs 64 bits
    (``write_offset_col``, tables.c:362-401)."""
    off = np.asarray(off)
    if len(off) != n_rows + 1:
        raise ValueError("offset column length must be the number of rows + 1")
    if len(off) and int(off[-1]) > 0xFFFFFFFF:
        return np.ascontiguousarray(off, dtype=np.uint64)
    return np.ascontiguousarray(off, dtype=np.uint32)


def _bytes_array(v, kas):
    if isinstance(v, str):
        v = v.encode("utf8")
    if isinstance(v, (bytes, bytearray)):
        return np.frombuffer(bytes(v), dtype=np.uint8).view(_KAS_DTYPES[kas])
    a = np.ascontiguousarray(v)
    if a.dtype.itemsize == 1:
        return a.view(_KAS_DTYPES[kas])
    return a.astype(_KAS_DTYPES[kas])


def _provenance_columns(rows):
    ts, rec = [], []
    for r in rows:
        if isinstance(r, dict):
            ts.append(r["timestamp"])
            rec.append(r["record"])
        elif isinstance(r, (tuple, list)):
            ts.append(r[0])
            rec.append(r[1])
        else:  # a bare record (what sim_mutations appends): stamped now, as tskit's add_row does
            ts.append(datetime.datetime.now().isoformat())
            rec.append(r)
    t, to = _tables.encode_ragged(ts)
    r, ro = _tables.encode_ragged(rec)
    return {"timestamp": t, "timestamp_offset": to, "record": r, "record_offset": ro}


def file_items(d):
    """``{key: array}`` of a ``.trees`` file for the ``asdict()``-style dict ``d``: the key set
    and array types of ``tsk_table_collection_dump``."""
    items = {
        "format/name": np.frombuffer(FORMAT_NAME.encode(), dtype=np.int8),
        "format/version": np.array(FORMAT_VERSION, dtype=np.uint32),
        "sequence_length": np.array([d["sequence_length"]], dtype=np.float64),
        "uuid": np.frombuffer(str(_uuid.uuid4()).encode(), dtype=np.int8),
        "time_units": _bytes_array(d.get("time_units", "unknown"), I8),
        "metadata": _bytes_array(d.get("metadata", b""), I8),
        "metadata_schema": _bytes_array(d.get("metadata_schema", ""), I8),
    }
    for tbl, (cols, ragged) in _SCHEMA.items():
        src = d.get(tbl)
        if tbl == "provenances":
            src = _provenance_columns(src or [])
        src = src or {}
        n = None
        for name, _, _ in cols:
            if name in src and src[name] is not None:
                n = len(src[name])
                break
        if n is None:
            for name, _ in ragged:
                if src.get(name + "_offset") is not None:
                    n = len(src[name + "_offset"]) - 1
                    break
        n = 0 if n is None else n
        for name, kas, default in cols:
            v = src.get(name)
            if v is None:
                if default is None and n > 0:
                    raise ValueError(f"column {tbl}/{name} is required")
                v = np.full(n, 0 if default is None else default, dtype=_KAS_DTYPES[kas])
            a = np.ascontiguousarray(v, dtype=_KAS_DTYPES[kas])
            if len(a) != n:
                raise ValueError(f"column {tbl}/{name} has {len(a)} rows, expected {n}")
            items[f"{tbl}/{name}"] = a
        for name, kas in ragged:
            v, off = src.get(name), src.get(name + "_offset")
            if v is None or off is None:
                v, off = np.zeros(0, dtype=_KAS_DTYPES[kas]), np.zeros(n + 1, dtype=np.uint32)
            a = _bytes_array(v, kas) if kas == U8 else np.ascontiguousarray(v, dtype=_KAS_DTYPES[kas])
            off = _offsets(off, n)
            if int(off[-1]) != len(a):
                raise ValueError(f"column {tbl}/{name}: last offset {int(off[-1])} != data length {len(a)}")
            items[f"{tbl}/{name}"] = a
            items[f"{tbl}/{name}_offset"] = off
        if tbl != "provenances":
            items[f"{tbl}/metadata_schema"] = _bytes_array(src.get("metadata_schema", ""), U8)
    idx = d.get("indexes") or {}
    if idx.get("edge_insertion_order") is not None and idx.get("edge_removal_order") is not None:
        for name in ("edge_insertion_order", "edge_removal_order"):
            items[f"indexes/{name}"] = np.ascontiguousarray(idx[name], dtype=np.int32)
    ref = d.get("reference_sequence")
    if ref:
        for name in ("data", "url", "metadata", "metadata_schema"):
            items[f"reference_sequence/{name}"] = _bytes_array(ref.get(name, b""), U8)
    return items


def write_items(path, items):
    """Writes ``{key: numpy array}`` as one kastore file (``mspb_store_write``)."""
    n = len(items)
    arr = (_Item * max(n, 1))()
    hold = []
    for j, (key, a) in enumerate(items.items()):
        a = np.ascontiguousarray(a)
        k = key.encode()
        hold.append((k, a))
        arr[j].key = C.cast(C.c_char_p(k), C.c_void_p)
        arr[j].key_len = len(k)
        arr[j].type = _KAS_OF[a.dtype]
        arr[j].array = a.ctypes.data if a.size else None
        arr[j].array_len = a.size
    _check(lib().mspb_store_write(os.fspath(path).encode(), arr, n), f"Cannot write '{path}'")


def dump(tree_sequence, path):
    """Writes a tree sequence (tskit or stand-in) or an ``asdict()``-style dict as ``.trees``;
    the counterpart of ``TreeSequence.dump`` on this path."""
    if isinstance(tree_sequence, dict):
        d = tree_sequence
    else:
        d, kind = _tables.from_tree_sequence(tree_sequence)
        if kind == "tskit":
            prov = tree_sequence.tables.provenances
            d = dict(d)
            d["provenances"] = [{"timestamp": p.timestamp, "record": p.record} for p in prov]
    write_items(path, file_items(d))
