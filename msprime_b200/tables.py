"""Table hand-off for the host layer.

The reference moves tables between Python and C as the tskit ``asdict()`` protocol:
a dict of per-table dicts of numpy columns (``tskit_lwt_interface.h:1412-1580``;
SURVEY.md section 8 b6).  This module speaks the same protocol.  When ``tskit`` is
importable, ``sim_mutations`` accepts and returns real ``tskit.TreeSequence``
objects; otherwise (tskit is not installed in this image) the light stand-ins below
carry the same dict and expose the handful of attributes the mutation path uses.
"""
from __future__ import annotations

import numpy as np

TIME_UNITS_UNCALIBRATED = "uncalibrated"

try:  # pragma: no cover - depends on the environment
    import tskit as _tskit
except Exception:  # noqa: BLE001
    _tskit = None


def have_tskit() -> bool:
    return _tskit is not None


_RAGGED = {
    "sites": ("ancestral_state", "metadata"),
    "mutations": ("derived_state", "metadata"),
}


def _as_bytes(col) -> np.ndarray:
    a = np.asarray(col)
    if a.dtype.itemsize != 1:
        a = a.astype(np.int8)
    return np.ascontiguousarray(a).view(np.uint8)


class Table:
    """Column store for one table: attribute access to numpy columns."""

    def __init__(self, name, columns):
        self._name = name
        self._columns = dict(columns)

    def __getattr__(self, key):
        try:
            return self.__dict__["_columns"][key]
        except KeyError:
            raise AttributeError(key) from None

    @property
    def num_rows(self) -> int:
        first = {"nodes": "time", "edges": "left", "sites": "position", "mutations": "site"}[self._name]
        return len(self._columns[first])

    def __len__(self):
        return self.num_rows

    def asdict(self):
        return dict(self._columns)

    def clear(self):
        for k, v in list(self._columns.items()):
            if k.endswith("_offset"):
                self._columns[k] = np.zeros(1, dtype=v.dtype)
            elif isinstance(v, np.ndarray):
                self._columns[k] = v[:0]


class TableCollection:
    """Minimal stand-in for ``tskit.TableCollection`` (dict protocol only)."""

    def __init__(self, sequence_length=0.0):
        self.sequence_length = float(sequence_length)
        self.time_units = "unknown"
        self._tables = {}
        self._extra = {}
        self.provenances = []

    @classmethod
    def fromdict(cls, d) -> "TableCollection":
        self = cls(d["sequence_length"])
        self.time_units = d.get("time_units", "unknown")
        for name in ("nodes", "edges", "sites", "mutations"):
            cols = d.get(name)
            if cols is None:
                from . import synth

                cols = {"sites": synth.empty_sites, "mutations": synth.empty_mutations}[name]()
            self._tables[name] = Table(name, cols)
        self._extra = {
            k: v for k, v in d.items()
            if k not in ("sequence_length", "time_units", "nodes", "edges", "sites", "mutations", "provenances")
        }
        self.provenances = list(d.get("provenances", []))
        return self

    nodes = property(lambda self: self._tables["nodes"])
    edges = property(lambda self: self._tables["edges"])
    sites = property(lambda self: self._tables["sites"])
    mutations = property(lambda self: self._tables["mutations"])

    def asdict(self):
        d = {"sequence_length": self.sequence_length, "time_units": self.time_units}
        for name, t in self._tables.items():
            d[name] = t.asdict()
        d.update(self._extra)
        d["provenances"] = list(self.provenances)
        return d

    def copy(self) -> "TableCollection":
        return TableCollection.fromdict(
            {k: ({c: np.array(v, copy=True) for c, v in t.items()} if isinstance(t, dict) else t)
             for k, t in self.asdict().items()})

    def tree_sequence(self) -> "TreeSequence":
        return TreeSequence(self)


class TreeSequence:
    """Minimal stand-in for ``tskit.TreeSequence``: immutable view of a TableCollection."""

    def __init__(self, tables: TableCollection):
        self._tables = tables

    @property
    def tables(self) -> TableCollection:
        return self._tables

    def dump_tables(self) -> TableCollection:
        return self._tables.copy()

    sequence_length = property(lambda self: self._tables.sequence_length)
    time_units = property(lambda self: self._tables.time_units)
    num_nodes = property(lambda self: self._tables.nodes.num_rows)
    num_edges = property(lambda self: self._tables.edges.num_rows)
    num_sites = property(lambda self: self._tables.sites.num_rows)
    num_mutations = property(lambda self: self._tables.mutations.num_rows)
    num_provenances = property(lambda self: len(self._tables.provenances))

    @property
    def num_samples(self) -> int:
        return int(np.sum(np.asarray(self._tables.nodes.flags) & 1))

    def sites_position(self):
        return self._tables.sites.position

    def alleles(self):
        """Per-site ancestral state and per-mutation derived state as Python strings."""
        s, m = self._tables.sites, self._tables.mutations
        return (decode_ragged(s.ancestral_state, s.ancestral_state_offset),
                decode_ragged(m.derived_state, m.derived_state_offset))


def decode_ragged(data, offset):
    b = _as_bytes(data).tobytes()
    off = np.asarray(offset, dtype=np.int64)
    return [b[off[i]:off[i + 1]].decode("utf8") for i in range(len(off) - 1)]


def encode_ragged(strings):
    bs = [s.encode("utf8") if isinstance(s, str) else bytes(s) for s in strings]
    off = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        off[1:] = np.cumsum([len(b) for b in bs])
    return np.frombuffer(b"".join(bs), dtype=np.uint8).copy(), off


def from_tree_sequence(tree_sequence):
    """Returns ``(tables_dict, kind)`` for a tskit or stand-in tree sequence.

    Raises ValueError for anything without ``.tables`` (``msprime/mutations.py:1637-1640``).
    """
    try:
        tables = tree_sequence.tables
    except AttributeError:
        raise ValueError("First argument must be a TreeSequence instance.") from None
    kind = "tskit" if (_tskit is not None and isinstance(tree_sequence, _tskit.TreeSequence)) else "light"
    return tables.asdict(), kind


def to_tree_sequence(d, kind, provenance_record=None):
    """Builds the return value of ``sim_mutations`` from an ``asdict()``-style dict."""
    if kind == "tskit":
        tables = _tskit.TableCollection.fromdict(d)
        if provenance_record is not None:
            tables.provenances.add_row(provenance_record)
        return tables.tree_sequence()
    tables = TableCollection.fromdict(d)
    if provenance_record is not None:
        tables.provenances.append(provenance_record)
    return tables.tree_sequence()


def ragged_u64(tbl, name, n):
    """(uint8 data, uint64 offsets) for ragged column ``name`` of table dict ``tbl``."""
    if tbl is not None and name in tbl and (name + "_offset") in tbl and tbl[name + "_offset"] is not None:
        data = _as_bytes(tbl[name])
        off = np.ascontiguousarray(np.asarray(tbl[name + "_offset"], dtype=np.uint64))
    else:
        data = np.zeros(0, dtype=np.uint8)
        off = np.zeros(n + 1, dtype=np.uint64)
    return data, off
