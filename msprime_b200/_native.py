"""ctypes binding of the C ABI in ``include/msprime_b200.h`` (``_mspb.so``).

This is the thin extension layer the reference has in ``msprime/_msprimemodule.c``:
it turns numpy columns into plain pointers, calls the CUDA library and maps status
codes to the reference's exception classes (``_msprimemodule.c:80-96``):
``LibraryError`` for MSP_ERR_* / tskit codes, ``InputError`` for rate-map input
errors.  There is no CPU fallback: if the extension is missing or no GPU is usable
the call fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import tables as _tables

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "_mspb.so")

MODEL_MATRIX = 0
MODEL_INFINITE_ALLELES = 1
MODEL_SLIM = 2
NUM_PHASES = 8

ERR_NO_DEVICE = -1001
_RATE_MAP_ERRORS = {-43, -44, -45, -46, -47, -48}


class LibraryError(Exception):
    """Error raised by the native library (mirrors ``_msprime.LibraryError``)."""

    def __init__(self, message, status=None):
        super().__init__(message)
        self.status = status


class InputError(Exception):
    """Bad user input detected by the native library (mirrors ``_msprime.InputError``)."""

    def __init__(self, message, status=None):
        super().__init__(message)
        self.status = status


class _Tables(C.Structure):
    _fields_ = [
        ("sequence_length", C.c_double),
        ("num_nodes", C.c_uint64), ("node_time", C.c_void_p),
        ("num_edges", C.c_uint64), ("edge_left", C.c_void_p), ("edge_right", C.c_void_p),
        ("edge_parent", C.c_void_p), ("edge_child", C.c_void_p),
        ("num_sites", C.c_uint64), ("site_position", C.c_void_p),
        ("site_ancestral_state", C.c_void_p), ("site_ancestral_state_offset", C.c_void_p),
        ("site_metadata", C.c_void_p), ("site_metadata_offset", C.c_void_p),
        ("num_mutations", C.c_uint64), ("mutation_site", C.c_void_p), ("mutation_node", C.c_void_p),
        ("mutation_time", C.c_void_p),
        ("mutation_derived_state", C.c_void_p), ("mutation_derived_state_offset", C.c_void_p),
        ("mutation_metadata", C.c_void_p), ("mutation_metadata_offset", C.c_void_p),
    ]


class _Params(C.Structure):
    _fields_ = [
        ("map_size", C.c_uint64), ("map_position", C.c_void_p), ("map_rate", C.c_void_p),
        ("model_kind", C.c_int32), ("num_alleles", C.c_uint64), ("alleles", C.c_void_p),
        ("allele_offset", C.c_void_p), ("root_distribution", C.c_void_p),
        ("transition_matrix", C.c_void_p), ("start_allele", C.c_uint64),
        ("discrete_genome", C.c_int32), ("keep", C.c_int32),
        ("start_time", C.c_double), ("end_time", C.c_double),
        ("seed", C.c_uint64), ("replay_stream", C.c_void_p), ("replay_len", C.c_uint64),
        ("slice_left", C.c_double), ("slice_right", C.c_double),
        ("offsets32", C.c_int32), ("slice_stream", C.c_uint32),
        ("slim_mutation_type_id", C.c_int32), ("slim_generation", C.c_int32),
    ]


class _Counts(C.Structure):
    _fields_ = [
        ("num_sites", C.c_uint64), ("num_mutations", C.c_uint64), ("num_new_mutations", C.c_uint64),
        ("ancestral_state_bytes", C.c_uint64), ("site_metadata_bytes", C.c_uint64),
        ("derived_state_bytes", C.c_uint64), ("mutation_metadata_bytes", C.c_uint64),
        ("next_allele", C.c_uint64), ("draws", C.c_uint64), ("num_new_sites", C.c_uint64),
        ("device_ms", C.c_double), ("offset_bytes", C.c_uint64),
    ]


_COLUMNS = [
    ("site_position", np.float64), ("site_ancestral_state", np.uint8),
    ("site_ancestral_state_offset", np.uint64), ("site_metadata", np.uint8),
    ("site_metadata_offset", np.uint64), ("mutation_site", np.int32), ("mutation_node", np.int32),
    ("mutation_parent", np.int32), ("mutation_time", np.float64), ("mutation_derived_state", np.uint8),
    ("mutation_derived_state_offset", np.uint64), ("mutation_metadata", np.uint8),
    ("mutation_metadata_offset", np.uint64),
]


class _Shard(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("bytes_per_rank", C.c_uint64)]


NUM_TABLE_SHARDS = 5


class _Columns(C.Structure):
    _fields_ = [(name, C.c_void_p) for name, _ in _COLUMNS]


EXPORTED_SYMBOLS = [
    "mspb_ctx_create", "mspb_ctx_destroy", "mspb_ctx_load_tables", "mspb_generate", "mspb_fetch",
    "mspb_device_columns", "mspb_kernel_launches", "mspb_phase_ms", "mspb_phase_name",
    "mspb_sim_mutations", "mspb_oneshot_release", "mspb_columns_free", "mspb_strerror", "mspb_version", "mspb_cuda_runtime_version",
    "mspb_ctx_stream", "mspb_kernel_timing", "mspb_kernel_time", "mspb_selftest_sort",
    "mspb_ctx_force_placement", "mspb_generate_begin", "mspb_generate_begin_device", "mspb_generate_end", "mspb_shift_offsets",
    "mspb_exchange_buffers", "mspb_exchange_stream", "mspb_exchange_post_counts", "mspb_generate_end_device",
    "mspb_generate_finish_device", "mspb_last_fetch_bytes", "mspb_ctx_load_tables_begin",
    "mspb_ctx_load_tables_end", "mspb_site_rows_noalloc", "mspb_ctx_build_sample_index", "mspb_samples_below",
    "mspb_site_afs", "mspb_pinned_alloc", "mspb_pinned_free", "mspb_decode_genotypes", "mspb_pinned_stats",
]

EXCHANGE_WORDS = 8  # MSPB_EXCHANGE_WORDS
EXCHANGE_MAX_WORLD = 64  # MSPB_EXCHANGE_MAX_WORLD
EXCHANGE_FIELDS = ("num_sites", "num_mutations", "alleles_used", "num_new_mutations", "ancestral_state_bytes",
                   "site_metadata_bytes", "derived_state_bytes", "mutation_metadata_bytes")


class _DeviceArray:
    """A device buffer of this library seen through ``__cuda_array_interface__`` (zero copy)."""

    def __init__(self, ptr, n, typestr="<i8"):
        self.__cuda_array_interface__ = {
            "shape": (int(n),), "typestr": typestr, "data": (int(ptr), False), "version": 2, "strides": None}

_lib = None


def lib():
    """Loads ``_mspb.so``; raises RuntimeError (never falls back) when it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise RuntimeError(
                f"msprime_b200: CUDA extension {SO_PATH} is not built "
                "(run `python -m msprime_b200.build`); there is no CPU fallback")
        L = C.CDLL(SO_PATH)
        L.mspb_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.mspb_ctx_destroy.argtypes = [C.c_void_p]
        L.mspb_ctx_destroy.restype = None
        L.mspb_ctx_load_tables.argtypes = [C.c_void_p, C.POINTER(_Tables)]
        L.mspb_ctx_load_tables_begin.argtypes = [C.c_void_p, C.POINTER(_Tables), C.c_uint32, C.c_uint32,
                                                 C.POINTER(_Shard)]
        L.mspb_ctx_load_tables_end.argtypes = [C.c_void_p]
        L.mspb_generate.argtypes = [C.c_void_p, C.POINTER(_Params), C.POINTER(_Counts)]
        L.mspb_generate_begin.argtypes = [C.c_void_p, C.POINTER(_Params), C.POINTER(_Counts)]
        L.mspb_generate_begin_device.argtypes = [C.c_void_p, C.POINTER(_Params)]
        L.mspb_ctx_force_placement.argtypes = [C.c_void_p, C.c_int]
        L.mspb_exchange_stream.argtypes = [C.c_void_p]
        L.mspb_exchange_stream.restype = C.c_void_p
        L.mspb_generate_end.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(_Counts)]
        L.mspb_shift_offsets.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64]
        L.mspb_exchange_buffers.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.mspb_exchange_post_counts.argtypes = [C.c_void_p]
        L.mspb_generate_end_device.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        L.mspb_generate_finish_device.argtypes = [
            C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(_Counts), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.mspb_fetch.argtypes = [C.c_void_p, C.POINTER(_Columns)]
        L.mspb_device_columns.argtypes = [C.c_void_p, C.POINTER(_Columns)]
        L.mspb_kernel_launches.argtypes = [C.c_void_p]
        L.mspb_kernel_launches.restype = C.c_uint64
        L.mspb_last_fetch_bytes.argtypes = [C.c_void_p]
        L.mspb_last_fetch_bytes.restype = C.c_uint64
        L.mspb_phase_ms.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.c_int]
        L.mspb_phase_name.argtypes = [C.c_int]
        L.mspb_phase_name.restype = C.c_char_p
        L.mspb_strerror.argtypes = [C.c_int]
        L.mspb_strerror.restype = C.c_char_p
        L.mspb_version.restype = C.c_char_p
        L.mspb_cuda_runtime_version.restype = C.c_int
        L.mspb_selftest_sort.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_uint64, C.c_uint64, C.c_int,
                                         C.POINTER(C.c_int)]
        L.mspb_ctx_stream.argtypes = [C.c_void_p]
        L.mspb_ctx_stream.restype = C.c_void_p
        L.mspb_kernel_timing.argtypes = [C.c_void_p, C.c_int]
        L.mspb_kernel_time.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_uint64),
                                       C.POINTER(C.c_double)]
        L.mspb_sim_mutations.argtypes = [C.c_int, C.POINTER(_Tables), C.POINTER(_Params), C.POINTER(_Counts),
                                         C.POINTER(_Columns)]
        L.mspb_columns_free.argtypes = [C.POINTER(_Columns)]
        L.mspb_columns_free.restype = None
        L.mspb_oneshot_release.argtypes = [C.c_int]
        L.mspb_oneshot_release.restype = None
        L.mspb_site_rows_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
        L.mspb_ctx_build_sample_index.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                                  C.POINTER(C.c_uint64), C.POINTER(C.c_double)]
        L.mspb_samples_below.argtypes = [C.c_void_p, C.c_void_p]
        L.mspb_site_afs.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_double)]
        L.mspb_decode_genotypes.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.POINTER(C.c_double)]
        L.mspb_pinned_alloc.argtypes = [C.c_uint64]
        L.mspb_pinned_alloc.restype = C.c_void_p
        L.mspb_pinned_free.argtypes = [C.c_void_p]
        L.mspb_pinned_free.restype = None
        _lib = L
    return _lib


def strerror(status: int) -> str:
    return lib().mspb_strerror(int(status)).decode()


ERR_RETRY_WITH_HOST = -1011  # MSPB_ERR_RETRY_WITH_HOST


class RetryWithHost(LibraryError):
    """``generate_finish_device`` after ``generate_begin_device``: some rank needs the host between the stages
    (every rank raises this together); make the call again with ``generate_begin``."""


def check(status: int):
    if status == 0:
        return
    msg = strerror(status)
    if status == ERR_RETRY_WITH_HOST:
        raise RetryWithHost(msg, status)
    if status in _RATE_MAP_ERRORS:
        raise InputError(f"Input error in mutation rate map: {msg}", status)
    raise LibraryError(msg, status)


def _c(a, dtype):
    return np.ascontiguousarray(np.asarray(a, dtype=dtype))


def _p(a):
    return a.ctypes.data


class _PinnedBlock:
    """One block of the library's pinned-memory pool; goes back to the pool when collected."""

    def __init__(self, nbytes):
        self.ptr = lib().mspb_pinned_alloc(int(nbytes))

    def __del__(self):
        try:
            if self.ptr:
                lib().mspb_pinned_free(self.ptr)
                self.ptr = None
        except Exception:  # noqa: BLE001  (interpreter shutdown)
            pass


def pinned_empty(n, dtype):
    """A numpy array of ``n`` items over pinned host memory from the library's pool (the block returns to the
    pool when the array and its views are collected); a plain ``np.empty`` when pinning fails."""
    dtype = np.dtype(dtype)
    nbytes = int(n) * dtype.itemsize
    if nbytes == 0:
        return np.empty(0, dtype=dtype)
    block = _PinnedBlock(nbytes)
    if not block.ptr:
        return np.empty(int(n), dtype=dtype)
    buf = (C.c_char * nbytes).from_address(block.ptr)
    buf._block = block  # the array below keeps `buf` alive, and `buf` the block
    return np.frombuffer(buf, dtype=dtype, count=int(n))


class Params:
    """Keeps the numpy arrays behind a ``mspb_params_t`` alive."""

    def __init__(self, *, map_position, map_rate, model, discrete_genome, keep, start_time, end_time,
                 seed=0, replay_stream=None, slice_left=0.0, slice_right=0.0, offsets32=False,
                 slice_stream=0):
        self._hold = []
        p = _Params()
        pos, rate = _c(map_position, np.float64), _c(map_rate, np.float64)
        if pos.ndim != 1 or rate.ndim != 1 or len(pos) != len(rate) + 1:
            # _msprimemodule.c:155-204 (parse_rate_map)
            raise ValueError("The position array must be one larger than the rate array")
        self._hold += [pos, rate]
        p.map_size, p.map_position, p.map_rate = len(rate), _p(pos), _p(rate)
        p.model_kind = model["kind"]
        if model["kind"] == MODEL_MATRIX:
            data, off = _tables.encode_ragged(model["alleles"])
            root = _c(model["root_distribution"], np.float64)
            mat = _c(model["transition_matrix"], np.float64)
            self._hold += [data, off, root, mat]
            p.num_alleles = len(model["alleles"])
            p.alleles, p.allele_offset = _p(data), _p(off)
            p.root_distribution, p.transition_matrix = _p(root), _p(mat)
        elif model["kind"] == MODEL_SLIM:
            # {"kind": 2, "type": mutation type id, "next_id": ..., "slim_generation": ...}
            p.start_allele = int(model.get("next_id", 0)) & 0xFFFFFFFFFFFFFFFF
            p.slim_mutation_type_id = int(model["type"])
            p.slim_generation = int(model.get("slim_generation", 1))
        else:
            p.start_allele = int(model["start_allele"]) & 0xFFFFFFFFFFFFFFFF
        p.discrete_genome, p.keep = int(bool(discrete_genome)), int(bool(keep))
        p.start_time, p.end_time = float(start_time), float(end_time)
        p.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        if replay_stream is not None:
            st = _c(replay_stream, np.float64)
            self._hold.append(st)
            p.replay_stream, p.replay_len = _p(st), len(st)
        p.slice_left, p.slice_right = float(slice_left), float(slice_right)
        p.offsets32 = int(bool(offsets32))
        p.slice_stream = int(slice_stream)
        self.struct = p


class Context:
    """One GPU context: upload a tree sequence once, overlay mutations many times."""

    def __init__(self, device: int = 0):
        self._h = C.c_void_p()
        check(lib().mspb_ctx_create(int(device), C.byref(self._h)))
        self.device = int(device)
        self.counts = None

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().mspb_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001  (interpreter shutdown: module globals may be gone)
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @staticmethod
    def _tables_struct(d):
        """``mspb_tables_t`` for the ``asdict()``-style dict ``d`` plus the arrays it points into."""
        hold = []
        t = _Tables()
        t.sequence_length = float(d["sequence_length"])
        nt = _c(d["nodes"]["time"], np.float64)
        e = d["edges"]
        el, er = _c(e["left"], np.float64), _c(e["right"], np.float64)
        ep, ec = _c(e["parent"], np.int32), _c(e["child"], np.int32)
        hold += [nt, el, er, ep, ec]
        t.num_nodes, t.node_time = len(nt), _p(nt)
        t.num_edges = len(el)
        t.edge_left, t.edge_right, t.edge_parent, t.edge_child = _p(el), _p(er), _p(ep), _p(ec)
        s = d.get("sites")
        ns = 0 if s is None else len(s["position"])
        t.num_sites = ns
        if ns > 0:
            sp = _c(s["position"], np.float64)
            sa, sao = _tables.ragged_u64(s, "ancestral_state", ns)
            sm, smo = _tables.ragged_u64(s, "metadata", ns)
            hold += [sp, sa, sao, sm, smo]
            t.site_position = _p(sp)
            t.site_ancestral_state, t.site_ancestral_state_offset = _p(sa), _p(sao)
            t.site_metadata, t.site_metadata_offset = _p(sm), _p(smo)
        m = d.get("mutations")
        nm = 0 if m is None else len(m["site"])
        t.num_mutations = nm
        if nm > 0:
            ms, mn = _c(m["site"], np.int32), _c(m["node"], np.int32)
            md, mdo = _tables.ragged_u64(m, "derived_state", nm)
            mm, mmo = _tables.ragged_u64(m, "metadata", nm)
            hold += [ms, mn, md, mdo, mm, mmo]
            t.mutation_site, t.mutation_node = _p(ms), _p(mn)
            if m.get("time") is not None:
                mt = _c(m["time"], np.float64)
                hold.append(mt)
                t.mutation_time = _p(mt)
            t.mutation_derived_state, t.mutation_derived_state_offset = _p(md), _p(mdo)
            t.mutation_metadata, t.mutation_metadata_offset = _p(mm), _p(mmo)
        return t, hold

    def load_tables(self, d):
        """``d`` is a tskit ``asdict()``-style dict of numpy columns."""
        t, hold = self._tables_struct(d)
        check(lib().mspb_ctx_load_tables(self._h, C.byref(t)))
        del hold

    def load_tables_begin(self, d, rank, world):
        """Sharded upload (``mspb_ctx_load_tables_begin``): copies this rank's share of the large
        columns and returns ``[(device pointer, bytes per rank)] * 5``: the buffers the caller must
        all-gather in place on this context's stream before :meth:`load_tables_end`."""
        t, hold = self._tables_struct(d)
        shards = (_Shard * NUM_TABLE_SHARDS)()
        check(lib().mspb_ctx_load_tables_begin(self._h, C.byref(t), int(rank), int(world), shards))
        del hold  # the call returns after its copies have completed
        return [(int(sh.ptr or 0), int(sh.bytes_per_rank)) for sh in shards]

    def load_tables_end(self):
        check(lib().mspb_ctx_load_tables_end(self._h))

    def load_native(self, t):
        """``t`` is an ``mspb_tables_t`` already filled in, e.g. by ``trees_io.Store.native_tables``
        (columns of a mapped ``.trees`` file go to the device without passing through a dict)."""
        check(lib().mspb_ctx_load_tables(self._h, C.byref(t)))

    def generate(self, params: Params):
        c = _Counts()
        check(lib().mspb_generate(self._h, C.byref(params.struct), C.byref(c)))
        self.counts = {name: getattr(c, name) for name, _ in _Counts._fields_}
        return self.counts

    def generate_begin(self, params: Params):
        """First phase of a genome-sharded call: counts of this slice (no tables yet)."""
        c = _Counts()
        check(lib().mspb_generate_begin(self._h, C.byref(params.struct), C.byref(c)))
        self.counts = {name: getattr(c, name) for name, _ in _Counts._fields_}
        return self.counts

    def generate_begin_device(self, params: Params):
        """``generate_begin`` without its wait (``mspb_generate_begin_device``): the counts stay on the device
        until ``generate_finish_device``, which may raise :class:`RetryWithHost`."""
        check(lib().mspb_generate_begin_device(self._h, C.byref(params.struct)))
        self.counts = None

    def force_k2(self, mode):
        """0: placement kernel chosen from the density; 1: one block per 256 edges; 2: one thread per mutation."""
        check(lib().mspb_ctx_force_placement(self._h, int(mode)))

    def exchange_stream(self):
        """The ``cudaStream_t`` (integer) the next all-gather of the exchange buffers must be enqueued on."""
        return int(lib().mspb_exchange_stream(self._h) or 0)

    def generate_end(self, site_base=0, mutation_base=0, allele_base=0):
        """Second phase: writes the tables with global site / mutation / allele numbering."""
        c = _Counts()
        check(lib().mspb_generate_end(self._h, int(site_base), int(mutation_base),
                                      int(allele_base) & 0xFFFFFFFFFFFFFFFF, C.byref(c)))
        self.counts = {name: getattr(c, name) for name, _ in _Counts._fields_}
        return self.counts

    def exchange_tensors(self, world):
        """(send, recv) of the device-side exchange as int64 torch tensors on this context's
        device: ``send`` has EXCHANGE_WORDS entries, ``recv`` has ``world`` times as many."""
        import torch

        if not 0 < world <= EXCHANGE_MAX_WORLD:
            raise ValueError("world size out of range")
        send, recv = C.c_void_p(), C.c_void_p()
        check(lib().mspb_exchange_buffers(self._h, C.byref(send), C.byref(recv)))
        dev = torch.device("cuda", self.device)
        return (torch.as_tensor(_DeviceArray(send.value, EXCHANGE_WORDS), device=dev),
                torch.as_tensor(_DeviceArray(recv.value, EXCHANGE_WORDS * world), device=dev))

    def exchange_post_counts(self):
        check(lib().mspb_exchange_post_counts(self._h))

    def generate_end_device(self, rank, world):
        """Second phase with the bases taken from the gathered counts on the device (no wait)."""
        check(lib().mspb_generate_end_device(self._h, int(rank), int(world)))

    def generate_finish_device(self, rank, world):
        """Shifts the byte offsets by the gathered byte totals and waits.  Returns
        ``(counts, totals, bases)``; the last two are dicts over EXCHANGE_FIELDS."""
        c = _Counts()
        tot = (C.c_uint64 * EXCHANGE_WORDS)()
        bas = (C.c_uint64 * EXCHANGE_WORDS)()
        check(lib().mspb_generate_finish_device(self._h, int(rank), int(world), C.byref(c), tot, bas))
        self.counts = {name: getattr(c, name) for name, _ in _Counts._fields_}
        return (self.counts, dict(zip(EXCHANGE_FIELDS, (int(v) for v in tot))),
                dict(zip(EXCHANGE_FIELDS, (int(v) for v in bas))))

    def shift_offsets(self, ancestral_state=0, site_metadata=0, derived_state=0, mutation_metadata=0):
        check(lib().mspb_shift_offsets(self._h, int(ancestral_state), int(site_metadata), int(derived_state),
                                       int(mutation_metadata)))

    def result_sizes(self):
        """Element counts of the output columns of the last ``generate``."""
        c = self.counts
        S, M = c["num_sites"], c["num_mutations"]
        return {
            "site_position": S, "site_ancestral_state": c["ancestral_state_bytes"],
            "site_ancestral_state_offset": S + 1, "site_metadata": c["site_metadata_bytes"],
            "site_metadata_offset": S + 1, "mutation_site": M, "mutation_node": M, "mutation_parent": M,
            "mutation_time": M, "mutation_derived_state": c["derived_state_bytes"],
            "mutation_derived_state_offset": M + 1, "mutation_metadata": c["mutation_metadata_bytes"],
            "mutation_metadata_offset": M + 1,
        }

    def fetch(self, out=None):
        """Copies the resident result to the host: ``{"sites": {...}, "mutations": {...}}``.

        ``out`` may map column names (``_COLUMNS``) to preallocated host arrays (pinned
        memory makes the copy a single DMA); each must be at least as long as the
        column.  Without it fresh numpy arrays are allocated."""
        sizes = self.result_sizes()
        cols = _Columns()
        arrays = {}
        off_dtype = np.uint32 if self.counts["offset_bytes"] == 4 else np.uint64
        for name, dtype in _COLUMNS:
            if name.endswith("_offset"):
                dtype = off_dtype
            if out is not None:
                buf = out[name]
                if buf.dtype != dtype or buf.size < sizes[name]:
                    raise ValueError(f"output buffer for {name} is too small or of the wrong dtype")
                a = buf[:sizes[name]]
            else:
                # fresh arrays over pinned memory from the library's pool: the DMA runs at the link's rate
                # (pageable destinations are staged by the driver at a fraction of it)
                a = pinned_empty(sizes[name], dtype)
            arrays[name] = a
            setattr(cols, name, _p(a) if a.size > 0 else None)
        check(lib().mspb_fetch(self._h, C.byref(cols)))
        sites = {k[len("site_"):]: v for k, v in arrays.items() if k.startswith("site_")}
        muts = {k[len("mutation_"):]: v for k, v in arrays.items() if k.startswith("mutation_")}
        return {"sites": sites, "mutations": muts}

    def last_fetch_bytes(self) -> int:
        """Device-to-host bytes of the last :meth:`fetch`."""
        return int(lib().mspb_last_fetch_bytes(self._h))

    def device_columns(self):
        """Raw device pointers (ints) of the resident result, with element counts and dtypes."""
        cols = _Columns()
        check(lib().mspb_device_columns(self._h, C.byref(cols)))
        return {name: getattr(cols, name) for name, _ in _COLUMNS}

    def device_tensors(self):
        """The resident result as torch tensors on this context's device, WITHOUT a copy
        (the table hand-off for consumers that stay on the GPU):
        ``{"sites": {...}, "mutations": {...}}`` with the column names of :meth:`fetch`.
        The tensors alias the context's buffers: they are valid until the next ``generate*``
        call on this context and are written on its stream (:meth:`stream`).  torch has no
        unsigned 32/64-bit arithmetic, so offset columns come back as int32 / int64 views
        and byte columns as uint8."""
        import torch

        ptrs = self.device_columns()
        sizes = self.result_sizes()
        dev = torch.device("cuda", self.device)
        off = "<i4" if self.counts["offset_bytes"] == 4 else "<i8"
        typestr = {np.float64: "<f8", np.int32: "<i4", np.uint8: "|u1", np.uint64: off}
        res = {"sites": {}, "mutations": {}}
        for name, dtype in _COLUMNS:
            tbl, col = ("sites", name[len("site_"):]) if name.startswith("site_") else (
                "mutations", name[len("mutation_"):])
            n = sizes[name]
            if n == 0 or not ptrs[name]:
                tdt = {"<f8": torch.float64, "<i4": torch.int32, "|u1": torch.uint8, "<i8": torch.int64}
                res[tbl][col] = torch.empty(0, dtype=tdt[typestr[dtype]], device=dev)
            else:
                res[tbl][col] = torch.as_tensor(_DeviceArray(ptrs[name], n, typestr[dtype]), device=dev)
        return res

    # ---- consumers of the resident result
    def build_sample_index(self, node_flags):
        """Builds, for the loaded tree sequence, the step functions below(u, x) = number of samples under
        node ``u`` in the tree covering ``x`` (``node_flags``: tskit's ``nodes.flags``, bit 0 = sample).
        Once per ``load_tables``; returns sizes and the device time of the build."""
        flags = _c(node_flags, np.uint32)
        ns, nl, nst, ms = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_double()
        check(lib().mspb_ctx_build_sample_index(self._h, C.c_void_p(_p(flags) if flags.size else None), C.byref(ns),
                                                C.byref(nl), C.byref(nst), C.byref(ms)))
        self.sample_index = {"num_samples": int(ns.value), "levels": int(nl.value), "steps": int(nst.value),
                             "build_ms": float(ms.value)}
        return self.sample_index

    def samples_below(self):
        """Samples below every mutation row of the resident result (``tsk_tree_get_num_samples`` of the
        mutation's node in the tree covering its site)."""
        out = np.empty(self.counts["num_mutations"], dtype=np.uint32)
        check(lib().mspb_samples_below(self._h, C.c_void_p(_p(out) if out.size else None) if out.size else _p(np.empty(1, np.uint32))))
        return out

    def site_afs(self):
        """Polarised, un-normalised site allele frequency spectrum of all samples of the resident result
        (``tsk_treeseq_allele_frequency_spectrum``, mode "site", polarised, span_normalise=False): a
        uint64 array of ``num_samples + 1`` counts.  ``self.afs_ms`` is the device time of the query."""
        n = self.sample_index["num_samples"]
        out = np.zeros(n + 1, dtype=np.uint64)
        ms = C.c_double()
        check(lib().mspb_site_afs(self._h, _p(out), n + 1, C.byref(ms)))
        self.afs_ms = float(ms.value)
        return out

    def genotypes(self, site_begin=0, site_end=None):
        """``tsk_variant_decode`` for the sites ``[site_begin, site_end)`` of the resident result: an int32 array of
        shape (sites, samples), sample nodes in increasing order, allele 0 = the ancestral state."""
        S = self.counts["num_sites"]
        site_end = S if site_end is None else int(site_end)
        n = self.sample_index["num_samples"]
        out = np.zeros((max(site_end - int(site_begin), 0), n), dtype=np.int32)
        ms = C.c_double()
        if out.size > 0:
            check(lib().mspb_decode_genotypes(self._h, int(site_begin), site_end, _p(out), C.byref(ms)))
        self.decode_ms = float(ms.value)
        return out

    def kernel_launches(self) -> int:
        return int(lib().mspb_kernel_launches(self._h))

    def selftest_sort(self, n, key_bytes=8, key_mask=0xFFFFFFFFFFFFFFFF, seed=1, measure_bits=True):
        """Runs the device radix sort on random keys and verifies it; returns the pass count."""
        passes = C.c_int(0)
        check(lib().mspb_selftest_sort(self._h, int(n), int(key_bytes), int(key_mask) & 0xFFFFFFFFFFFFFFFF,
                                       int(seed), int(bool(measure_bits)), C.byref(passes)))
        return passes.value

    def stream(self) -> int:
        """The context's cudaStream_t as an integer (wrap with ``torch.cuda.ExternalStream``)."""
        return int(lib().mspb_ctx_stream(self._h) or 0)

    def kernel_timing(self, enable=True):
        check(lib().mspb_kernel_timing(self._h, int(bool(enable))))

    def kernel_times(self):
        """``{kernel name: (launches, total ms)}`` accumulated since ``kernel_timing(True)``."""
        out = {}
        i = 0
        name, n, ms = C.c_char_p(), C.c_uint64(), C.c_double()
        while lib().mspb_kernel_time(self._h, i, C.byref(name), C.byref(n), C.byref(ms)) == 0:
            out[name.value.decode()] = (int(n.value), float(ms.value))
            i += 1
        return out

    def phase_ms(self):
        buf = (C.c_float * NUM_PHASES)()
        check(lib().mspb_phase_ms(self._h, buf, NUM_PHASES))
        return {lib().mspb_phase_name(i).decode(): float(buf[i]) for i in range(NUM_PHASES)}


def sim_mutations_oneshot(tables_dict, params: Params, device: int = 0):
    """The one-shot entry point ``mspb_sim_mutations`` (the call INTEGRATION.md's stub binds where the
    reference calls ``mutgen_alloc`` .. ``mutgen_generate``, ``msprime/_msprimemodule.c:3084-3109``):
    host tables in, malloc'd host columns out.  Returns ``(result, counts)`` with the columns copied
    into numpy arrays and the library's buffers released (``mspb_columns_free``)."""
    t, hold = Context._tables_struct(tables_dict)
    c, cols = _Counts(), _Columns()
    check(lib().mspb_sim_mutations(int(device), C.byref(t), C.byref(params.struct), C.byref(c), C.byref(cols)))
    del hold
    try:
        counts = {name: getattr(c, name) for name, _ in _Counts._fields_}
        S, M = counts["num_sites"], counts["num_mutations"]
        sizes = {
            "site_position": S, "site_ancestral_state": counts["ancestral_state_bytes"],
            "site_ancestral_state_offset": S + 1, "site_metadata": counts["site_metadata_bytes"],
            "site_metadata_offset": S + 1, "mutation_site": M, "mutation_node": M, "mutation_parent": M,
            "mutation_time": M, "mutation_derived_state": counts["derived_state_bytes"],
            "mutation_derived_state_offset": M + 1, "mutation_metadata": counts["mutation_metadata_bytes"],
            "mutation_metadata_offset": M + 1,
        }
        off_dtype = np.uint32 if counts["offset_bytes"] == 4 else np.uint64
        arrays = {}
        for name, dtype in _COLUMNS:
            if name.endswith("_offset"):
                dtype = off_dtype
            n = int(sizes[name])
            ptr = getattr(cols, name)
            if n == 0 or not ptr:
                arrays[name] = np.zeros(0, dtype=dtype)
            else:
                arrays[name] = np.frombuffer(C.string_at(ptr, n * np.dtype(dtype).itemsize), dtype=dtype).copy()
    finally:
        lib().mspb_columns_free(C.byref(cols))
    sites = {k[len("site_"):]: v for k, v in arrays.items() if k.startswith("site_")}
    muts = {k[len("mutation_"):]: v for k, v in arrays.items() if k.startswith("mutation_")}
    return {"sites": sites, "mutations": muts}, counts


NOALLOC_MAX_RECORDS = 1 << 26  # MSPB_NOALLOC_MAX_RECORDS
ERR_NOALLOC_TOO_LARGE = -1010


def site_rows_noalloc(device, stream, sorted_positions_ptr, n, site_row_ptr, num_sites_ptr):
    """``mspb_site_rows_noalloc``: raw device pointers in, work enqueued on ``stream`` (an integer
    ``cudaStream_t``, 0 = default stream); nothing is allocated.  Raises on a non-zero status."""
    check(lib().mspb_site_rows_noalloc(int(device), C.c_void_p(int(stream) or None),
                                       C.c_void_p(int(sorted_positions_ptr) or None), int(n),
                                       C.c_void_p(int(site_row_ptr) or None), C.c_void_p(int(num_sites_ptr) or None)))


def pinned_stats():
    """(blocks pinned, blocks reused, bytes pinned so far, blocks unpinned) of the pinned-memory pool."""
    out = (C.c_uint64 * 4)()
    lib().mspb_pinned_stats(out)
    return tuple(int(v) for v in out)
