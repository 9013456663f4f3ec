"""Synthetic tree-sequence inputs (benchmark / test infrastructure).

The reference generates its benchmark inputs with ``msprime.sim_ancestry``, which
needs GSL and the whole Hudson simulator (``lib/msprime.c``) -- neither exists in
this image nor on the GPU box.  This module builds
coalescent-with-recombination tree sequences itself, with the SMC' left-to-right
construction: a Kingman tree at position 0, then for each recombination breakpoint
one subtree-prune-and-regraft move whose prune point is uniform on the tree's
branch length and whose regraft time follows the coalescent hazard of the
remaining lineages.  The output obeys tskit's table requirements (edges sorted by
parent time, parent, child, left; parent older than child).

``sim_ancestry_smc`` here is the small-input, numpy-vectorised O(n log n)-per-event
version used by the tests; the large benchmark inputs come from the C
implementation of the same construction (``csrc/treegen.c``, ``native_treeseq``).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_TREEGEN_SO = os.path.join(_HERE, "_treegen.so")


class _TreegenResult(C.Structure):
    _fields_ = [
        ("num_nodes", C.c_uint64), ("num_samples", C.c_uint64), ("num_edges", C.c_uint64),
        ("num_trees", C.c_uint64), ("node_time", C.POINTER(C.c_double)),
        ("edge_left", C.POINTER(C.c_double)), ("edge_right", C.POINTER(C.c_double)),
        ("edge_parent", C.POINTER(C.c_int32)), ("edge_child", C.POINTER(C.c_int32)),
    ]


_treegen = None


def _treegen_lib():
    global _treegen
    if _treegen is None:
        if not os.path.exists(_TREEGEN_SO):
            from . import build

            build.build_treegen()
        lib = C.CDLL(_TREEGEN_SO)
        lib.treegen_smc.argtypes = [C.c_uint64, C.c_double, C.c_double, C.c_double, C.c_uint64, C.c_int,
                                    C.POINTER(_TreegenResult)]
        lib.treegen_smc.restype = C.c_int
        lib.treegen_free.argtypes = [C.POINTER(_TreegenResult)]
        lib.treegen_free.restype = None
        _treegen = lib
    return _treegen


def native_treeseq(samples, sequence_length, recombination_rate, population_size, random_seed, ploidy=2,
                   discrete_genome=True):
    """Large SMC' tree sequence from the C generator (``csrc/treegen.c``); same arguments
    and output layout as :func:`sim_ancestry_smc`.  Adds ``num_trees`` to the dict."""
    lib = _treegen_lib()
    res = _TreegenResult()
    n = int(samples) * int(ploidy)
    ret = lib.treegen_smc(n, float(sequence_length), float(recombination_rate),
                          float(ploidy) * float(population_size), int(random_seed), int(bool(discrete_genome)),
                          C.byref(res))
    if ret != 0:
        raise RuntimeError(f"treegen_smc failed with status {ret}")
    try:
        N, E = res.num_nodes, res.num_edges
        flags = np.zeros(N, dtype=np.uint32)
        flags[:n] = 1

        def arr(ptr, count, dtype):
            if count == 0:
                return np.zeros(0, dtype=dtype)
            return np.ctypeslib.as_array(ptr, shape=(count,)).astype(dtype, copy=True)

        d = {
            "sequence_length": float(sequence_length),
            "time_units": "generations",
            "nodes": {"time": arr(res.node_time, N, np.float64), "flags": flags},
            "edges": {
                "left": arr(res.edge_left, E, np.float64), "right": arr(res.edge_right, E, np.float64),
                "parent": arr(res.edge_parent, E, np.int32), "child": arr(res.edge_child, E, np.int32),
            },
            "sites": empty_sites(),
            "mutations": empty_mutations(),
            "num_trees": int(res.num_trees),
        }
    finally:
        lib.treegen_free(C.byref(res))
    return d


def _kingman(n, two_ne, rng):
    """Random binary coalescent tree over leaves 0..n-1; returns (parent, time)."""
    parent = np.full(2 * n - 1, -1, dtype=np.int64)
    time = np.zeros(2 * n - 1, dtype=np.float64)
    live = list(range(n))
    t = 0.0
    nxt = n
    for k in range(n, 1, -1):
        t += rng.exponential(two_ne / (k * (k - 1) / 2.0))
        i, j = rng.choice(k, size=2, replace=False)
        a, b = live[i], live[j]
        parent[a] = nxt
        parent[b] = nxt
        time[nxt] = t
        live = [x for idx, x in enumerate(live) if idx not in (i, j)]
        live.append(nxt)
        nxt += 1
    return parent, time


def sim_ancestry_smc(
    samples,
    sequence_length,
    recombination_rate,
    population_size,
    random_seed,
    ploidy=2,
    discrete_genome=True,
):
    """SMC' tree sequence; returns the tskit ``asdict()``-style dict of numpy columns.

    Mirrors the argument meaning of ``msprime.sim_ancestry`` (``msprime/ancestry.py:963``):
    ``samples`` diploid individuals -> ``ploidy * samples`` sample nodes at time 0.
    """
    rng = np.random.default_rng(random_seed)
    n = int(samples) * ploidy
    L = float(sequence_length)
    two_ne = ploidy * float(population_size)
    parent0, time0 = _kingman(n, two_ne, rng)
    cap = 4 * n
    parent = np.full(cap, -1, dtype=np.int64)
    time = np.zeros(cap, dtype=np.float64)
    edge_left = np.zeros(cap, dtype=np.float64)  # left end of the open edge above each node
    active = np.zeros(cap, dtype=bool)
    child = np.full((cap, 2), -1, dtype=np.int64)
    size = 2 * n - 1
    parent[:size] = parent0
    time[:size] = time0
    active[:size] = True
    for v in range(size):
        if parent0[v] >= 0:
            child[parent0[v], 0 if child[parent0[v], 0] < 0 else 1] = v
    out_l, out_r, out_p, out_c = [], [], [], []

    def close_edge(c, x):
        if parent[c] >= 0 and edge_left[c] < x:
            out_l.append(edge_left[c])
            out_r.append(x)
            out_p.append(parent[c])
            out_c.append(c)

    def replace_child(p, old, new):
        if p >= 0:
            child[p, 0 if child[p, 0] == old else 1] = new

    x = 0.0
    while True:
        idx = np.flatnonzero(active[:size] & (parent[:size] >= 0))
        blen = time[parent[idx]] - time[idx]
        total = blen.sum()
        if recombination_rate <= 0 or total <= 0:
            break
        x += rng.exponential(1.0 / (recombination_rate * total))
        bp = np.ceil(x) if discrete_genome else x
        if bp >= L:
            break
        # prune point: uniform on branch length
        pick = int(np.searchsorted(np.cumsum(blen), rng.random() * total, side="right"))
        u = int(idx[min(pick, idx.size - 1)])
        p = int(parent[u])
        t = time[u] + rng.random() * (time[p] - time[u])
        # nodes in the subtree of u (excluded from regraft targets)
        in_sub = np.zeros(size, dtype=bool)
        stack = [u]
        while stack:
            v = stack.pop()
            in_sub[v] = True
            for w in child[v]:
                if w >= 0:
                    stack.append(int(w))
        s = int(child[p, 1] if child[p, 0] == u else child[p, 0])
        g = int(parent[p])
        # detach: remove p, splice s onto g
        close_edge(u, bp)
        close_edge(s, bp)
        close_edge(p, bp)
        parent[s] = g
        replace_child(g, p, s)
        edge_left[s] = bp
        active[p] = False
        parent[u] = -1
        parent[p] = -1
        child[p] = -1
        # lineages available to the floating lineage
        cand = np.flatnonzero(active[:size] & ~in_sub)
        tc = time[cand]
        tp = np.where(parent[cand] >= 0, time[np.maximum(parent[cand], 0)], np.inf)
        bounds = np.unique(np.concatenate([[t], tc[tc > t], tp[(tp > t) & np.isfinite(tp)]]))
        tc_s, tp_s = np.sort(tc), np.sort(tp)
        kk = np.searchsorted(tc_s, bounds, side="right") - np.searchsorted(tp_s, bounds, side="right")
        widths = np.diff(np.concatenate([bounds, [np.inf]]))
        with np.errstate(invalid="ignore"):
            haz = np.cumsum(np.where(kk > 0, kk * widths, 0.0)) / two_ne
        e = rng.exponential(1.0)
        i = int(np.searchsorted(haz, e, side="left"))
        prev = haz[i - 1] if i > 0 else 0.0
        T = bounds[i] + (e - prev) * two_ne / kk[i]
        alive = cand[(tc <= T) & (T < tp)]
        c = int(alive[rng.integers(alive.size)])
        if T <= time[c]:
            T = np.nextafter(time[c], np.inf)
        # regraft: new node q above c at time T
        if size == parent.size:
            k = parent.size
            parent = np.concatenate([parent, np.full(k, -1, dtype=np.int64)])
            time = np.concatenate([time, np.zeros(k)])
            edge_left = np.concatenate([edge_left, np.zeros(k)])
            active = np.concatenate([active, np.zeros(k, dtype=bool)])
            child = np.concatenate([child, np.full((k, 2), -1, dtype=np.int64)])
        q = size
        size += 1
        time[q] = T
        active[q] = True
        pc = int(parent[c])
        close_edge(c, bp)
        parent[q] = pc
        replace_child(pc, c, q)
        edge_left[q] = bp
        parent[c] = q
        edge_left[c] = bp
        parent[u] = q
        edge_left[u] = bp
        child[q, 0] = c
        child[q, 1] = u
    parent, time = parent[:size], time[:size]
    for c in np.flatnonzero(parent >= 0):
        close_edge(c, L)
    el = np.array(out_l, dtype=np.float64)
    er = np.array(out_r, dtype=np.float64)
    ep = np.array(out_p, dtype=np.int64)
    ec = np.array(out_c, dtype=np.int64)
    return finalise_tables(L, time, n, el, er, ep, ec)


def finalise_tables(L, time, n_samples, el, er, ep, ec):
    """Drops unused nodes, renumbers by time, sorts edges the tskit way, returns the dict."""
    used = np.zeros(time.size, dtype=bool)
    used[:n_samples] = True
    used[ep] = True
    used[ec] = True
    old = np.flatnonzero(used)
    # samples keep ids 0..n-1; the rest in increasing time
    rest = old[old >= n_samples]
    rest = rest[np.argsort(time[rest], kind="stable")]
    order = np.concatenate([np.arange(n_samples), rest])
    new_id = np.full(time.size, -1, dtype=np.int64)
    new_id[order] = np.arange(order.size)
    t = time[order]
    p, c = new_id[ep], new_id[ec]
    key = np.lexsort((el, c, p, t[p]))
    flags = np.zeros(order.size, dtype=np.uint32)
    flags[:n_samples] = 1
    return {
        "sequence_length": float(L),
        "time_units": "generations",
        "nodes": {"time": t, "flags": flags},
        "edges": {
            "left": el[key],
            "right": er[key],
            "parent": p[key].astype(np.int32),
            "child": c[key].astype(np.int32),
        },
        "sites": empty_sites(),
        "mutations": empty_mutations(),
    }


def empty_sites():
    return {
        "position": np.zeros(0, dtype=np.float64),
        "ancestral_state": np.zeros(0, dtype=np.int8),
        "ancestral_state_offset": np.zeros(1, dtype=np.uint64),
        "metadata": np.zeros(0, dtype=np.int8),
        "metadata_offset": np.zeros(1, dtype=np.uint64),
    }


def empty_mutations():
    return {
        "site": np.zeros(0, dtype=np.int32),
        "node": np.zeros(0, dtype=np.int32),
        "parent": np.zeros(0, dtype=np.int32),
        "time": np.zeros(0, dtype=np.float64),
        "derived_state": np.zeros(0, dtype=np.int8),
        "derived_state_offset": np.zeros(1, dtype=np.uint64),
        "metadata": np.zeros(0, dtype=np.int8),
        "metadata_offset": np.zeros(1, dtype=np.uint64),
    }
