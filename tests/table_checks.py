"""Structural checks of an overlay result against its input tree sequence (numpy only)."""
import numpy as np


def ragged(col, off):
    b = np.asarray(col).astype(np.uint8).tobytes()
    off = np.asarray(off, dtype=np.int64)
    return [b[off[i]:off[i + 1]] for i in range(len(off) - 1)]


def edge_of(ts, node, pos):
    """Index of the edge whose child is node[k] and that covers pos[k] (-1 if none): numpy
    lookup through edges sorted by (child, left)."""
    e = ts["edges"]
    order = np.lexsort((e["left"], e["child"]))
    c, l, r = e["child"][order].astype(np.int64), e["left"][order], e["right"][order]
    key = c.astype(np.float64) * (ts["sequence_length"] + 1.0) + l
    q = np.asarray(node, dtype=np.float64) * (ts["sequence_length"] + 1.0) + np.asarray(pos, dtype=np.float64)
    i = np.searchsorted(key, q, side="right") - 1
    ok = (i >= 0) & (c[np.maximum(i, 0)] == node) & (l[np.maximum(i, 0)] <= pos) & (pos < r[np.maximum(i, 0)])
    return np.where(ok, order[np.maximum(i, 0)], -1)


def check_result(ts, res, *, discrete, start_time=None, end_time=None, kept_mutations=0):
    """Asserts the invariants every valid overlay obeys (lib/mutgen.c:1027-1074 and tskit's
    table requirements): sites sorted and unique, mutations grouped by site in site order,
    parent before child at the same site, each new mutation on an edge of its node that
    covers the site, with its time on that branch (clamped to the window)."""
    s, m = res["sites"], res["mutations"]
    pos = np.asarray(s["position"])
    S, M = len(pos), len(m["site"])
    assert np.all(np.diff(pos) > 0)
    assert S == 0 or (pos[0] >= 0 and pos[-1] < ts["sequence_length"])
    if discrete:
        assert np.all(pos == np.floor(pos))
    site = np.asarray(m["site"]).astype(np.int64)
    assert np.all(np.diff(site) >= 0)
    assert M == 0 or (site[0] >= 0 and site[-1] < S)
    for tbl, col, n in (("sites", "ancestral_state", S), ("sites", "metadata", S), ("mutations", "derived_state", M),
                        ("mutations", "metadata", M)):
        off = np.asarray(res[tbl][col + "_offset"]).astype(np.int64)
        assert len(off) == n + 1 and off[0] == 0 and np.all(np.diff(off) >= 0)
        assert off[-1] == len(res[tbl][col])
    par = np.asarray(m["parent"]).astype(np.int64)
    has = par >= 0
    assert np.all(par[has] < np.flatnonzero(has))
    assert np.all(site[par[has]] == site[has])
    t = np.asarray(m["time"])
    # within a site, time is non-increasing (new mutations; kept rows keep their own order)
    if kept_mutations == 0 and M > 1:
        same = site[1:] == site[:-1]
        assert np.all(t[1:][same] <= t[:-1][same])
    node = np.asarray(m["node"]).astype(np.int64)
    if kept_mutations == 0 and M > 0:
        e = edge_of(ts, node, pos[site])
        assert np.all(e >= 0)
        nt = ts["nodes"]["time"]
        lo = nt[node]
        hi = nt[ts["edges"]["parent"][e]]
        if start_time is not None:
            lo = np.maximum(lo, start_time)
        if end_time is not None:
            hi = np.minimum(hi, end_time)
        assert np.all(t >= lo) and np.all(t < hi)
        # the parent mutation sits on an ancestor-or-self branch, hence is not younger
        assert np.all(t[par[has]] >= t[has])
    return True
