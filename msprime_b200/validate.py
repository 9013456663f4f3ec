"""Structural checks of an overlay result against its input tree sequence (numpy only).

These are the invariants every valid result of ``mutgen_generate`` obeys, whatever random stream
produced it (``lib/mutgen.c:1027-1074, 1103-1151, 1160-1224`` and tskit's table requirements).  They are
what ``bench.py`` runs on the last timed overlay (its ``"validated"`` key) and what the size tests in
``tests/test_gpu_baseline_sizes.py`` run at BASELINE.json's full sizes, where no CPU oracle run is
compared row by row."""
from __future__ import annotations

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


def check_result(ts, res, *, discrete, start_time=None, end_time=None, kept_mutations=0, new_rows=None):
    """Asserts the invariants every valid overlay obeys (lib/mutgen.c:1027-1074 and tskit's
    table requirements): sites sorted and unique, mutations grouped by site in site order,
    parent before child at the same site, each new mutation on an edge of its node that
    covers the site, with its time on that branch (clamped to the window).

    ``new_rows`` (boolean mask over the mutation rows) names the rows this call drew when rows were
    kept: the branch / window checks then apply to those rows only."""
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
    if (kept_mutations == 0 or new_rows is not None) and M > 0:
        rows = np.ones(M, dtype=bool) if new_rows is None else np.asarray(new_rows, dtype=bool)
        e = edge_of(ts, node[rows], pos[site[rows]])
        assert np.all(e >= 0)
        nt = ts["nodes"]["time"]
        lo = nt[node[rows]]
        hi = nt[ts["edges"]["parent"][e]]
        if start_time is not None:
            lo = np.maximum(lo, start_time)
        if end_time is not None:
            hi = np.minimum(hi, end_time)
        assert np.all(t[rows] >= lo) and np.all(t[rows] < hi)
    if kept_mutations == 0 and M > 0:
        # the parent mutation sits on an ancestor-or-self branch, hence is not younger
        assert np.all(t[par[has]] >= t[has])
    return True


def expected_new_mutations(ts, map_position, map_rate, *, discrete, start_time=None, end_time=None):
    """Mean of the number of new mutations: the sum over edges of branch length (inside the time window)
    times the rate mass of the edge's span (lib/mutgen.c:1109-1119; in the discrete genome the mass of
    [a, b) sits on the integers it contains)."""
    e, nt = ts["edges"], np.asarray(ts["nodes"]["time"])
    lo, hi = nt[e["child"]], nt[e["parent"]]
    if start_time is not None:
        lo = np.maximum(lo, start_time)
    if end_time is not None:
        hi = np.minimum(hi, end_time)
    branch = np.maximum(hi - lo, 0.0)
    mp, mr = np.asarray(map_position, dtype=np.float64), np.asarray(map_rate, dtype=np.float64)
    coord = np.ceil if discrete else (lambda x: x)
    cum = np.concatenate([[0.0], np.cumsum((coord(mp[1:]) - coord(mp[:-1])) * mr)])

    def mass_upto(x):
        i = np.clip(np.searchsorted(mp, x, side="right") - 1, 0, len(mr) - 1)
        return cum[i] + mr[i] * (coord(x) - coord(mp[i]))

    return float(np.sum(branch * (mass_upto(np.asarray(e["right"])) - mass_upto(np.asarray(e["left"])))))


def check_count(ts, num_new, map_position, map_rate, *, discrete, start_time=None, end_time=None, sigmas=5.0):
    """|new mutations - mean| within ``sigmas`` Poisson standard deviations; returns (mean, z)."""
    mean = expected_new_mutations(ts, map_position, map_rate, discrete=discrete, start_time=start_time,
                                  end_time=end_time)
    z = (num_new - mean) / np.sqrt(max(mean, 1.0))
    assert abs(z) < sigmas, f"{num_new} new mutations against a mean of {mean:.1f} (z = {z:.2f})"
    return mean, float(z)


def check_infinite_alleles_numbering(res, start_allele=0):
    """tests/test_mutations.py:1704-1711 for a call without kept rows: allele numbers count up from
    ``start_allele`` in site order, a site's ancestral state first, then its mutations in row order.
    Checked without building Python strings: the decimal text of every number is rebuilt from the
    byte columns with numpy."""
    s, m = res["sites"], res["mutations"]
    S, M = len(s["position"]), len(m["site"])
    site = np.asarray(m["site"]).astype(np.int64)
    # number of the ancestral state of site j: start + j + (mutations before site j)
    first = np.searchsorted(site, np.arange(S))
    anc_num = start_allele + np.arange(S, dtype=np.int64) + first
    der_num = start_allele + np.arange(M, dtype=np.int64) + site + 1

    def decode(col, off, n):
        off = np.asarray(off).astype(np.int64)
        b = np.asarray(col).astype(np.int64) - 48
        assert len(b) == 0 or (b.min() >= 0 and b.max() <= 9)
        length = np.diff(off)
        assert len(length) == n and (n == 0 or length.min() >= 1)
        # value = sum of digit * 10^(digits after it)
        row = np.repeat(np.arange(n), length)
        after = off[1:][row] - 1 - np.arange(len(b))
        # (float weights are exact here: every partial sum is an integer below 2^53)
        assert n == 0 or length.max() <= 15
        val = np.bincount(row, weights=(b * 10 ** after).astype(np.float64), minlength=n).astype(np.int64)
        # no leading zeros (except the number 0 itself)
        lead = b[off[:-1]]
        assert np.all((lead != 0) | (length == 1))
        return val

    assert np.array_equal(decode(s["ancestral_state"], s["ancestral_state_offset"], S), anc_num)
    assert np.array_equal(decode(m["derived_state"], m["derived_state_offset"], M), der_num)
    return True
