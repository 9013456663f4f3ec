"""PROTOTYPE (numpy, CPU) of the level-synchronous sample-count algorithm planned for the device
consumers (plan item 4 (v)).  Not product code and not on
any product path: it exists to pin the algorithm down against tskit before kernels are written
(tests/test_below_prototype.py) and to measure, on the bench inputs, the two numbers that size the
kernels: how many levels there are and how many break points the step functions hold.

below(u, x) = number of samples under node u in the tree covering x.  As a function of x it is a step
function f_u, and

    f_u = [u is a sample] on the stretches where u is in a tree ... + sum over edges (u, c, l, r) of f_c clipped to [l, r)

so nodes can be processed children-first.  A LEVEL is the set of nodes whose children are all done
(level(u) = 1 + max level of its children, samples without children are level 0).  One level is one
data-parallel step: every edge into a node of the level contributes the events of its child's function
inside [l, r) (value at l, the child's own changes, minus the value at r); all events of the level are
sorted by (node, x) and a segmented running sum gives the functions; points where nothing changes are
dropped.  A query (u, x) is a binary search in f_u.

Functions are stored CSR-style: for node u, breaks[start[u]:start[u+1]] are the x where f_u changes and
values[...] the value from that x on (0 before the first break).
"""
import numpy as np


def node_levels(num_nodes, parent, child):
    """level[u] = 0 for nodes without children, else 1 + max level over its children (edges are
    in tskit order, i.e. by parent time: a parent's children were all seen as parents earlier,
    so one pass over the edges in order is enough)."""
    level = np.zeros(num_nodes, dtype=np.int64)
    for p, c in zip(parent, child):  # plain loop: prototype; the device version relaxes with scatter-max
        if level[p] < level[c] + 1:
            level[p] = level[c] + 1
    return level


class StepFunctions:
    def __init__(self, num_nodes):
        self.start = np.zeros(num_nodes + 1, dtype=np.int64)
        self.breaks = np.zeros(0, dtype=np.float64)
        self.values = np.zeros(0, dtype=np.int64)

    def value_at(self, nodes, x):
        """f_node(x) for arrays of queries."""
        nodes, x = np.asarray(nodes), np.asarray(x, dtype=np.float64)
        out = np.zeros(len(nodes), dtype=np.int64)
        for i, (u, q) in enumerate(zip(nodes, x)):
            a, b = self.start[u], self.start[u + 1]
            k = np.searchsorted(self.breaks[a:b], q, side="right")
            out[i] = self.values[a + k - 1] if k > 0 else 0
        return out


def build(tables):
    """Step functions of every node of the tree sequence (``asdict()``-style numpy columns)."""
    flags = np.asarray(tables["nodes"]["flags"])
    e = tables["edges"]
    left, right = np.asarray(e["left"], dtype=np.float64), np.asarray(e["right"], dtype=np.float64)
    parent, child = np.asarray(e["parent"], dtype=np.int64), np.asarray(e["child"], dtype=np.int64)
    N, L = len(flags), float(tables["sequence_length"])
    level = node_levels(N, parent, child)
    is_sample = (flags & 1) != 0

    # per-node function storage while building (lists of arrays; CSR at the end)
    fb = [None] * N
    fv = [None] * N
    stats = {"levels": int(level.max()) + 1 if N else 0, "events_sorted": 0, "break_points": 0}
    edge_order = np.argsort(level[parent], kind="stable")
    bounds = np.searchsorted(level[parent][edge_order], np.arange(stats["levels"] + 1))
    for lv in range(stats["levels"]):
        nodes_here = np.flatnonzero(level == lv)
        ev_node, ev_x, ev_d = [], [], []
        # a sample counts itself wherever it is in a tree; as a leaf that is the whole genome
        s = nodes_here[is_sample[nodes_here]]
        ev_node += [s, s]
        ev_x += [np.zeros(len(s)), np.full(len(s), L)]
        ev_d += [np.ones(len(s), dtype=np.int64), -np.ones(len(s), dtype=np.int64)]
        for j in edge_order[bounds[lv]:bounds[lv + 1]]:  # edges into this level (device: one thread block each)
            c, p, l, r = child[j], parent[j], left[j], right[j]
            b, v = fb[c], fv[c]
            lo = np.searchsorted(b, l, side="right")  # breaks <= l are summarised by the value at l
            hi = np.searchsorted(b, r, side="left")  # breaks in (l, r) are the child's own changes
            at_l = v[lo - 1] if lo > 0 else 0
            inner_x = b[lo:hi]
            inner_v = v[lo:hi]
            prev = np.concatenate([[at_l], inner_v[:-1]]) if hi > lo else np.zeros(0, dtype=np.int64)
            at_r = inner_v[-1] if hi > lo else at_l
            ev_node.append(np.full(hi - lo + 2, p))
            ev_x.append(np.concatenate([[l], inner_x, [r]]))
            ev_d.append(np.concatenate([[at_l], inner_v - prev, [-at_r]]).astype(np.int64))
        if not ev_node:
            continue
        node = np.concatenate(ev_node)
        x = np.concatenate(ev_x)
        d = np.concatenate(ev_d)
        stats["events_sorted"] += len(node)
        order = np.lexsort((x, node))  # device: one radix sort of the level's events by (node, x)
        node, x, d = node[order], x[order], d[order]
        # merge events at the same (node, x), drop the ones that change nothing
        first = np.ones(len(node), dtype=bool)
        first[1:] = (node[1:] != node[:-1]) | (x[1:] != x[:-1])
        group = np.cumsum(first) - 1
        dsum = np.bincount(group, weights=d).astype(np.int64)
        gnode, gx = node[first], x[first]
        keep = dsum != 0
        gnode, gx, dsum = gnode[keep], gx[keep], dsum[keep]
        # segmented running sum per node
        total = np.cumsum(dsum)
        seg_first = np.ones(len(gnode), dtype=bool)
        seg_first[1:] = gnode[1:] != gnode[:-1]
        seg_begin = np.maximum.accumulate(np.where(seg_first, np.arange(len(gnode)), 0))
        run = total - (total - dsum)[seg_begin]
        cuts = np.flatnonzero(seg_first)
        ends = np.append(cuts[1:], len(gnode))
        for a, b_ in zip(cuts, ends):
            u = gnode[a]
            fb[u], fv[u] = gx[a:b_], run[a:b_]
        for u in nodes_here:  # nodes of this level with no events at all
            if fb[u] is None:
                fb[u], fv[u] = np.zeros(0), np.zeros(0, dtype=np.int64)
    out = StepFunctions(N)
    sizes = np.array([0 if b is None else len(b) for b in fb], dtype=np.int64)
    out.start[1:] = np.cumsum(sizes)
    out.breaks = np.concatenate([b for b in fb if b is not None and len(b)]) if sizes.sum() else np.zeros(0)
    out.values = np.concatenate([v for v in fv if v is not None and len(v)]) if sizes.sum() else np.zeros(0, np.int64)
    stats["break_points"] = int(sizes.sum())
    return out, stats


def site_afs(tables, fns=None):
    """Polarised, un-normalised site AFS of all samples from `below` alone (no genotype matrix):
    every mutation gives below(node, position) to its allele and takes it from its parent
    mutation's allele (or the ancestral state); alleles are compared as byte strings."""
    if fns is None:
        fns, _ = build(tables)
    s, m = tables["sites"], tables["mutations"]
    n = int(np.sum((np.asarray(tables["nodes"]["flags"]) & 1) != 0))
    pos = np.asarray(s["position"], dtype=np.float64)
    site = np.asarray(m["site"], dtype=np.int64)
    below = fns.value_at(np.asarray(m["node"], dtype=np.int64), pos[site])
    parent = np.asarray(m["parent"], dtype=np.int64)

    def ragged(data, off):
        raw = np.asarray(data).view(np.uint8).tobytes()
        off = np.asarray(off).astype(np.int64)
        return [raw[off[i]:off[i + 1]] for i in range(len(off) - 1)]

    anc = ragged(s["ancestral_state"], s["ancestral_state_offset"])
    der = ragged(m["derived_state"], m["derived_state_offset"])
    afs = np.zeros(n + 1)
    first = np.searchsorted(site, np.arange(len(pos) + 1))
    for j in range(len(pos)):
        count = {anc[j]: n}
        for k in range(first[j], first[j + 1]):
            count[der[k]] = count.get(der[k], 0) + below[k]
            source = anc[j] if parent[k] < 0 else der[parent[k]]
            count[source] -= below[k]
        for name, c in count.items():
            if name != anc[j] and 0 < c < n:
                afs[c] += 1
    return afs


def _range_search(keys, begin, count, target, right_side):
    """Per query i: how many of keys[begin[i] : begin[i] + count[i]] (ascending) are <= target[i]
    (``right_side``) or < target[i].  Vectorised bisection."""
    lo = np.zeros(len(begin), dtype=np.int64)
    hi = count.astype(np.int64).copy()
    while True:
        active = lo < hi
        if not active.any():
            return lo
        mid = (lo + hi) // 2
        k = keys[np.where(active, begin + mid, 0)]
        go_right = (k <= target) if right_side else (k < target)
        lo = np.where(active & go_right, mid + 1, lo)
        hi = np.where(active & ~go_right, mid, hi)


def build_fast(tables):
    """The same functions as :func:`build`, a whole level at a time in numpy (no per-edge Python), the
    way the device version will work: CSR storage that only grows, and every (node, x) pair folded into
    one float64 key ``node * 2L + x`` (exact while break points are integers and ``N * 2L < 2**53``:
    true of the discrete-genome bench inputs).  Used to measure levels / events / break points at full
    size; returns ``(StepFunctions, stats)`` with per-level event counts in ``stats["per_level"]``."""
    flags = np.asarray(tables["nodes"]["flags"])
    e = tables["edges"]
    left, right = np.asarray(e["left"], dtype=np.float64), np.asarray(e["right"], dtype=np.float64)
    parent, child = np.asarray(e["parent"], dtype=np.int64), np.asarray(e["child"], dtype=np.int64)
    N, L = len(flags), float(tables["sequence_length"])
    W = 2.0 * L
    assert N * W < 2.0**53 and np.all(left == np.floor(left)) and np.all(right == np.floor(right))
    # levels: relax until nothing changes (device: one scatter-max kernel per round)
    level = np.zeros(N, dtype=np.int64)
    while True:
        cand = np.zeros(N, dtype=np.int64)
        np.maximum.at(cand, parent, level[child] + 1)
        cand = np.maximum(cand, level)
        if np.array_equal(cand, level):
            break
        level = cand
    num_levels = int(level.max()) + 1 if N else 0
    is_sample = (flags & 1) != 0
    edge_order = np.argsort(level[parent], kind="stable")
    bounds = np.searchsorted(level[parent][edge_order], np.arange(num_levels + 1))

    start = np.zeros(N + 1, dtype=np.int64)  # CSR over nodes, filled level by level
    size = np.zeros(N, dtype=np.int64)
    keys_all = np.zeros(0)  # node * W + x of every break point stored so far, ascending within a node
    vals_all = np.zeros(0, dtype=np.int64)
    where = np.full(N, -1, dtype=np.int64)  # offset of node's function inside keys_all
    stats = {"levels": num_levels, "events_sorted": 0, "break_points": 0, "per_level": []}
    for lv in range(num_levels):
        nodes_here = np.flatnonzero(level == lv)
        s = nodes_here[is_sample[nodes_here]]
        ev_key = [s * W, s * W + L]
        ev_d = [np.ones(len(s), dtype=np.int64), -np.ones(len(s), dtype=np.int64)]
        je = edge_order[bounds[lv]:bounds[lv + 1]]
        if len(je):
            c, p, l, r = child[je], parent[je], left[je], right[je]
            a, n_c = where[c], size[c]
            # binary search inside each child's own range of keys_all (what a device thread would do)
            lo = _range_search(keys_all, a, n_c, c * W + l, right_side=True)
            hi = _range_search(keys_all, a, n_c, c * W + r, right_side=False)
            lo = np.where(n_c > 0, lo, 0)
            hi = np.where(n_c > 0, hi, 0)
            at_l = np.where(lo > 0, vals_all[np.clip(a + lo - 1, 0, max(len(vals_all) - 1, 0))] if len(vals_all) else 0, 0)
            at_r = np.where(hi > 0, vals_all[np.clip(a + hi - 1, 0, max(len(vals_all) - 1, 0))] if len(vals_all) else 0, 0)
            at_r = np.where(hi > lo, at_r, at_l)
            inner = hi - lo
            tot = int(inner.sum())
            if tot:
                owner = np.repeat(np.arange(len(je)), inner)
                offs = np.arange(tot) - np.repeat(np.cumsum(inner) - inner, inner)
                src = a[owner] + lo[owner] + offs
                x_in = keys_all[src] - c[owner] * W
                v_in = vals_all[src]
                prev = np.where(offs == 0, at_l[owner], vals_all[np.maximum(src - 1, 0)])
                ev_key.append(p[owner] * W + x_in)
                ev_d.append(v_in - prev)
            ev_key += [p * W + l, p * W + r]
            ev_d += [at_l.astype(np.int64), -at_r.astype(np.int64)]
        key = np.concatenate(ev_key)
        d = np.concatenate(ev_d)
        stats["events_sorted"] += len(key)
        order = np.argsort(key, kind="stable")
        key, d = key[order], d[order]
        first = np.ones(len(key), dtype=bool)
        first[1:] = key[1:] != key[:-1]
        dsum = np.add.reduceat(d, np.flatnonzero(first)) if len(key) else d
        gkey = key[first]
        keep = dsum != 0
        gkey, dsum = gkey[keep], dsum[keep]
        gnode = np.floor(gkey / W).astype(np.int64)
        total = np.cumsum(dsum)
        seg_first = np.ones(len(gnode), dtype=bool)
        seg_first[1:] = gnode[1:] != gnode[:-1]
        seg_begin = np.maximum.accumulate(np.where(seg_first, np.arange(len(gnode)), 0))
        run = total - (total - dsum)[seg_begin]
        cuts = np.flatnonzero(seg_first)
        counts = np.diff(np.append(cuts, len(gnode)))
        where[gnode[cuts]] = len(keys_all) + cuts
        size[gnode[cuts]] = counts
        keys_all = np.concatenate([keys_all, gkey])
        vals_all = np.concatenate([vals_all, run])
        stats["per_level"].append((len(nodes_here), len(je), len(key), len(gkey)))
    # CSR in node order
    out = StepFunctions(N)
    out.start[1:] = np.cumsum(size)
    order = np.argsort(np.where(size > 0, where, np.iinfo(np.int64).max), kind="stable")
    order = order[size[order] > 0]
    idx = np.repeat(where[order], size[order]) + (np.arange(int(size.sum())) - np.repeat(np.cumsum(size[order]) - size[order], size[order]))
    node_of = np.repeat(order, size[order])
    pos = np.argsort(node_of, kind="stable")
    out.breaks = (keys_all[idx] - node_of * W)[pos]
    out.values = vals_all[idx][pos]
    stats["break_points"] = int(size.sum())
    return out, stats
