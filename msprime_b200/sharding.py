"""Genome-sharded overlays: one GPU context per rank, each owning a position slice.

The path shards by genome position: edges, nodes and the rate map
are replicated, a site never straddles two slices, and the native stream is keyed by
(seed, edge, ordinal) rather than by rank, so the union of the slices' tables is identical
to the single-GPU tables.  What has to be exchanged is tiny: per-rank counts, to turn
rank-local site ids, parent ids, allele numbers and byte offsets into global ones.

Two small all-gathers per call (``allgather`` is any callable that maps this rank's list
of ints to the list of every rank's lists, e.g. :func:`dist_allgather`):

1. after ``generate_begin``: (sites, mutations, allele numbers used)  -> id / allele bases;
2. after ``generate_end``:   (ancestral, site-metadata, derived, mutation-metadata bytes)
   -> byte-offset bases, applied on the device by ``shift_offsets``.
"""
from __future__ import annotations

import numpy as np

SITE_COLUMNS = ("position", "ancestral_state", "ancestral_state_offset", "metadata", "metadata_offset")
MUTATION_COLUMNS = ("site", "node", "parent", "time", "derived_state", "derived_state_offset", "metadata",
                    "metadata_offset")


def slice_cuts(sequence_length, world, discrete_genome=True):
    """``world + 1`` increasing cut points from 0 to L; integers when the genome is discrete
    (a site must belong to exactly one slice)."""
    L = float(sequence_length)
    cuts = [L * r / world for r in range(world + 1)]
    if discrete_genome:
        cuts = [float(np.floor(c)) for c in cuts]
    cuts[0], cuts[-1] = 0.0, L
    for a, b in zip(cuts[:-1], cuts[1:]):
        if not a < b:
            raise ValueError("sequence too short to cut into that many slices")
    return cuts


def balanced_cuts(tables, map_position, map_rate, world, *, discrete_genome=True, start_time=None, end_time=None,
                  bins=16384):
    """``world + 1`` cut points such that every slice holds about the same expected number of new mutations
    ("the cuts are chosen so that the total branch mass is balanced"), where equal-length
    slices (:func:`slice_cuts`) differ as much as the rate map does.

    The mutation mass of [0, x) is the integral of rate(x) times the total branch length (inside the time
    window) of the edges covering x.  It is accumulated over ``bins`` equal stretches of the genome: the branch
    length alive inside a stretch comes from the edges' end points (two weighted histograms), the rate mass
    of a stretch from the rate map; inside a stretch both are taken as uniform, which is exact for the
    purpose (cuts only have to balance the slices, any cuts give the same distribution of results)."""
    L = float(tables["sequence_length"])
    if world == 1:
        return [0.0, L]
    e, nt = tables["edges"], np.asarray(tables["nodes"]["time"], dtype=np.float64)
    lo, hi = nt[np.asarray(e["child"])], nt[np.asarray(e["parent"])]
    if start_time is not None:
        lo = np.maximum(lo, start_time)
    if end_time is not None:
        hi = np.minimum(hi, end_time)
    branch = np.maximum(hi - lo, 0.0)
    left, right = np.asarray(e["left"], dtype=np.float64), np.asarray(e["right"], dtype=np.float64)
    w = L / bins
    edges_ = np.arange(bins + 1) * w
    # covered(x) = sum over edges of branch * length of (edge intersected with [0, x)) at the bin boundaries
    #            = sum branch * min(x, right) - sum branch * min(x, left); for a set of points p with weights b,
    # sum b * min(X_j, p) = X_j * (weight of the points in bins >= j) + (sum of b * p over the bins < j)
    def capped_sum(points):
        k = np.minimum((points / w).astype(np.int64), bins - 1)
        weight = np.bincount(k, weights=branch, minlength=bins)
        moment = np.bincount(k, weights=branch * points, minlength=bins)
        beyond = np.concatenate([np.cumsum(weight[::-1])[::-1], [0.0]])
        before = np.concatenate([[0.0], np.cumsum(moment)])
        return edges_ * beyond + before

    covered = capped_sum(right) - capped_sum(left)
    alive = np.diff(covered) / w  # mean branch length alive inside every bin
    mp, mr = np.asarray(map_position, dtype=np.float64), np.asarray(map_rate, dtype=np.float64)
    cum_rate = np.concatenate([[0.0], np.cumsum(np.diff(mp) * mr)])
    rate_mass = np.diff(np.interp(edges_, mp, cum_rate))
    mass = np.concatenate([[0.0], np.cumsum(alive * rate_mass)])
    if not mass[-1] > 0:
        return slice_cuts(L, world, discrete_genome)
    cuts = np.interp(np.arange(world + 1) * mass[-1] / world, mass, edges_)
    cuts = [float(np.floor(c)) if discrete_genome else float(c) for c in cuts]
    cuts[0], cuts[-1] = 0.0, L
    for a, b in zip(cuts[:-1], cuts[1:]):
        if not a < b:
            return slice_cuts(L, world, discrete_genome)
    return cuts


def exclusive_bases(all_values, rank):
    """Sum over ranks before ``rank`` of each column of ``all_values`` (world x k ints)."""
    a = np.asarray(all_values, dtype=object)
    k = a.shape[1]
    return [int(sum(int(a[r][j]) for r in range(rank))) for j in range(k)]


def dist_allgather(group=None, device=None):
    """An ``allgather`` callable on top of ``torch.distributed`` (gloo on CPU tensors, nccl on
    ``device``).  Values travel as int64."""
    import torch
    import torch.distributed as dist

    def allgather(values):
        world = dist.get_world_size(group)
        # uint64 values (allele counters) travel as their two's-complement int64 image
        signed = [((int(v) & 0xFFFFFFFFFFFFFFFF) ^ (1 << 63)) - (1 << 63) for v in values]
        t = torch.tensor(signed, dtype=torch.int64, device=device)
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t, group=group)
        return [[int(x) & 0xFFFFFFFFFFFFFFFF for x in o.cpu().tolist()] for o in out]

    return allgather


def run_slice(engine, params, *, rank, world, allgather, start_allele=0, fetch=True, out=None):
    """Runs this rank's slice on ``engine`` (a ``_native.Context`` with tables loaded, or
    anything with the same four methods) and returns ``(local_result, info)``.

    ``params`` must already carry this rank's ``slice_left/right``.  The local result has
    global site ids, parent ids, allele numbers and byte offsets; ``info`` holds the global
    totals (``num_sites``, ``num_mutations``, ``next_allele``, byte totals).  ``fetch=False``
    leaves the slice resident on the device; ``out`` passes preallocated host buffers."""
    c1 = engine.generate_begin(params)
    used = (int(c1["next_allele"]) - int(start_allele)) & 0xFFFFFFFFFFFFFFFF
    all1 = allgather([c1["num_sites"], c1["num_mutations"], used, c1["num_new_mutations"]])
    site_base, mut_base, allele_base, _ = exclusive_bases(all1, rank)
    c2 = engine.generate_end(site_base, mut_base, (int(start_allele) + allele_base) & 0xFFFFFFFFFFFFFFFF)
    all2 = allgather([c2["ancestral_state_bytes"], c2["site_metadata_bytes"], c2["derived_state_bytes"],
                      c2["mutation_metadata_bytes"]])
    anc, smd, der, mmd = exclusive_bases(all2, rank)
    engine.shift_offsets(anc, smd, der, mmd)
    local = None
    if fetch:
        local = engine.fetch(out) if out is not None else engine.fetch()
    tot1 = [sum(int(r[j]) for r in all1) for j in range(4)]
    tot2 = [sum(int(r[j]) for r in all2) for j in range(4)]
    info = {
        "num_sites": tot1[0], "num_mutations": tot1[1], "num_new_mutations": tot1[3],
        "next_allele": (int(start_allele) + tot1[2]) & 0xFFFFFFFFFFFFFFFF,
        "ancestral_state_bytes": tot2[0], "site_metadata_bytes": tot2[1], "derived_state_bytes": tot2[2],
        "mutation_metadata_bytes": tot2[3], "site_base": site_base, "mutation_base": mut_base,
    }
    return local, info


def nccl_device_allgather(ctx, world, group=None):
    """``gather()`` for :func:`run_slice_device`: all-gathers the context's send buffer into
    its recv buffer with NCCL, enqueued on the stream the context names (no host wait)."""
    import torch
    import torch.distributed as dist

    send, recv = ctx.exchange_tensors(world)
    streams = {}

    def gather():
        # the context's own stream, or its second one while a call's exchanges run beside K4
        handle = ctx.exchange_stream()
        if handle not in streams:
            streams[handle] = torch.cuda.ExternalStream(handle, device=send.device)
        with torch.cuda.stream(streams[handle]):
            dist.all_gather_into_tensor(recv, send, group=group)

    return gather


def load_tables_sharded(ctx, tables_dict, *, rank, world, group=None, all_gather=None):
    """Replicates the tree sequence on every rank's GPU with ONE copy of it crossing PCIe in total:
    each rank uploads its 1 / world share of the node and edge columns
    (``mspb_ctx_load_tables_begin``), the shares are all-gathered in place over NVLink (NCCL,
    enqueued on the context's stream), then every rank validates and builds its index
    (``mspb_ctx_load_tables_end``).  ``all_gather(full, mine)`` replaces the NCCL call in tests."""
    import torch

    from . import _native

    shards = ctx.load_tables_begin(tables_dict, rank, world)
    if world > 1:
        dev = torch.device("cuda", ctx.device)
        stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
        if all_gather is None:
            import torch.distributed as dist

            def all_gather(full, mine):
                dist.all_gather_into_tensor(full, mine, group=group)

        with torch.cuda.stream(stream):
            for ptr, per in shards:
                if per == 0:
                    continue
                full = torch.as_tensor(_native._DeviceArray(ptr, per * world, "|u1"), device=dev)
                all_gather(full, full[rank * per:(rank + 1) * per])
    ctx.load_tables_end()


def run_slice_device(ctx, params, *, rank, world, gather, fetch=True, out=None, host_between_stages=False):
    """:func:`run_slice` with the exchange kept on the device: ``gather()`` must all-gather the
    context's send buffer into its recv buffer on the context's stream (see
    :func:`nccl_device_allgather`).  The host waits once per overlay, at the end: the first stage is
    enqueued without its closing wait (``generate_begin_device``) and its counts go from the device totals
    into the exchange.  When some rank needs the host between the stages after all (an error, a dense
    genome's second sort, positions to draw again) every rank learns it from the exchanged counts and the
    call is made again the slower way, all ranks together.  ``host_between_stages=True`` (or the SLiM
    model) takes that way at once.  ``params.start_allele`` is the global allele counter before the call."""
    from . import _native

    totals = None
    if not host_between_stages and params.struct.model_kind != 2:
        ctx.generate_begin_device(params)
        ctx.exchange_post_counts()
        gather()
        ctx.generate_end_device(rank, world)
        gather()
        try:
            _, totals, bases = ctx.generate_finish_device(rank, world)
        except _native.RetryWithHost:
            totals = None
    if totals is None:
        ctx.generate_begin(params)
        ctx.exchange_post_counts()
        gather()
        ctx.generate_end_device(rank, world)
        gather()
        _, totals, bases = ctx.generate_finish_device(rank, world)
    local = None
    if fetch:
        local = ctx.fetch(out) if out is not None else ctx.fetch()
    info = {
        "num_sites": totals["num_sites"], "num_mutations": totals["num_mutations"],
        "num_new_mutations": totals["num_new_mutations"], "next_allele": ctx.counts["next_allele"],
        "ancestral_state_bytes": totals["ancestral_state_bytes"],
        "site_metadata_bytes": totals["site_metadata_bytes"],
        "derived_state_bytes": totals["derived_state_bytes"],
        "mutation_metadata_bytes": totals["mutation_metadata_bytes"],
        "site_base": bases["num_sites"], "mutation_base": bases["num_mutations"],
    }
    return local, info


def merge_slices(results):
    """Concatenates the per-slice results (in slice order) into one result.  Offsets are already
    global: every slice but the last drops its closing offset entry."""
    out = {"sites": {}, "mutations": {}}
    last = len(results) - 1
    for tbl, cols in (("sites", SITE_COLUMNS), ("mutations", MUTATION_COLUMNS)):
        for col in cols:
            parts = []
            for r, res in enumerate(results):
                a = np.asarray(res[tbl][col])
                if col.endswith("_offset") and r != last:
                    a = a[:-1]
                parts.append(a)
            out[tbl][col] = np.concatenate(parts) if parts else np.zeros(0)
    return out


def gather_results(local, *, rank, world, group=None):
    """Final table gather to rank 0 (returns the merged result there, None elsewhere)."""
    import torch.distributed as dist

    gathered = [None] * world if rank == 0 else None
    dist.gather_object(local, gathered, dst=0, group=group)
    if rank != 0:
        return None
    return merge_slices(gathered)
