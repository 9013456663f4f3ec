"""GPU, native Philox stream: structural validity of the tables, determinism per seed,
slice invariance (the union of genome slices equals the unsliced result, bit for bit),
model bookkeeping (infinite-alleles numbering) and the Python surface end to end."""
import numpy as np
import pytest

import cases
import table_checks
import msprime_b200 as msp
from msprime_b200 import _native, sharding, synth, tables as mtables

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ts_small():
    return synth.native_treeseq(200, 2e5, 1e-8, 1e4, 3)


def run(ctx, ts, *, model, rate, discrete=True, keep=False, seed=1, load=True, **kw):
    if load:
        ctx.load_tables(ts)
    L = ts["sequence_length"]
    p = _native.Params(map_position=kw.pop("map_position", [0, L]), map_rate=kw.pop("map_rate", [rate]),
                       model=cases.native_model(model), discrete_genome=discrete, keep=keep,
                       start_time=kw.pop("start_time", -cases.FMAX), end_time=kw.pop("end_time", cases.FMAX),
                       seed=seed, **kw)
    c = ctx.generate(p)
    return ctx.fetch(), c


@pytest.fixture(scope="module")
def ctx(native_built):
    c = _native.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("model", ["jc69", "binary", "infinite_alleles", "multichar"])
@pytest.mark.parametrize("discrete", [True, False])
@pytest.mark.parametrize("rate", [1e-8, 3e-7, 2e-5])
def test_tables_are_valid(ctx, ts_small, model, discrete, rate):
    res, c = run(ctx, ts_small, model=cases.MODELS[model], rate=rate, discrete=discrete, seed=11)
    assert c["num_mutations"] == len(res["mutations"]["site"]) == c["num_new_mutations"]
    table_checks.check_result(ts_small, res, discrete=discrete)
    if rate >= 3e-7:
        assert c["num_mutations"] > 100


def test_time_window(ctx, ts_small):
    res, _ = run(ctx, ts_small, model=cases.JC69, rate=1e-6, start_time=500.0, end_time=5000.0, seed=5)
    t = res["mutations"]["time"]
    assert len(t) > 100 and t.min() >= 500.0 and t.max() < 5000.0
    table_checks.check_result(ts_small, res, discrete=True, start_time=500.0, end_time=5000.0)


def test_seed_determinism(ctx, ts_small):
    a, _ = run(ctx, ts_small, model=cases.JC69, rate=1e-6, seed=7)
    b, _ = run(ctx, ts_small, model=cases.JC69, rate=1e-6, seed=7, load=False)
    c, _ = run(ctx, ts_small, model=cases.JC69, rate=1e-6, seed=8, load=False)
    for tbl in a:
        for col in a[tbl]:
            assert np.array_equal(a[tbl][col], b[tbl][col])
    assert not np.array_equal(a["sites"]["position"], c["sites"]["position"])


@pytest.mark.parametrize("model", ["jc69", "infinite_alleles"])
def test_continuous_genome_collisions_are_redrawn(ctx, ts_small, model):
    """lib/mutgen.c:1127-1134, tests/test_mutations.py:1340: a second layer with the SAME seed on
    a continuous genome draws every position of the first layer again; each such mutation must
    move to a fresh position instead of sharing (or failing on) the kept site's."""
    mdl = cases.MODELS[model]
    first, c1 = run(ctx, ts_small, model=mdl, rate=3e-7, discrete=False, seed=11)
    assert c1["num_new_mutations"] > 1000
    ts2 = cases.with_kept(ts_small, first)
    second, c2 = run(ctx, ts2, model=mdl, rate=3e-7, discrete=False, keep=True, seed=11)
    assert c2["num_new_mutations"] == c1["num_new_mutations"]  # same seed, same counts
    assert c2["num_mutations"] == 2 * c1["num_mutations"]
    pos = second["sites"]["position"]
    assert np.all(np.diff(pos) > 0)  # sorted, no two sites share a position
    assert c2["num_sites"] == 2 * c1["num_sites"]  # nothing merged into a kept site
    assert np.isin(first["sites"]["position"], pos).all()
    table_checks.check_result(ts2, second, discrete=False)


def test_infinite_alleles_numbering_and_counter(ctx, ts_small):
    m = dict(cases.INF_ALLELES, start_allele=123456789012)
    res, c = run(ctx, ts_small, model=m, rate=1e-6, seed=3)
    anc = [int(x) for x in table_checks.ragged(res["sites"]["ancestral_state"], res["sites"]["ancestral_state_offset"])]
    der = [int(x) for x in table_checks.ragged(res["mutations"]["derived_state"],
                                               res["mutations"]["derived_state_offset"])]
    nxt, mi, site = 123456789012, 0, res["mutations"]["site"]
    for s, a in enumerate(anc):
        assert a == nxt
        nxt += 1
        while mi < len(der) and site[mi] == s:
            assert der[mi] == nxt
            nxt += 1
            mi += 1
    assert c["next_allele"] == nxt


def test_counter_wraps(ctx):
    # tests/test_mutations.py:1774-1794: the uint64 counter wraps around
    ts = cases.four_leaf_tree()
    m = dict(cases.INF_ALLELES, start_allele=2**64 - 3)
    res, c = run(ctx, ts, model=m, rate=1.0, seed=2)
    anc = [int(x) for x in table_checks.ragged(res["sites"]["ancestral_state"], res["sites"]["ancestral_state_offset"])]
    assert anc[0] == 2**64 - 3 and c["next_allele"] < 2**63 and 0 in anc + [int(x) for x in table_checks.ragged(
        res["mutations"]["derived_state"], res["mutations"]["derived_state_offset"])]


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("model,keep", [("infinite_alleles", False), ("jc69", False), ("jc69", True),
                                        ("infinite_alleles", True), ("slim", False), ("slim", True)])
def test_union_of_slices_equals_whole(native_built, ts_small, world, model, keep):
    """Every rank of a sharded call is emulated by its own context on this GPU."""
    ts = ts_small
    L = ts["sequence_length"]
    mdl = cases.MODELS[model]
    whole_ctx = _native.Context(0)
    if keep:
        first, _ = run(whole_ctx, ts, model=mdl, rate=5e-7, seed=21)
        ts = cases.with_kept(ts, first)
    # (the SLiM model's ids run on across the slices like the allele numbers do: its derived states stack the
    # ids along parent chains, so a slice's byte counts depend on the base it learns from the other slices)
    start = 1000 if model == "infinite_alleles" else (mdl["next_id"] if model == "slim" else 0)
    mdl = dict(mdl, start_allele=start) if model == "infinite_alleles" else mdl
    whole, cw = run(whole_ctx, ts, model=mdl, rate=5e-7, keep=keep, seed=22)
    whole_ctx.close()
    cuts = sharding.slice_cuts(L, world)
    ctxs = [_native.Context(0) for _ in range(world)]
    try:
        params = []
        for r, c in enumerate(ctxs):
            c.load_tables(ts)
            params.append(_native.Params(map_position=[0, L], map_rate=[5e-7], model=cases.native_model(mdl),
                                         discrete_genome=True, keep=keep, start_time=-cases.FMAX,
                                         end_time=cases.FMAX, seed=22, slice_left=cuts[r], slice_right=cuts[r + 1]))
        # the all-gathers of a real run, played back serially
        first_phase = [c.generate_begin(p) for c, p in zip(ctxs, params)]
        all1 = [[f["num_sites"], f["num_mutations"], (f["next_allele"] - start) % 2**64, f["num_new_mutations"]]
                for f in first_phase]
        second = []
        for r, c in enumerate(ctxs):
            sb, mb, ab, _ = sharding.exclusive_bases(all1, r)
            second.append(c.generate_end(sb, mb, start + ab))
        all2 = [[f["ancestral_state_bytes"], f["site_metadata_bytes"], f["derived_state_bytes"],
                 f["mutation_metadata_bytes"]] for f in second]
        parts = []
        for r, c in enumerate(ctxs):
            c.shift_offsets(*sharding.exclusive_bases(all2, r))
            parts.append(c.fetch())
        merged = sharding.merge_slices(parts)
    finally:
        for c in ctxs:
            c.close()
    for tbl in whole:
        for col in whole[tbl]:
            assert np.array_equal(np.asarray(whole[tbl][col]), np.asarray(merged[tbl][col])), (tbl, col)
    assert sum(a[3] for a in all1) == cw["num_new_mutations"]


@pytest.mark.parametrize("begin", ["host", "device"])
@pytest.mark.parametrize("world", [2, 5])
@pytest.mark.parametrize("model,keep", [("infinite_alleles", False), ("jc69", True), ("multichar", False),
                                        ("infinite_alleles", True), ("jc69", False)])
def test_device_exchange_equals_whole(native_built, ts_small, world, model, keep, begin):
    device_exchange_case(ts_small, world, model, keep, begin, 5e-7)


@pytest.mark.parametrize("begin", ["host", "device"])
@pytest.mark.parametrize("model", ["infinite_alleles", "jc69", "multichar"])
def test_device_exchange_with_empty_slices(native_built, ts_small, model, begin):
    """A handful of mutations over eight slices: some ranks have nothing to write (and take the plain order of
    stages while their neighbours run the exchanges beside K4)."""
    device_exchange_case(ts_small, 8, model, False, begin, 1e-10)


def device_exchange_case(ts_small, world, model, keep, begin, rate):
    """The exchange kept on the device (mspb_exchange_* / _end_device / _finish_device): every
    rank is a context on this GPU and the all-gather is a device copy between their buffers.
    begin="device": the first stage returns without waiting (mspb_generate_begin_device) and the counts go
    from the device totals into the exchange."""
    import torch

    ts = ts_small
    L = ts["sequence_length"]
    mdl = cases.MODELS[model]
    whole_ctx = _native.Context(0)
    if keep:
        first, _ = run(whole_ctx, ts, model=mdl, rate=rate, seed=31)
        ts = cases.with_kept(ts, first)
    start = 77 if model == "infinite_alleles" else 0
    mdl = dict(mdl, start_allele=start) if model == "infinite_alleles" else mdl
    whole, cw = run(whole_ctx, ts, model=mdl, rate=rate, keep=keep, seed=32)
    whole_ctx.close()
    cuts = sharding.slice_cuts(L, world)
    ctxs = [_native.Context(0) for _ in range(world)]
    try:
        bufs = [c.exchange_tensors(world) for c in ctxs]

        def gather_all():
            # what NCCL does in a real run; here the ranks share the device, so wait and copy
            torch.cuda.synchronize()
            sends = torch.cat([b[0] for b in bufs])
            for b in bufs:
                b[1].copy_(sends)
            torch.cuda.synchronize()

        for r, c in enumerate(ctxs):
            c.load_tables(ts)
            first_stage = c.generate_begin if begin == "host" else c.generate_begin_device
            first_stage(_native.Params(
                map_position=[0, L], map_rate=[rate], model=cases.native_model(mdl), discrete_genome=True,
                keep=keep, start_time=-cases.FMAX, end_time=cases.FMAX, seed=32, slice_left=cuts[r],
                slice_right=cuts[r + 1]))
            c.exchange_post_counts()
        gather_all()
        for r, c in enumerate(ctxs):
            c.generate_end_device(r, world)
        gather_all()
        parts, infos = [], []
        for r, c in enumerate(ctxs):
            counts, totals, bases = c.generate_finish_device(r, world)
            infos.append((counts, totals, bases))
            parts.append(c.fetch())
        merged = sharding.merge_slices(parts)
    finally:
        for c in ctxs:
            c.close()
    for tbl in whole:
        for col in whole[tbl]:
            assert np.array_equal(np.asarray(whole[tbl][col]), np.asarray(merged[tbl][col])), (tbl, col)
    counts, totals, bases = infos[-1]
    assert totals["num_sites"] == cw["num_sites"] and totals["num_mutations"] == cw["num_mutations"]
    assert totals["num_new_mutations"] == cw["num_new_mutations"]
    assert totals["derived_state_bytes"] == cw["derived_state_bytes"]
    assert counts["next_allele"] == cw["next_allele"]
    assert bases["num_sites"] == sum(i[0]["num_sites"] for i in infos[:-1])


@pytest.mark.parametrize("why", ["dense", "collisions"])
def test_device_begin_hands_back_to_the_host(native_built, why):
    """What the host looks at between the stages of a call -- a dense genome's second sort, positions that are
    taken -- is raised through the exchanged counts when the first stage did not wait: every rank's
    generate_finish_device asks for the call to be made again (RetryWithHost), and sharding.run_slice_device then
    makes it with the host in the loop, all ranks together, with the result of the one-GPU call."""
    import torch

    world = 2
    if why == "dense":
        # a few positions and thousands of mutations on each: KEY32_POS reports a dense genome
        ts = cases.four_leaf_tree()
        kw = dict(model=cases.JC69, rate=5.0, discrete=True)
    else:
        # continuous genome under a discrete-looking rate: same-position draws are impossible natively, so take
        # positions that ARE taken: a kept layer drawn from the same seed (every new position collides)
        ts = cases.four_leaf_tree()
        kw = dict(model=cases.BINARY, rate=2.0, discrete=False)
    L = ts["sequence_length"]
    keep = why == "collisions"
    whole_ctx = _native.Context(0)
    if keep:
        first, _ = run(whole_ctx, ts, model=kw["model"], rate=kw["rate"], seed=61, discrete=False)
        ts = cases.with_kept(ts, first)
    seed = 61 if keep else 62
    whole, cw = run(whole_ctx, ts, model=kw["model"], rate=kw["rate"], keep=keep, seed=seed, discrete=kw["discrete"])
    whole_ctx.close()
    cuts = sharding.slice_cuts(L, world, kw["discrete"])
    ctxs = [_native.Context(0) for _ in range(world)]
    try:
        bufs = [c.exchange_tensors(world) for c in ctxs]

        def gather_all():
            torch.cuda.synchronize()
            sends = torch.cat([b[0] for b in bufs])
            for b in bufs:
                b[1].copy_(sends)
            torch.cuda.synchronize()

        def params(r):
            return _native.Params(
                map_position=[0, L], map_rate=[kw["rate"]], model=cases.native_model(kw["model"]),
                discrete_genome=kw["discrete"], keep=keep, start_time=-cases.FMAX, end_time=cases.FMAX, seed=seed,
                slice_left=cuts[r], slice_right=cuts[r + 1])

        for r, c in enumerate(ctxs):
            c.load_tables(ts)
            c.generate_begin_device(params(r))
            c.exchange_post_counts()
        gather_all()
        for r, c in enumerate(ctxs):
            c.generate_end_device(r, world)
        gather_all()
        asked = 0
        for r, c in enumerate(ctxs):
            with pytest.raises(_native.RetryWithHost):
                c.generate_finish_device(r, world)
            asked += 1
        assert asked == world
        # the call again, with the host between the stages
        try:
            for r, c in enumerate(ctxs):
                c.generate_begin(params(r))
                c.exchange_post_counts()
        except _native.LibraryError as e:
            # invariant slices cannot move a colliding position: the one-call path owns that case
            assert why == "collisions" and e.status == -1004
            return
        gather_all()
        for r, c in enumerate(ctxs):
            c.generate_end_device(r, world)
        gather_all()
        parts = []
        for r, c in enumerate(ctxs):
            c.generate_finish_device(r, world)
            parts.append(c.fetch())
        merged = sharding.merge_slices(parts)
    finally:
        for c in ctxs:
            c.close()
    for tbl in whole:
        for col in whole[tbl]:
            assert np.array_equal(np.asarray(whole[tbl][col]), np.asarray(merged[tbl][col])), (tbl, col)


@pytest.mark.parametrize("discrete", [True, False])
@pytest.mark.parametrize("keep", [False, True])
def test_both_placement_kernels_write_the_same_records(native_built, ts_small, discrete, keep):
    """K2 as one block per 256 edges and as one thread per mutation (the form sparse slices take): whole calls and
    clipped slices, first calls and calls on warm buffers (placed ahead of the K1 read-back, the sparse launch sized
    from the call before: smaller, equal and larger counts), with and without kept rows."""
    ts = ts_small
    L = ts["sequence_length"]
    if keep:
        with _native.Context(0) as c0:
            first, _ = run(c0, ts, model=cases.JC69, rate=2e-7, seed=71, discrete=discrete)
        ts = cases.with_kept(ts, first)
    plan = [(1e-8, 0.0, 0.0, 0), (3e-7, 0.0, 0.0, 0), (3e-7, 0.0, 0.0, 0), (2e-8, 0.0, 0.0, 0),
            (3e-7, 2e4, 9e4, 3), (3e-7, 2e4, 9e4, 3), (1e-6, 2e4, 9e4, 3)]
    results = {}
    for mode in (1, 2):
        out = []
        with _native.Context(0) as c:
            c.force_k2(mode)
            c.load_tables(ts)
            for k, (rate, lo, hi, stream) in enumerate(plan):
                cnt = c.generate(_native.Params(
                    map_position=[0, L], map_rate=[rate], model=cases.native_model(cases.JC69), discrete_genome=discrete,
                    keep=keep, start_time=-cases.FMAX, end_time=cases.FMAX, seed=80 + k, slice_left=lo, slice_right=hi,
                    slice_stream=stream))
                out.append((cnt["num_new_mutations"], c.fetch()))
        results[mode] = out
    assert [n for n, _ in results[1]] == [n for n, _ in results[2]]
    assert results[1][1][0] > 5 * results[1][0][0] > 0  # the second call outgrows what the first one sized
    for (n, a), (_, b) in zip(results[1], results[2]):
        for tbl in a:
            for col in a[tbl]:
                assert np.array_equal(np.asarray(a[tbl][col]), np.asarray(b[tbl][col])), (tbl, col, n)


def test_device_tensors_alias_the_result(ctx, ts_small):
    """the result handed over as torch tensors on the device, no copy."""
    import torch

    host, c = run(ctx, ts_small, model=cases.INF_ALLELES, rate=1e-6, seed=5)
    dev = ctx.device_tensors()
    torch.cuda.synchronize()
    for tbl in host:
        for col in host[tbl]:
            t = dev[tbl][col]
            assert t.is_cuda and t.numel() == len(host[tbl][col]), (tbl, col)
            got = t.cpu().numpy()
            want = np.asarray(host[tbl][col])
            assert np.array_equal(got.view(np.uint8), want.astype(got.dtype).view(np.uint8)), (tbl, col)
    # a consumer on the GPU: per-site mutation counts straight from the resident columns
    counts = torch.bincount(dev["mutations"]["site"].long(), minlength=c["num_sites"])
    assert int(counts.sum()) == c["num_mutations"] and int(counts.min()) >= 1


def test_clipped_slices(native_built, ts_small):
    """slice_stream != 0: edges are cut to the slice and drawn on the slice's own stream."""
    ts = ts_small
    L = ts["sequence_length"]
    world, rate = 4, 1e-6
    cuts = sharding.slice_cuts(L, world)
    e, nt = ts["edges"], ts["nodes"]["time"]
    expect = ((e["right"] - e["left"]) * (nt[e["parent"]] - nt[e["child"]])).sum() * rate
    ctx = _native.Context(0)
    ctx.load_tables(ts)
    total, parts, streams = 0, [], []
    try:
        for r in range(world):
            p = _native.Params(map_position=[0, L], map_rate=[rate], model=cases.native_model(cases.JC69),
                               discrete_genome=True, keep=False, start_time=-cases.FMAX, end_time=cases.FMAX,
                               seed=9, slice_left=cuts[r], slice_right=cuts[r + 1], slice_stream=r + 1)
            c = ctx.generate(p)
            res = ctx.fetch()
            pos = res["sites"]["position"]
            assert len(pos) > 0 and pos.min() >= cuts[r] and pos.max() < cuts[r + 1]
            table_checks.check_result(ts, res, discrete=True)
            total += c["num_new_mutations"]
            parts.append(res)
            streams.append(res["mutations"]["time"][:50].copy())
    finally:
        ctx.close()
    assert abs(total - expect) < 5 * np.sqrt(expect)
    assert not np.array_equal(streams[0], streams[1])


def test_python_surface_layers_and_provenance(native_built, ts_small):
    ts = mtables.TableCollection.fromdict(ts_small).tree_sequence()
    a = msp.sim_mutations(ts, rate=2e-7, random_seed=3, model="jc69")
    assert a.num_mutations > 0 and a.num_provenances == ts.num_provenances + 1
    b = msp.sim_mutations(a, rate=2e-7, random_seed=4, model=msp.HKY(kappa=2.0))  # keep=True by default
    assert b.num_mutations > a.num_mutations and b.num_sites >= a.num_sites
    c = msp.sim_mutations(a, rate=0, random_seed=4)
    assert c.num_mutations == a.num_mutations
    d = msp.sim_mutations(a, rate=0, random_seed=4, keep=False, record_provenance=False)
    assert d.num_mutations == 0 and d.num_sites == 0 and d.num_provenances == a.num_provenances
    assert ts.num_mutations == 0  # the input is never modified
    ia = msp.InfiniteAlleles()
    e = msp.sim_mutations(ts, rate=2e-7, random_seed=5, model=ia)
    assert ia.next_allele == e.num_sites + e.num_mutations
    f = msp.mutate(ts, rate=2e-7, random_seed=6)
    assert set(f.alleles()[0]) == {"0"} and set(f.alleles()[1]) <= {"0", "1"}


def test_slim_model_native_stream(ctx, ts_small):
    """SLiMMutationModel on the native stream (tests/test_mutations.py:1554-1696): empty
    ancestral states, ids numbered in table order, stacked derived states = the chain of ids up
    to the root, 17 bytes of metadata per stacked id; a second layer stacks onto kept states."""
    import struct

    L = ts_small["sequence_length"]
    m1 = {"kind": "slim", "type": 2, "next_id": 100, "slim_generation": 7}
    first, c1 = run(ctx, ts_small, model=m1, rate=2e-6, seed=41)
    M = c1["num_mutations"]
    assert M > 500 and c1["next_allele"] == 100 + M and c1["ancestral_state_bytes"] == 0
    table_checks.check_result(ts_small, first, discrete=True)
    ds = table_checks.ragged(first["mutations"]["derived_state"], first["mutations"]["derived_state_offset"])
    md = table_checks.ragged(first["mutations"]["metadata"], first["mutations"]["metadata_offset"])
    par = first["mutations"]["parent"]
    t = first["mutations"]["time"]
    for j in range(M):
        ids = [int(x) for x in bytes(ds[j]).decode().split(",")]
        assert ids[-1] == 100 + j  # own id last, numbered in row order
        assert len(bytes(md[j])) == 17 * len(ids)
        if par[j] >= 0:
            assert bytes(ds[j]).decode().rsplit(",", 1)[0] == bytes(ds[par[j]]).decode()
            assert bytes(md[j])[:-17] == bytes(md[par[j]])
        else:
            assert len(ids) == 1
        ty, sel, sub, origin, nuc = struct.unpack("<ifiib", bytes(md[j])[-17:])
        assert (ty, sel, sub, nuc) == (2, 0.0, -1, -1) and origin == 7 - int(t[j])
    assert (par >= 0).sum() > 0  # the rate stacks some mutations
    # second layer on top: ids continue, stacked onto kept states where a kept mutation is above
    ts2 = cases.with_kept(ts_small, first)
    m2 = dict(m1, next_id=c1["next_allele"], type=5)
    second, c2 = run(ctx, ts2, model=m2, rate=2e-6, seed=42, keep=True)
    assert c2["next_allele"] == c1["next_allele"] + c2["num_new_mutations"]
    ds2 = table_checks.ragged(second["mutations"]["derived_state"], second["mutations"]["derived_state_offset"])
    md2 = table_checks.ragged(second["mutations"]["metadata"], second["mutations"]["metadata_offset"])
    par2 = second["mutations"]["parent"]
    seen_new = 0
    for j in range(c2["num_mutations"]):
        ids = bytes(ds2[j]).decode().split(",")
        assert len(bytes(md2[j])) == 17 * len(ids)
        if int(ids[-1]) >= c1["next_allele"]:
            seen_new += 1
            if par2[j] >= 0:
                assert ",".join(ids[:-1]) == bytes(ds2[par2[j]]).decode()
    assert seen_new == c2["num_new_mutations"]
    # bad parameters surface the reference's error (lib/mutgen.c:326-343)
    with pytest.raises(_native.LibraryError, match="nonnegative"):
        run(ctx, ts_small, model=dict(m1, type=-1), rate=1e-6, seed=1)


def test_slim_model_python_surface(native_built, ts_small):
    ts = mtables.TableCollection.fromdict(ts_small).tree_sequence()
    model = msp.SLiMMutationModel(type=1, next_id=5)
    a = msp.sim_mutations(ts, rate=2e-7, random_seed=8, model=model)
    assert a.num_mutations > 0 and model.next_id == 5 + a.num_mutations
    b = msp.sim_mutations(a, rate=2e-7, random_seed=9, model=model)
    assert model.next_id == 5 + b.num_mutations



@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_sharded_table_upload(native_built, ts_small, world):
    """mspb_ctx_load_tables_begin / _end: every "rank" (contexts on one device here) uploads its
    1 / world share of the node and edge columns, the shares are exchanged with device copies (the
    stand-in for the in-place NCCL all-gather), and each context then overlays exactly what a
    context loaded the ordinary way does, kept rows included."""
    import torch

    first, _ = run(_native.Context(0), ts_small, model=cases.JC69, rate=2e-7, seed=3)
    ts = cases.with_kept(ts_small, first)
    ref = _native.Context(0)
    want, _ = run(ref, ts, model=cases.JC69, rate=3e-7, keep=True, seed=9)
    dev = torch.device("cuda", 0)
    ctxs = [_native.Context(0) for _ in range(world)]
    shards = [c.load_tables_begin(ts, r, world) for r, c in enumerate(ctxs)]
    views = [[torch.as_tensor(_native._DeviceArray(ptr, per * world, "|u1"), device=dev) if per else None
              for ptr, per in sh] for sh in shards]
    torch.cuda.synchronize()
    for k in range(_native.NUM_TABLE_SHARDS):
        per = shards[0][k][1]
        assert all(sh[k][1] == per for sh in shards)
        for r in range(world):
            for q in range(world):
                if q != r and per:
                    views[r][k][q * per:(q + 1) * per].copy_(views[q][k][q * per:(q + 1) * per])
    torch.cuda.synchronize()
    for r, c in enumerate(ctxs):
        c.load_tables_end()
        got, _ = run(c, ts, model=cases.JC69, rate=3e-7, keep=True, seed=9, load=False)
        for tbl in want:
            for col in want[tbl]:
                assert np.array_equal(np.asarray(want[tbl][col]), np.asarray(got[tbl][col])), (r, tbl, col)
    # the hand-off is checked: _end without _begin is refused, and bad rows are still rejected
    with pytest.raises(_native.LibraryError):
        ctxs[0].load_tables_end()
    bad = dict(ts, edges=dict(ts["edges"], child=np.array(ts["edges"]["child"], copy=True)))
    bad["edges"]["child"][5] = len(ts["nodes"]["time"]) + 7
    ctxs[0].load_tables_begin(bad, 0, 1)
    with pytest.raises(_native.LibraryError):
        ctxs[0].load_tables_end()
    for c in ctxs + [ref]:
        c.close()
