"""GPU: every error path of the reference's C tests (lib/tests/test_mutations.c:141-222, 586-592,
1152-1265) and of its Python-C tests (tests/test_python_c.py:2847-3075), driven through the C ABI
(`mspb_ctx_load_tables` + `mspb_generate`, and once each through the one-shot `mspb_sim_mutations`).
For every bad input the status must equal the one the reference returns for the SAME input: the
compiled reference where it was built (oracle/_ref), else the C restatement, which
tests/test_oracle_fuzz.py holds to it.  The codes are the reference's own numbering (lib/util.h:48-136),
tskit codes with bit 13 flipped (lib/util.c:337)."""
import numpy as np
import pytest

import cases
from msprime_b200 import _native

pytestmark = pytest.mark.gpu

FMAX = cases.FMAX
TSK = lambda code: code ^ (1 << 13)  # noqa: E731  msp_set_tsk_error
UNKNOWN_TIME = np.frombuffer(np.array([0x7FF874736B697421], dtype=np.uint64).tobytes(), dtype=np.float64)[0]  # TSK_UNKNOWN_TIME


def reference_status(oracle_mod, ts, **kw):
    which = "ref" if oracle_mod.have("ref") else "port"
    kw.setdefault("seed", 1)
    try:
        oracle_mod.sim_mutations(ts, which=which, **kw)
    except oracle_mod.OracleError as e:
        return e.status
    return 0


def device_status(ts, *, oneshot=False, **kw):
    model = kw.pop("model")
    okw = dict(map_position=kw["map_position"], map_rate=kw["map_rate"], model=cases.native_model(model),
               discrete_genome=kw.get("discrete_genome", True), keep=kw.get("keep", True),
               start_time=kw.get("start_time", -FMAX), end_time=kw.get("end_time", FMAX), seed=kw.get("seed", 1))
    try:
        params = _native.Params(**okw)
        if oneshot:
            _native.sim_mutations_oneshot(ts, params)
        else:
            with _native.Context(0) as ctx:
                ctx.load_tables(ts)
                ctx.generate(params)
    except (_native.LibraryError, _native.InputError) as e:
        return e.status
    return 0


def with_sites(ts, position, anc, muts=(), times=None):
    """`ts` with hand-made existing rows: sites at `position` with ancestral states `anc`, mutations
    (site, node, derived)."""
    d = dict(ts)
    a = [x.encode() for x in anc]
    d["sites"] = {"position": np.array(position, dtype=np.float64),
                  "ancestral_state": np.frombuffer(b"".join(a), dtype=np.int8),
                  "ancestral_state_offset": np.concatenate([[0], np.cumsum([len(x) for x in a])]).astype(np.uint64),
                  "metadata": np.zeros(0, dtype=np.int8), "metadata_offset": np.zeros(len(a) + 1, dtype=np.uint64)}
    dr = [m[2].encode() for m in muts]
    nt = ts["nodes"]["time"]
    d["mutations"] = {"site": np.array([m[0] for m in muts], dtype=np.int32),
                      "node": np.array([m[1] for m in muts], dtype=np.int32),
                      "parent": np.full(len(muts), -1, dtype=np.int32),
                      "time": np.array([nt[m[1]] + 0.25 for m in muts] if times is None else times, dtype=np.float64),
                      "derived_state": np.frombuffer(b"".join(dr), dtype=np.int8),
                      "derived_state_offset": np.concatenate([[0], np.cumsum([len(x) for x in dr])]).astype(np.uint64),
                      "metadata": np.zeros(0, dtype=np.int8),
                      "metadata_offset": np.zeros(len(muts) + 1, dtype=np.uint64)}
    return d


def edit(ts, table, column, index, value):
    d = dict(ts)
    d[table] = dict(ts[table])
    col = np.array(ts[table][column])
    col[index] = value
    d[table][column] = col
    return d


def _cases():
    ts = cases.four_leaf_tree()  # L = 10, lib/tests/test_mutations.c:22-94
    ok = dict(map_position=[0, 10], map_rate=[0.5], model=cases.JC69)
    out = []

    def add(name, want, ts_=ts, **kw):
        out.append(pytest.param(ts_, {**ok, **kw}, want, id=name))

    # --- mutgen_set_rate_map / rate_map_alloc (lib/mutgen.c:796-813, lib/rate_map.c:53-88)
    add("map-length-differs", -44, map_position=[0, 11])                      # test_mutations.c:163-166
    add("map-length-shorter", -44, map_position=[0, 5, 9], map_rate=[1, 1])
    add("negative-rate", -43, map_rate=[-1])                                  # test_mutations.c:168-174
    add("nan-rate", -43, map_rate=[np.nan])
    add("infinite-rate", -43, map_rate=[np.inf])
    add("map-start-nonzero", -46, map_position=[1, 10])
    add("map-positions-unsorted", -48, map_position=[0, 6, 4, 10], map_rate=[1, 1, 1])
    add("map-positions-repeat", -48, map_position=[0, 4, 4, 10], map_rate=[1, 1, 1])
    add("map-position-nonfinite", -47, map_position=[0, np.inf], map_rate=[1])
    add("map-position-nan", -47, map_position=[0, np.nan, 10], map_rate=[1, 1])
    # --- mutgen_set_time_interval (lib/mutgen.c:781-794)
    add("end-before-start", -4, start_time=2.0, end_time=1.0)
    add("empty-window-is-fine", 0, start_time=1.0, end_time=1.0)
    # --- matrix_mutation_model_alloc (lib/mutgen.c:501-524); test_mutations.c:1152-1265
    add("one-allele", -52, model={"kind": "matrix", "alleles": ["A"], "root_distribution": [1.0],
                                  "transition_matrix": [[1.0]]})
    add("root-does-not-sum-to-one", -53, model={**cases.BINARY, "root_distribution": [0.5, 0.6]})
    add("root-negative", -53, model={**cases.BINARY, "root_distribution": [1.5, -0.5]})
    add("matrix-row-does-not-sum-to-one", -54, model={**cases.BINARY, "transition_matrix": [[0.0, 1.0], [0.5, 0.6]]})
    add("matrix-row-negative", -54, model={**cases.BINARY, "transition_matrix": [[1.5, -0.5], [1.0, 0.0]]})
    add("matrix-sum-within-tolerance", 0, model={**cases.BINARY, "root_distribution": [0.5, 0.5 + 1e-13]})
    # --- slim_mutation_model_alloc (lib/mutgen.c:552-589)
    add("slim-negative-type", -56, model={"kind": "slim", "type": -1, "next_id": 0, "slim_generation": 1})
    # --- mutgen_initialise_sites (lib/mutgen.c:974-1025); test_mutations.c:586-592
    kept = with_sites(ts, [1.0, 3.0], ["A", "C"], [(0, 0, "C"), (1, 2, "G")])
    add("kept-rows-are-fine", 0, ts_=kept)
    add("unknown-time-under-keep", -69, ts_=edit(kept, "mutations", "time", 1, UNKNOWN_TIME))
    add("unknown-time-without-keep-is-fine", 0, keep=False, ts_=edit(kept, "mutations", "time", 1, UNKNOWN_TIME))
    add("duplicate-site-position", -21, ts_=with_sites(ts, [1.0, 1.0], ["A", "C"]))          # lib/mutgen.c:877
    # --- an allele the model does not know (lib/mutgen.c:225-230; test_mutations.c:176-201): an error
    # only when a NEW mutation has to read it
    foreign = with_sites(ts, [2.0, 4.5], ["0", "1"], [(0, 1, "1")])
    add("foreign-alleles-rate-zero", 0, ts_=foreign, map_rate=[0.0])
    add("foreign-alleles-under-new-mutations", -51, ts_=foreign, map_rate=[20.0])
    add("foreign-alleles-not-kept", 0, ts_=foreign, map_rate=[20.0], keep=False)
    add("foreign-allele-off-the-integers", 0, ts_=with_sites(ts, [4.5], ["xyz"]), map_rate=[20.0])
    add("foreign-alleles-continuous-genome", 0, ts_=foreign, map_rate=[20.0], discrete_genome=False)
    # --- tsk_table_collection_check_integrity(tables, 0) (lib/mutgen.c:1316; test_mutations.c:176-179)
    add("edge-beyond-sequence-length", TSK(-309), ts_={**ts, "sequence_length": 0.1}, map_position=[0, 0.1])
    add("edge-parent-out-of-bounds", TSK(-202), ts_=edit(ts, "edges", "parent", 2, 99))
    add("edge-child-out-of-bounds", TSK(-202), ts_=edit(ts, "edges", "child", 2, -7))
    add("edge-null-parent", TSK(-300), ts_=edit(ts, "edges", "parent", 0, -1))
    add("edge-null-child", TSK(-301), ts_=edit(ts, "edges", "child", 0, -1))
    add("edge-left-not-below-right", TSK(-307), ts_=edit(ts, "edges", "left", 3, 10.0))
    add("edge-left-negative", TSK(-310), ts_=edit(ts, "edges", "left", 3, -1.0))
    add("edge-nonfinite", TSK(-211), ts_=edit(ts, "edges", "right", 3, np.inf))
    add("parent-not-older-than-child", TSK(-306), ts_=edit(ts, "nodes", "time", 4, 0.0))
    add("node-time-nonfinite", TSK(-210), ts_=edit(ts, "nodes", "time", 6, np.inf))
    add("site-position-out-of-range", TSK(-402), ts_=with_sites(ts, [1.0, 10.0], ["A", "C"]))
    add("site-position-negative", TSK(-402), ts_=with_sites(ts, [-1.0, 3.0], ["A", "C"]))
    add("mutation-site-out-of-bounds", TSK(-205), ts_=edit(kept, "mutations", "site", 1, 2))
    add("mutation-node-out-of-bounds", TSK(-202), ts_=edit(kept, "mutations", "node", 1, 7))
    add("mutation-time-nonfinite", TSK(-210), ts_=edit(kept, "mutations", "time", 0, np.inf))
    add("mutation-younger-than-node", TSK(-506), ts_=edit(kept, "mutations", "time", 1, -1.0))
    # --- mutgen_alloc (lib/mutgen.c:760); test_mutations.c:213-216
    add("sequence-length-negative", -4, ts_={**ts, "sequence_length": -1.0})
    return out


@pytest.mark.parametrize("ts,kw,want", _cases())
def test_error_table_through_the_c_abi(native_built, oracle_mod, ts, kw, want):
    assert reference_status(oracle_mod, ts, **{k: v for k, v in kw.items() if k != "seed"}) == want, \
        "the reference disagrees with this table"
    assert device_status(ts, **kw) == want
    assert device_status(ts, oneshot=True, **kw) == want


def test_one_shot_entry_point_equals_the_context_calls(native_built, oracle_mod):
    """`mspb_sim_mutations` (host tables in, malloc'd host columns out: what INTEGRATION.md's stub binds)
    against load_tables + generate + fetch, both against the oracle under one replay stream; then the native
    stream, where the two must agree bit for bit for one seed."""
    import test_gpu_replay as replay

    ts = cases.smc_ts(50, 1e5, 1)
    L = ts["sequence_length"]
    stream = np.random.default_rng(3).random(400_000)
    first = None
    for keep, model in ((False, cases.JC69), (True, cases.JC69), (False, cases.INF_ALLELES), (False, cases.SLIM)):
        base = ts if not keep else cases.with_kept(ts, first)
        want = oracle_mod.sim_mutations(base, which="port", map_position=[0, L], map_rate=[2e-6], model=model,
                                        discrete_genome=True, keep=keep, stream=stream)
        params = _native.Params(map_position=[0, L], map_rate=[2e-6], model=cases.native_model(model),
                                discrete_genome=True, keep=keep, start_time=-FMAX, end_time=FMAX, replay_stream=stream)
        got, counts = _native.sim_mutations_oneshot(base, params)
        assert not replay.describe_diff(want, got)
        assert counts["num_mutations"] == len(want["mutations"]["site"])
        if first is None:
            first = want
    params = _native.Params(map_position=[0, L], map_rate=[2e-6], model=cases.native_model(cases.JC69),
                            discrete_genome=False, keep=False, start_time=-FMAX, end_time=FMAX, seed=77)
    one, _ = _native.sim_mutations_oneshot(ts, params)
    with _native.Context(0) as ctx:
        ctx.load_tables(ts)
        ctx.generate(params)
        two = ctx.fetch()
    assert not replay.describe_diff(one, two) and len(one["mutations"]["site"]) > 100


def test_one_shot_calls_share_a_context_and_survive_errors(native_built):
    """The one-shot call keeps one context per device: tables of different sizes one after the other, a failing
    call in between (which drops the context), mspb_oneshot_release, and the same answers as a fresh context."""
    import test_gpu_replay as replay

    small, large = cases.smc_ts(20, 2e4, 5), cases.smc_ts(60, 2e5, 6)

    def params(ts, rate=2e-6, seed=9):
        L = ts["sequence_length"]
        return _native.Params(map_position=[0, L], map_rate=[rate], model=cases.native_model(cases.JC69),
                              discrete_genome=True, keep=False, start_time=-FMAX, end_time=FMAX, seed=seed)

    want = {}
    for name, ts in (("small", small), ("large", large)):
        with _native.Context(0) as ctx:
            ctx.load_tables(ts)
            ctx.generate(params(ts))
            want[name] = ctx.fetch()
    order = ["small", "large", "small", "bad", "large", "release", "small"]
    for step in order:
        if step == "bad":
            with pytest.raises((_native.LibraryError, _native.InputError)):
                _native.sim_mutations_oneshot(small, params(small, rate=-1.0))
        elif step == "release":
            _native.lib().mspb_oneshot_release(-1)
            _native.lib().mspb_oneshot_release(0)  # nothing left: a no-op
        else:
            ts = small if step == "small" else large
            got, _ = _native.sim_mutations_oneshot(ts, params(ts))
            assert not replay.describe_diff(want[step], got), step
    _native.lib().mspb_oneshot_release(-1)


def test_pageable_and_pinned_tables_load_alike(native_built):
    """Columns in pageable memory above 4 MB are staged through the pinned pool by host threads; the same columns
    in page-locked memory go to the device directly.  Same index, same overlay."""
    import test_gpu_replay as replay
    from msprime_b200 import synth

    ts = synth.native_treeseq(60_000, 3e7, 1e-8, 1e4, 4)  # 8e5 edges: 6 MB per float64 column
    assert len(ts["edges"]["left"]) * 8 > (4 << 20)
    L = ts["sequence_length"]
    pinned = {k: (dict(v) if isinstance(v, dict) else v) for k, v in ts.items()}
    hold = []
    for tbl, cols in (("edges", ("left", "right", "parent", "child")), ("nodes", ("time", "flags"))):
        for col in cols:
            a = ts[tbl][col]
            b = _native.pinned_empty(len(a), a.dtype)
            b[:] = a
            pinned[tbl][col] = b
            hold.append(b)
    prm = _native.Params(map_position=[0, L], map_rate=[1e-8], model=cases.native_model(cases.JC69),
                         discrete_genome=True, keep=False, start_time=-FMAX, end_time=FMAX, seed=3)
    res = []
    for tables in (ts, pinned, ts):
        with _native.Context(0) as ctx:
            ctx.load_tables(tables)
            ctx.generate(prm)
            res.append(ctx.fetch())
    assert len(res[0]["mutations"]["site"]) > 10_000
    assert not replay.describe_diff(res[0], res[1]) and not replay.describe_diff(res[0], res[2])


def hand_kept(ts, rows):
    """``ts`` with ONE existing site at position 3 (ancestral state A) and the given existing mutations
    ``(node, time, derived_state)`` in the order given."""
    d = dict(ts)
    d["sites"] = {"position": np.array([3.0]), "ancestral_state": np.frombuffer(b"A", dtype=np.uint8).copy(),
                  "ancestral_state_offset": np.array([0, 1], dtype=np.uint64),
                  "metadata": np.zeros(0, dtype=np.uint8), "metadata_offset": np.zeros(2, dtype=np.uint64)}
    der = "".join(r[2] for r in rows).encode()
    d["mutations"] = {"site": np.zeros(len(rows), dtype=np.int32), "node": np.array([r[0] for r in rows], dtype=np.int32),
                      "parent": np.full(len(rows), -1, dtype=np.int32),
                      "time": np.array([r[1] for r in rows], dtype=np.float64),
                      "derived_state": np.frombuffer(der, dtype=np.uint8).copy(),
                      "derived_state_offset": np.arange(len(rows) + 1, dtype=np.uint64),
                      "metadata": np.zeros(0, dtype=np.uint8), "metadata_offset": np.zeros(len(rows) + 1, dtype=np.uint64)}
    return d


def test_existing_mutations_out_of_order_under_a_new_one_are_refused(native_built, oracle_mod):
    """Existing mutations of a site that are not listed oldest first, with a new mutation on the site: the
    reference's within-site comparator (existing rows by id, the rest by time, lib/mutgen.c:36-49) is then not a
    total order, what it returns is undefined and its debug build asserts (lib/mutgen.c:1200).  The device refuses
    such a call with MSPB_ERR_KEPT_MUTATIONS_UNSORTED; the same rows oldest first, or out of order but with no new
    mutation on their site, go through (and match the reference's algorithm under a shared stream)."""
    import test_gpu_replay as replay

    tree = cases.four_leaf_tree()
    young, old = (0, 0.5, "C"), (4, 1.5, "G")  # node 0 sits below node 4
    sorted_ts, unsorted_ts = hand_kept(tree, [old, young]), hand_kept(tree, [young, old])

    def status(ts, rate, seed=3):
        with _native.Context(0) as ctx:
            try:
                ctx.load_tables(ts)
                ctx.generate(_native.Params(map_position=[0, 10.0], map_rate=[rate], model=cases.native_model(cases.JC69),
                                            discrete_genome=True, keep=True, start_time=-FMAX, end_time=FMAX, seed=seed))
            except (_native.LibraryError, _native.InputError) as e:
                return e.status
        return 0

    assert status(unsorted_ts, 5.0) == -1012  # every position takes new mutations at this rate
    assert status(unsorted_ts, 0.0) == 0      # none lands on the site: the rows are kept as they are
    assert status(sorted_ts, 5.0) == 0
    # and the accepted forms agree with the reference's algorithm
    stream = np.random.default_rng(8).random(200_000)
    for ts, rate in ((sorted_ts, 5.0), (unsorted_ts, 0.0)):
        want = oracle_mod.sim_mutations(ts, which="port", map_position=[0, 10.0], map_rate=[rate], model=cases.JC69,
                                        discrete_genome=True, keep=True, stream=stream)
        with _native.Context(0) as ctx:
            got = replay.gpu_run(ctx, ts, model=cases.JC69, stream=stream, map_position=[0, 10.0], map_rate=[rate],
                                 discrete=True, keep=True)
        assert not replay.describe_diff(want, got)


def test_too_many_new_mutations_is_an_error_not_a_wrap(native_built):
    """A rate that asks for more new mutations than 31-bit ids can name (a mistyped rate) must fail with
    MSPB_ERR_TOO_MANY_ROWS before anything is placed; the count must not wrap around 2^32."""
    ts = cases.four_leaf_tree(L=1e6)
    with _native.Context(0) as ctx:
        ctx.load_tables(ts)
        for rate in (400.0, 5e4, 1e12):  # 9 branch units x 1e6 bp x rate: 3.6e9, 4.5e11, 9e18 expected
            with pytest.raises(_native.LibraryError) as e:
                ctx.generate(_native.Params(map_position=[0, 1e6], map_rate=[rate], model=cases.native_model(cases.BINARY),
                                            discrete_genome=True, keep=False, start_time=-FMAX, end_time=FMAX, seed=5))
            assert e.value.status == -1005, rate
        # and the context is still usable
        c = ctx.generate(_native.Params(map_position=[0, 1e6], map_rate=[1e-5], model=cases.native_model(cases.BINARY),
                                        discrete_genome=True, keep=False, start_time=-FMAX, end_time=FMAX, seed=5))
        assert 20 < c["num_mutations"] < 200
