"""What mutations MEAN under a model, after the reference's TestFiniteSites / TestInterval
(tests/test_mutations.py:837-1229): parents found on the tree, times on the branch, every transition
possible under the matrix, one-way / flip-flop / silent / circular models, integer sites, windows.
These are invariants of any correct stream, so the same bodies run twice through the public
`sim_mutations`: on the CPU with the device call answered by the oracle stand-in
(tests/oracle_engine.py; `-m "not gpu"`), and on the B200 with the native Philox stream (`-m gpu`).
The second half does the same for keep=True over hand-made existing tables (the reference's TestKeep)."""
import numpy as np
import pytest

import cases
import msprime_b200 as msp
from msprime_b200 import mutations, tables as mtables

NUC = ["A", "C", "G", "T"]

def _engine(request, monkeypatch):
    if request.param == "oracle":
        import oracle_engine

        eng = oracle_engine.OracleContext()
        monkeypatch.setattr(mutations, "_context", lambda device: eng)
    else:
        request.getfixturevalue("native_built")
    monkeypatch.setattr(mutations, "_tables", mtables)
    return request.param


@pytest.fixture(params=["oracle", pytest.param("device", marks=[pytest.mark.gpu])])
def engine(request, monkeypatch):
    return _engine(request, monkeypatch)


def small_ts(seed=1, samples=10, L=10.0):
    """Like sim_ancestry(10, sequence_length=10, recombination_rate=1): times of order one, several trees."""
    return cases.smc_ts(samples, L, seed, 0.5, 0.5)


def overlay(d, model, rate, discrete, seed=42, **kw):
    ts = mtables.TableCollection.fromdict(d).tree_sequence()
    return msp.sim_mutations(ts, rate=rate, random_seed=seed, model=model, discrete_genome=discrete, **kw)


def rows(out):
    t = out.tables
    anc, der = out.alleles()
    return (np.asarray(t.sites.position), anc, np.asarray(t.mutations.site).astype(np.int64),
            np.asarray(t.mutations.node).astype(np.int64), np.asarray(t.mutations.parent).astype(np.int64),
            np.asarray(t.mutations.time), der)


def parent_array(d, x):
    e = d["edges"]
    parent = np.full(len(d["nodes"]["time"]), -1, dtype=np.int64)
    cover = (e["left"] <= x) & (x < e["right"])
    parent[e["child"][cover]] = e["parent"][cover]
    return parent


def verify_mutations(d, out, discrete, model=None):
    """TestFiniteSites.verify_mutations (tests/test_mutations.py:844-873) on plain columns."""
    pos, anc, site, node, par, time, der = rows(out)
    node_time = d["nodes"]["time"]
    first = np.searchsorted(site, np.arange(len(pos) + 1))
    for j, x in enumerate(pos):
        tree = parent_array(d, x)
        if not discrete:
            assert first[j + 1] - first[j] == 1
        seen = {}
        for k in range(first[j], first[j + 1]):
            n = node[k]
            assert node_time[n] <= time[k] < node_time[tree[n]]
            while n != -1 and n not in seen:
                n = tree[n]
            if n == -1:
                assert par[k] == -1
                inherited = anc[j]
            else:
                assert par[k] == seen[n] and time[k] < time[seen[n]]
                inherited = der[seen[n]]
            if model is not None:
                a, b = model.alleles.index(inherited), model.alleles.index(der[k])
                assert model.transition_matrix[a][b] > 0
            seen[node[k]] = k  # last: parents are listed before their children


def matrix_model(alleles, matrix, root=None):
    root = [1 / len(alleles)] * len(alleles) if root is None else root
    return msp.MatrixMutationModel(alleles, root, matrix)


def max_per_site(out):
    _, _, site, *_ = rows(out)
    return np.bincount(site).max()


@pytest.mark.parametrize("discrete", [True, False])
def test_default_binary_and_nucleotide_alphabets(engine, discrete):
    d = small_ts()
    out = overlay(d, matrix_model(["0", "1"], [[0, 1], [1, 0]], [1, 0]), 1.0, discrete)
    _, anc, _, _, _, _, der = rows(out)
    assert out.num_sites > 0 and set(anc) == {"0"} and set(der) <= {"0", "1"}
    verify_mutations(d, out, discrete)
    model = matrix_model(NUC, [[0.25] * 4] * 4)
    out = overlay(d, model, 1.0, discrete)
    _, anc, _, _, _, _, der = rows(out)
    assert set(anc) <= set(NUC) and set(der) <= set(NUC) and len(set(anc)) > 1
    verify_mutations(d, out, discrete, model)


@pytest.mark.parametrize("discrete", [True, False])
def test_one_way_mutation(engine, discrete):
    d = small_ts()
    model = matrix_model(["0", "1"], [[0.0, 1.0], [0.0, 1.0]], [1.0, 0.0])
    out = overlay(d, model, 1.0, discrete)
    _, anc, site, _, _, _, der = rows(out)
    assert out.num_sites > 0 and set(anc) == {"0"} and set(der) == {"1"}
    assert np.bincount(site, minlength=out.num_sites).min() >= 1
    if discrete:
        assert max_per_site(out) > 1
    verify_mutations(d, out, discrete, model)


@pytest.mark.parametrize("discrete", [True, False])
def test_flip_flop_mutation(engine, discrete):
    d = small_ts()
    model = matrix_model(NUC, [[0, 0, 0.5, 0.5], [0, 0, 0.5, 0.5], [0.5, 0.5, 0, 0], [0.5, 0.5, 0, 0]])
    out = overlay(d, model, 5.0, discrete)
    _, anc, site, _, par, _, der = rows(out)
    assert out.num_sites > 0
    if discrete:
        assert max_per_site(out) > 1
    for k in range(len(der)):
        inherited = anc[site[k]] if par[k] == -1 else der[par[k]]
        assert der[k] in ("GT" if inherited in "AC" else "AC")
    verify_mutations(d, out, discrete, model)


@pytest.mark.parametrize("discrete", [True, False])
def test_do_nothing_mutations(engine, discrete):
    d = small_ts()
    model = matrix_model(NUC, np.eye(4).tolist())
    out = overlay(d, model, 5.0, discrete)
    _, anc, site, _, _, _, der = rows(out)
    assert out.num_sites > 0 and len(set(anc)) > 1
    assert all(der[k] == anc[site[k]] for k in range(len(der)))  # silent, but recorded
    verify_mutations(d, out, discrete, model)


@pytest.mark.parametrize("discrete", [True, False])
def test_uniform_mutations_away_from_the_root_state(engine, discrete):
    d = small_ts()
    model = matrix_model(NUC, [[0, 1 / 3, 1 / 3, 1 / 3], [0, 0, 0.5, 0.5], [0, 0.5, 0, 0.5], [0, 0.5, 0.5, 0]],
                         [1.0, 0, 0, 0])
    out = overlay(d, model, 20.0, discrete)
    _, anc, _, _, _, _, der = rows(out)
    assert set(anc) == {"A"}
    counts = {a: der.count(a) for a in NUC}
    assert counts["A"] == 0 and min(counts["C"], counts["G"], counts["T"]) > 0
    if discrete:
        assert max_per_site(out) > 1
    verify_mutations(d, out, discrete, model)


@pytest.mark.parametrize("discrete", [True, False])
def test_circular_mutations(engine, discrete):
    d = small_ts()
    model = matrix_model(NUC, [[0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1], [1, 0, 0, 0]], [1.0, 0, 0, 0])
    out = overlay(d, model, 2.0, discrete)
    _, anc, site, _, par, _, der = rows(out)
    assert out.num_sites > 0 and set(anc) == {"A"}
    if discrete:
        assert max_per_site(out) > 1
    for k in range(len(der)):
        inherited = "A" if par[k] == -1 else der[par[k]]
        assert NUC.index(der[k]) == (NUC.index(inherited) + 1) % 4
    verify_mutations(d, out, discrete, model)


def test_integer_sites(engine):
    d = small_ts(seed=5)
    out = overlay(d, matrix_model(["0", "1"], [[0, 1], [1, 0]], [1, 0]), 5.0, True)
    pos = rows(out)[0]
    assert pos[0] == 0.0 and np.all(pos == np.floor(pos)) and pos[-1] < d["sequence_length"]
    assert len(pos) == 10  # at this rate every one of the ten positions is hit


@pytest.mark.parametrize("keep", [True, False])
def test_zero_rate_on_mutated_input(engine, keep):
    d = small_ts()
    first = overlay(d, "binary", 2.0, False, seed=1)
    again = msp.sim_mutations(first, rate=0, keep=keep, model="binary", random_seed=2, discrete_genome=False)
    if keep:
        for a, b in zip(rows(first), rows(again)):
            assert np.array_equal(np.asarray(a), np.asarray(b))
    else:
        assert again.num_sites == 0 and again.num_mutations == 0
    assert again.num_provenances == first.num_provenances + 1


def test_stick_tree_windows_pick_the_branch(engine):
    # TestInterval.test_stick_tree, tests/test_mutations.py:1134-1200
    d = cases.stick_tree(5, 1.0)
    for k in range(4):
        out = overlay(d, "binary", 50.0, False, seed=3 + k, start_time=k, end_time=k + 1)
        _, _, _, node, _, time, _ = rows(out)
        assert len(node) > 0 and set(node.tolist()) == {k}
        assert (time >= k).all() and (time < k + 1).all()
    out = overlay(d, "binary", 50.0, False, seed=9, start_time=1.5, end_time=2.5)
    _, _, _, node, _, time, _ = rows(out)
    assert set(node.tolist()) == {1, 2} and (time >= 1.5).all() and (time < 2.5).all()
    assert (time[node == 1] < 2).all() and (time[node == 2] >= 2).all()
    for empty in ((4.0, 10.0), (-5.0, 0.0)):  # above the root and below the leaves: no branch at all
        assert overlay(d, "binary", 50.0, False, start_time=empty[0], end_time=empty[1]).num_mutations == 0


def test_coalescent_trees_windows(engine):
    # TestInterval.test_coalescent_trees / test_negative_time, :1202-1229
    d = small_ts(seed=2, samples=8, L=20.0)
    tmax = float(d["nodes"]["time"].max())
    for lo, hi in ((0.0, tmax / 4), (tmax / 4, tmax / 2), (tmax / 8, tmax)):
        out = overlay(d, "jc69", 8.0 / tmax, True, seed=4, start_time=lo, end_time=hi)
        time = rows(out)[5]
        assert len(time) > 0 and (time >= lo).all() and (time < hi).all()
        verify_mutations(d, out, True)
    assert overlay(d, "jc69", 1.0, True, start_time=-10.0, end_time=-1.0).num_mutations == 0


def rate_map_overlay(d, position, rate, discrete, seed=1):
    ts = mtables.TableCollection.fromdict(d).tree_sequence()
    return msp.sim_mutations(ts, rate=msp.RateMap(position=position, rate=rate), random_seed=seed,
                             discrete_genome=discrete)


def test_rate_map_zero_stretches_and_integer_sites(engine):
    """TestMutateRateMap, tests/test_mutations.py:527-590: on a discrete genome an interval of the map
    covers the INTEGERS inside it, whatever its break points; a stretch of rate 0 stays empty."""
    d = small_ts(seed=6)  # ten positions, several trees
    assert np.array_equal(rows(rate_map_overlay(d, [0, 3, 7, 10], [4.0, 0.0, 4.0], True))[0], [0, 1, 2, 7, 8, 9])
    # [0, 2.9) holds 0, 1, 2; [4.3, 4.9) and [9.5, 10) hold no integer at all, however high their rates
    out = rate_map_overlay(d, [0, 2.9, 4.3, 4.9, 9.5, 10], [4.0, 0.0, 10.0, 0.0, 10.0], True, seed=2)
    assert np.array_equal(rows(out)[0], [0, 1, 2])
    out = rate_map_overlay(d, [0, 0.5, 1.0, 4.5, 5.0, 9.5, 10], [0, 10, 0, 10, 0, 10], True, seed=2)
    assert out.num_mutations == 0 and out.num_sites == 0
    # continuous genome: every stretch with a rate gets mutations, the others none
    position, rate = [0, 3, 4.3, 4.9, 10], [5.0, 0.0, 5.0, 0.0]
    pos = rows(rate_map_overlay(d, position, rate, False, seed=2))[0]
    for left, right, r in zip(position[:-1], position[1:], rate):
        assert (np.sum((pos >= left) & (pos < right)) > 0) == (r > 0)


def cut_above(d, t):
    """The tree sequence with every edge into a node older than `t` removed: trees with several roots,
    as `sim_ancestry(..., end_time=t)` leaves them (tests/test_mutations.py:1678, 1758)."""
    e = d["edges"]
    keep = d["nodes"]["time"][e["parent"]] <= t
    out = dict(d)
    out["edges"] = {k: v[keep] for k, v in e.items()}
    return out


@pytest.mark.parametrize("model", ["jc69", "infinite_alleles", "slim"])
def test_incomplete_trees(engine, model):
    full = small_ts(seed=4, samples=8)
    d = cut_above(full, float(np.quantile(full["nodes"]["time"], 0.7)))
    e = d["edges"]
    cover = (e["left"] <= 0.0) & (0.0 < e["right"])
    in_tree = np.union1d(e["parent"][cover], e["child"][cover])
    assert np.sum(parent_array(d, 0.0)[in_tree] == -1) > 1  # several roots in the first tree
    m = {"jc69": "jc69", "infinite_alleles": msp.InfiniteAlleles(start_allele=3),
         "slim": msp.SLiMMutationModel(type=1, next_id=10)}[model]
    out = overlay(d, m, 2.0, True, seed=23)
    assert out.num_mutations > 10
    verify_mutations(d, out, True)
    pos, anc, site, node, par, time, der = rows(out)
    if model == "infinite_alleles":
        expect = 3
        for j in range(len(pos)):
            assert int(anc[j]) == expect
            expect += 1
            for k in np.flatnonzero(site == j):
                assert int(der[k]) == expect
                expect += 1
        assert m.next_allele == expect
    if model == "slim":
        assert set(anc) == {""} and m.next_id == 10 + len(der)
        for k in range(len(der)):
            ids = der[k].split(",")
            assert int(ids[-1]) == 10 + k
            assert (",".join(ids[:-1]) == der[par[k]]) if par[k] >= 0 else len(ids) == 1


# ---------------------------------------------------------------------------------------------
# keep=True over inputs that did not come out of an overlay (TestKeep, tests/test_mutations.py:1231-1520):
# hand-placed mutations on every branch, unique metadata on the existing rows, multi-character alleles,
# a site without mutations, rows whose parent column has to be worked out again.
# (The parity tests on the device layer only over overlay output so far, so these are new ground there.)


@pytest.fixture(params=["oracle", pytest.param("device", marks=[pytest.mark.gpu])])
def keep_engine(request, monkeypatch):
    return _engine(request, monkeypatch)


def pack(strings):
    data, off = mtables.encode_ragged(strings)
    return data.view(np.int8), off


def with_existing(d, sites, muts):
    """`d` with hand-made site / mutation tables: sites = [(position, ancestral_state, metadata)],
    muts = [dict(site, node, derived, time, parent, metadata)], already in table order."""
    out = dict(d)
    anc, anc_off = pack([s[1] for s in sites])
    smd, smd_off = pack([s[2] for s in sites])
    out["sites"] = {"position": np.array([s[0] for s in sites], dtype=np.float64), "ancestral_state": anc,
                    "ancestral_state_offset": anc_off, "metadata": smd, "metadata_offset": smd_off}
    der, der_off = pack([m["derived"] for m in muts])
    mmd, mmd_off = pack([m["metadata"] for m in muts])
    out["mutations"] = {
        "site": np.array([m["site"] for m in muts], dtype=np.int32),
        "node": np.array([m["node"] for m in muts], dtype=np.int32),
        "parent": np.array([m["parent"] for m in muts], dtype=np.int32),
        "time": np.array([m["time"] for m in muts], dtype=np.float64),
        "derived_state": der, "derived_state_offset": der_off, "metadata": mmd, "metadata_offset": mmd_off}
    return out


def tag(k):
    return int(k).to_bytes(4, "little")


def branch_mutations(d, oldest_first=False):
    """tsutil.insert_branch_mutations (tests/tsutil.py:52-84): one site per tree at its left end, a
    mutation above every non-root node flipping 0 <-> 1 along each lineage, parents before children;
    times at the middle of the branch; unique metadata on every row.  ``oldest_first`` lists each
    site's rows by decreasing time, the order an overlay writes (needed where new rows are merged
    into the site: the reference orders existing rows by id and new ones by time)."""
    lefts = np.unique(d["edges"]["left"])
    node_time = d["nodes"]["time"]
    sites, muts = [], []
    for j, x in enumerate(lefts):
        parent = parent_array(d, x)
        sites.append((float(x), "0", b"site%d" % j))
        children = {}
        for c in np.flatnonzero(parent >= 0):
            children.setdefault(int(parent[c]), []).append(int(c))
        roots = [u for u in children if parent[u] == -1]
        for root in roots:
            state, above, stack = {root: 0}, {root: -1}, [root]
            while stack:
                u = stack.pop()
                stack.extend(children.get(u, []))
                v = parent[u]
                if v != -1:
                    state[u] = (state[v] + 1) % 2
                    above[u] = len(muts)
                    muts.append(dict(site=j, node=u, derived=str(state[u]), parent=above[v],
                                     time=(node_time[u] + node_time[v]) / 2, metadata=tag(len(muts))))
    if oldest_first:
        order = sorted(range(len(muts)), key=lambda k: (muts[k]["site"], -muts[k]["time"]))
        new_id = {old: new for new, old in enumerate(order)}
        muts = [dict(muts[k], parent=new_id.get(muts[k]["parent"], -1)) for k in order]
    return with_existing(d, sites, muts)


def table_rows(ts_or_dict):
    ts = ts_or_dict if hasattr(ts_or_dict, "tables") else mtables.TableCollection.fromdict(ts_or_dict).tree_sequence()
    t = ts.tables
    pos, anc, site, node, par, time, der = rows(ts)
    md = [bytes(x) for x in table_bytes(t.mutations.metadata, t.mutations.metadata_offset)]
    smd = [bytes(x) for x in table_bytes(t.sites.metadata, t.sites.metadata_offset)]
    return dict(pos=pos, anc=anc, site=site, node=node, par=par, time=time, der=der, md=md, smd=smd)


def table_bytes(data, offset):
    raw = np.asarray(data).astype(np.int8).view(np.uint8).tobytes()
    off = np.asarray(offset).astype(np.int64)
    return [raw[off[i]:off[i + 1]] for i in range(len(off) - 1)]


def verify_sites(original, updated, discrete):
    """TestKeep.verify_sites (tests/test_mutations.py:1258-1298): every existing site and row is still
    there with its node, state and metadata, and the parent of an existing row, followed up through
    any new rows in between, is the existing row it had before."""
    o, u = table_rows(original), table_rows(updated)
    old_site = {p: j for j, p in enumerate(o["pos"])}
    old_row = {m: k for k, m in enumerate(o["md"])}
    assert len(old_row) == len(o["md"])  # the metadata identifies the rows
    found = 0
    for j, p in enumerate(u["pos"]):
        if p not in old_site:
            continue
        found += 1
        oj = old_site[p]
        assert u["anc"][j] == o["anc"][oj] and u["smd"][j] == o["smd"][oj]
        mine = np.flatnonzero(u["site"] == j)
        if not discrete:
            assert len(mine) == int(np.sum(o["site"] == oj))
        for k in mine:
            if u["md"][k] not in old_row:
                continue
            ok = old_row[u["md"][k]]
            assert o["site"][ok] == oj and u["node"][k] == o["node"][ok] and u["der"][k] == o["der"][ok]
            assert u["time"][k] == o["time"][ok]
            q = u["par"][k]
            while q != -1 and u["md"][q] not in old_row:
                q = u["par"][q]
            assert (o["par"][ok] == -1) if q == -1 else (old_row[u["md"][q]] == o["par"][ok])
    assert found == len(o["pos"])
    kept = [m for m in u["md"] if m in old_row]
    assert sorted(kept) == sorted(o["md"])


def check_keep(d_existing, rate, seed):
    """TestKeep.verify (:1237-1256): on a continuous genome keep=True gives exactly the existing rows
    plus what keep=False would have drawn."""
    ts = mtables.TableCollection.fromdict(d_existing).tree_sequence()
    fresh = msp.sim_mutations(ts, rate=rate, random_seed=seed, discrete_genome=False, keep=False)
    both = msp.sim_mutations(ts, rate=rate, random_seed=seed, discrete_genome=False, keep=True)
    assert fresh.num_sites > 0
    assert ts.num_sites + fresh.num_sites == both.num_sites
    assert ts.num_mutations + fresh.num_mutations == both.num_mutations
    old, new, union = (set(rows(x)[0].tolist()) for x in (ts, fresh, both))
    assert old | new == union
    verify_sites(ts, both, discrete=False)
    return both


def test_keep_adds_to_a_first_layer(keep_engine):
    d = small_ts(seed=2)
    first = overlay(d, "binary", 1.0, False, seed=11)
    r = table_rows(first)
    existing = with_existing(
        d, [(p, a, b"") for p, a in zip(r["pos"], r["anc"])],
        [dict(site=int(s), node=int(n), derived=x, time=float(t), parent=int(p), metadata=tag(k))
         for k, (s, n, x, t, p) in enumerate(zip(r["site"], r["node"], r["der"], r["time"], r["par"]))])
    assert len(r["pos"]) > 5
    check_keep(existing, 1.0, 2)


def test_keep_over_a_mutation_on_every_branch(keep_engine):
    d = small_ts(seed=2, samples=5)
    existing = branch_mutations(d)
    assert len(existing["sites"]["position"]) > 1
    both = check_keep(existing, 3.0, 7)
    assert both.num_sites > len(existing["sites"]["position"])
    # with no new mutations at all the tables come back exactly, parents included
    ts = mtables.TableCollection.fromdict(existing).tree_sequence()
    same = msp.sim_mutations(ts, rate=0, random_seed=1, keep=True, discrete_genome=False)
    a, b = table_rows(ts), table_rows(same)
    for key in a:
        assert np.array_equal(np.asarray(a[key]), np.asarray(b[key])), key


def test_keep_parents_are_worked_out_again_on_a_discrete_genome(keep_engine):
    # new mutations land ON the existing sites (integer positions 0 .. 9), above and below existing rows
    d = small_ts(seed=3, samples=6)
    existing = branch_mutations(d, oldest_first=True)
    ts = mtables.TableCollection.fromdict(existing).tree_sequence()
    model = matrix_model(["0", "1"], [[0, 1], [1, 0]], [1, 0])
    both = msp.sim_mutations(ts, rate=1.0, random_seed=5, keep=True, discrete_genome=True, model=model)
    assert both.num_mutations > ts.num_mutations
    verify_sites(ts, both, discrete=True)
    u = table_rows(both)
    new_on_old = [k for k in range(len(u["md"])) if u["md"][k] == b"" and u["pos"][u["site"][k]] in
                  set(existing["sites"]["position"].tolist())]
    assert len(new_on_old) > 0
    # every row's state differs from the state it inherits (flip model), new and existing alike at new rows
    for k in new_on_old:
        inherited = u["anc"][u["site"][k]] if u["par"][k] == -1 else u["der"][u["par"][k]]
        assert u["der"][k] != inherited


def test_keep_multichar_alleles_under_a_nucleotide_layer(keep_engine):
    # tests/test_mutations.py:1356-1374: existing alleles the model does not know are fine as long as
    # no new mutation has to read them (continuous genome: new sites only)
    d = small_ts(seed=3, samples=12)
    first = table_rows(overlay(d, "binary", 1.0, False, seed=1))
    assert len(first["pos"]) > 2
    existing = with_existing(
        d, [(p, "A" * j, b"") for j, p in enumerate(first["pos"])],
        [dict(site=int(s), node=int(n), derived="T" * int(s), time=float(t), parent=int(p), metadata=tag(k))
         for k, (s, n, t, p) in enumerate(zip(first["site"], first["node"], first["time"], first["par"]))])
    ts = mtables.TableCollection.fromdict(existing).tree_sequence()
    both = msp.sim_mutations(ts, rate=1.0, random_seed=1, keep=True, discrete_genome=False)
    verify_sites(ts, both, discrete=False)
    assert both.num_sites > ts.num_sites


def test_keep_a_site_without_mutations(keep_engine):
    # tests/test_mutations.py:1335-1338 (tsutil.insert_site): the empty site stays
    d = small_ts(seed=3, samples=12)
    existing = with_existing(d, [(0.5, "XY", b"meta")], [])
    ts = mtables.TableCollection.fromdict(existing).tree_sequence()
    both = msp.sim_mutations(ts, rate=3.0, random_seed=7, keep=True, discrete_genome=False)
    u = table_rows(both)
    j = int(np.flatnonzero(u["pos"] == 0.5)[0])
    assert u["anc"][j] == "XY" and u["smd"][j] == b"meta" and not np.any(u["site"] == j)
    assert both.num_sites > 1 and both.num_mutations == both.num_sites - 1
    gone = msp.sim_mutations(ts, rate=3.0, random_seed=7, keep=False, discrete_genome=False)
    assert 0.5 not in rows(gone)[0]


def test_same_seed_twice_does_not_repeat_sites(keep_engine):
    # tests/test_mutations.py:1340-1354: the second layer redraws positions that are taken
    d = small_ts(seed=3, samples=12)
    first = overlay(d, "jc69", 1.0, False, seed=1)
    second = msp.sim_mutations(first, rate=1.0, random_seed=1, keep=True, discrete_genome=False)
    pos = rows(second)[0]
    assert len(set(pos.tolist())) == len(pos) and set(rows(first)[0].tolist()) < set(pos.tolist())


def test_layers_in_overlapping_windows(keep_engine):
    # tests/test_mutations.py:1434-1454
    d = small_ts(seed=3, samples=12)
    one = overlay(d, "jc69", 1.0, True, seed=1, start_time=1.0, end_time=2.0)
    two = msp.sim_mutations(one, rate=1.0, random_seed=3, discrete_genome=True, start_time=0.0, end_time=1.5)
    assert two.num_mutations > one.num_mutations > 0
    verify_mutations(d, two, True)


def test_layering_a_model_with_other_alleles(keep_engine):
    # tests/test_mutations.py:1478-1520: a second model that turns everything into one of two identical
    # multi-byte alleles, over binary rows; new rows at old sites must read the old states
    d = small_ts(seed=3, samples=5, L=100.0)
    tree = "\U0001F384"
    model = msp.MatrixMutationModel(["0", "1", tree, tree], [0, 0, 0.5, 0.5],
                                    [[0, 0, 1, 0], [0, 0, 0, 1], [0, 0, 0.5, 0.5], [0, 0, 0.5, 0.5]])
    r = table_rows(overlay(d, "binary", 0.05, True, seed=1))
    assert set(r["anc"]) | set(r["der"]) <= {"0", "1"} and len(r["pos"]) > 3
    existing = with_existing(
        d, [(p, a, b"") for p, a in zip(r["pos"], r["anc"])],
        [dict(site=int(s), node=int(n), derived=x, time=float(t), parent=int(p), metadata=tag(k))
         for k, (s, n, x, t, p) in enumerate(zip(r["site"], r["node"], r["der"], r["time"], r["par"]))])
    ts = mtables.TableCollection.fromdict(existing).tree_sequence()
    both = msp.sim_mutations(ts, rate=0.2, random_seed=2, keep=True, model=model)
    verify_sites(ts, both, discrete=True)
    u = table_rows(both)
    old_positions = set(r["pos"].tolist())
    assert old_positions < set(u["pos"].tolist())
    new_at_old = 0
    for k in range(len(u["md"])):
        at_old = u["pos"][u["site"][k]] in old_positions
        if not at_old:
            assert u["anc"][u["site"][k]] == tree and u["der"][k] == tree
        elif u["md"][k] == b"":
            assert u["der"][k] == tree
            new_at_old += 1
    assert new_at_old > 0


# ---------------------------------------------------------------------------------------------
# The same hand-made existing tables through the replay harness, bit for bit against the oracle (the
# inputs of tests/test_oracle_fuzz.py::test_port_equals_reference_over_hand_made_existing_tables, where
# the oracle's restatement is held to the compiled reference).
@pytest.mark.gpu
@pytest.mark.parametrize("case", range(12))
def test_replay_parity_over_hand_made_existing_tables(native_built, oracle_mod, case):
    import test_gpu_replay as replay
    from msprime_b200 import _native

    rng = np.random.default_rng(7000 + case)
    d = small_ts(seed=int(rng.integers(1, 6)), samples=int(rng.integers(2, 9)), L=float(rng.choice([5, 10, 30])))
    oldest_first = bool(case % 2)
    existing = branch_mutations(d, oldest_first=oldest_first)
    if case in (4, 5):  # plus a site without mutations, between the others
        x = float(np.setdiff1d(np.arange(d["sequence_length"]) + 0.5, existing["sites"]["position"])[0])
        t = table_rows(existing)
        j = int(np.searchsorted(t["pos"], x))
        sites = [(p, a, m) for p, a, m in zip(t["pos"], t["anc"], t["smd"])]
        sites.insert(j, (x, "empty", b"no mutations here"))
        muts = [dict(site=int(s) + (1 if s >= j else 0), node=int(n), derived=x_, time=float(tm), parent=int(p),
                     metadata=m) for s, n, x_, tm, p, m in zip(t["site"], t["node"], t["der"], t["time"], t["par"],
                                                               t["md"])]
        existing = with_existing(d, sites, muts)
    model = cases.BINARY if case % 3 else cases.JC69_SI  # JC69 cannot read "0" / "1": an error, the same one
    with _native.Context(0) as ctx:
        kw = dict(model=model, map_position=[0, d["sequence_length"]], map_rate=[float(rng.choice([0.0, 0.2, 1.0, 4.0]))],
                  discrete=oldest_first and case % 4 != 3, keep=True, seed=case + 1)
        try:
            replay.check_case(ctx, oracle_mod, existing, **kw)
        except oracle_mod.OracleError as e:
            with pytest.raises(_native.LibraryError) as got:
                replay.gpu_run(ctx, existing, model=model, stream=np.random.default_rng(case + 1).random(200_000),
                               map_position=kw["map_position"], map_rate=kw["map_rate"], discrete=kw["discrete"],
                               keep=True)
            assert got.value.status == e.status


@pytest.mark.gpu
def test_existing_offset_columns_are_checked_at_the_boundary(native_built):
    """tskit refuses offset columns that do not start at 0 or that decrease (TSK_ERR_BAD_OFFSET, when the
    columns are set); the C boundary does the same before anything is uploaded, because the kernels take
    lengths as differences of neighbouring offsets."""
    from msprime_b200 import _native

    d = small_ts(seed=2, samples=5)
    good = branch_mutations(d)
    with _native.Context(0) as ctx:
        ctx.load_tables(good)
        for table, column in (("sites", "ancestral_state_offset"), ("sites", "metadata_offset"),
                              ("mutations", "derived_state_offset"), ("mutations", "metadata_offset")):
            for damage in ("start", "order"):
                bad = dict(good)
                bad[table] = dict(good[table])
                off = np.array(good[table][column], dtype=np.uint64)
                if damage == "start":
                    off[0] = 1
                else:
                    off[1], off[2] = max(off[1], off[2]) + 1, min(off[1], off[2])
                    if off[1] <= off[2]:
                        continue
                bad[table][column] = off
                with pytest.raises(_native.LibraryError) as e:
                    ctx.load_tables(bad)
                assert e.value.status == (-200 ^ (1 << 13)), (table, column, damage)
        ctx.load_tables(good)  # the context is still usable
