"""CPU: the oracle's C restatement (oracle/mutgen_port.c) against the golden vectors
that the unmodified reference mutgen produced (tests/golden/make_golden.py), plus the
known-answer cases of the reference's own tests.  Bit-exact."""
import numpy as np
import pytest

import cases
import golden_cases
import golden_util

ALL = golden_cases.all_cases()


def run_port(oracle_mod, c, which="port"):
    ts, stream = golden_util.case_inputs(c)
    return oracle_mod.sim_mutations(
        ts, which=which, map_position=c["map_position"], map_rate=c["map_rate"], model=cases.MODELS[c["model"]],
        discrete_genome=c["discrete"], keep=c["keep"], start_time=c["start_time"], end_time=c["end_time"],
        stream=stream)


@pytest.mark.parametrize("c", ALL, ids=[c["name"] for c in ALL])
def test_port_matches_reference_golden(oracle_mod, c):
    want = golden_util.golden(c["name"])
    got = run_port(oracle_mod, c)
    assert not golden_util.diff_tables(want, got)
    assert got["draws"] == want["draws"]
    assert got["next_allele"] == want["next_allele"]


def test_compiled_reference_still_matches_golden(oracle_mod):
    if not oracle_mod.have("ref"):
        pytest.skip("compiled reference absent (only built where /root/reference exists)")
    for c in ALL[::7]:
        want = golden_util.golden(c["name"])
        got = run_port(oracle_mod, c, which="ref")
        assert not golden_util.diff_tables(want, got), c["name"]


def test_zero_rate_intervals_give_exact_positions():
    # tests/test_mutations.py:528-546
    g = golden_util.golden("four_leaf-zero_rate_intervals")
    assert sorted(set(g["sites"]["position"].tolist())) == [0, 1, 2, 7, 8, 9]


def test_stick_tree_window_picks_the_branch():
    # tests/test_mutations.py:1134-1165: window [k, k+1) can only hit the branch above node k
    for k in range(4):
        g = golden_util.golden(f"stick-window-{k}")
        assert len(g["mutations"]["node"]) > 0
        assert set(g["mutations"]["node"].tolist()) == {k}
        t = g["mutations"]["time"]
        assert (t >= k).all() and (t < k + 1).all()


def test_infinite_alleles_numbering():
    # tests/test_mutations.py:1704-1711: alleles are 0, 1, 2, ... in site then mutation order
    g = golden_util.golden("smc-infinite_alleles-d-a")
    s, m = g["sites"], g["mutations"]
    anc = [bytes(s["ancestral_state"][a:b]).decode() for a, b in
           zip(s["ancestral_state_offset"][:-1], s["ancestral_state_offset"][1:])]
    der = [bytes(m["derived_state"][a:b]).decode() for a, b in
           zip(m["derived_state_offset"][:-1], m["derived_state_offset"][1:])]
    nxt = 0
    mi = 0
    for site, a in enumerate(anc):
        assert int(a) == nxt
        nxt += 1
        while mi < len(der) and m["site"][mi] == site:
            assert int(der[mi]) == nxt
            nxt += 1
            mi += 1
    assert nxt == g["next_allele"]


def test_keep_with_zero_rate_reproduces_input(oracle_mod):
    # lib/tests/test_mutations.c:354-369
    base = golden_util.golden("smc-jc69-d-a")
    ts = cases.with_kept(golden_cases.build_ts(("smc", 20, 2e4, 2)), base)
    got = oracle_mod.sim_mutations(ts, which="port", map_position=[0, 2e4], map_rate=[0.0], model=cases.JC69,
                                   keep=True, stream=np.random.default_rng(1).random(50_000))
    assert not golden_util.diff_tables(base, got)


def test_keep_false_discards_input(oracle_mod):
    base = golden_util.golden("smc-jc69-d-a")
    ts = cases.with_kept(golden_cases.build_ts(("smc", 20, 2e4, 2)), base)
    got = oracle_mod.sim_mutations(ts, which="port", map_position=[0, 2e4], map_rate=[0.0], model=cases.JC69,
                                   keep=False, stream=np.random.default_rng(1).random(50_000))
    assert len(got["sites"]["position"]) == 0 and len(got["mutations"]["site"]) == 0


@pytest.mark.parametrize("kw,status", [
    (dict(map_position=[0, 5], map_rate=[1.0]), -44),       # lib/mutgen.c:807 map length
    (dict(map_position=[1, 10], map_rate=[1.0]), -46),      # rate_map.c start non-zero
    (dict(map_position=[0, 5, 5, 10], map_rate=[1, 1, 1.0]), -48),  # unsorted
    (dict(map_position=[0, 10], map_rate=[-1.0]), -43),     # negative rate
    (dict(map_position=[0, 10], map_rate=[1.0], start_time=2.0, end_time=1.0), -4),  # lib/mutgen.c:786
])
def test_error_codes(oracle_mod, kw, status):
    # lib/tests/test_mutations.c:141-222
    ts = cases.four_leaf_tree()
    with pytest.raises(oracle_mod.OracleError) as e:
        oracle_mod.sim_mutations(ts, which="port", model=cases.JC69, stream=np.zeros(10) + 0.5, **kw)
    assert e.value.status == status
