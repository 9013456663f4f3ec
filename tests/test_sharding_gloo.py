"""CPU, world_size 2 over gloo: the host side of a genome-sharded call
(msprime_b200/sharding.py): cuts, the two count all-gathers, id / allele / byte-offset
bases and the final table gather.  The per-rank engine is a stand-in with the same four
methods as ``_native.Context`` that computes its slice with the CPU oracle (a rate map
that is zero outside the slice), because no GPU exists here; the collective plumbing under
test is the real one."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

import cases
import table_checks
from msprime_b200 import sharding


class OracleSliceEngine:
    """generate_begin / generate_end / shift_offsets / fetch on top of the oracle."""

    def __init__(self, ts, model, rate, seed):
        from oracle import oracle

        oracle.build(("port",))
        self.oracle, self.ts, self.model, self.rate, self.seed = oracle, ts, model, rate, seed

    def _run(self, start_allele):
        L = self.ts["sequence_length"]
        pos, rate = [0.0], []
        for a, b, r in ((0.0, self.left, 0.0), (self.left, self.right, self.rate), (self.right, L, 0.0)):
            if b > a:
                pos.append(b)
                rate.append(r)
        model = dict(self.model)
        if model["kind"] == "infinite_alleles":
            model["start_allele"] = start_allele
        return self.oracle.sim_mutations(self.ts, which="port", map_position=pos, map_rate=rate, model=model,
                                         keep=False, seed=self.seed)

    def generate_begin(self, params):
        self.left, self.right, self.start = params["slice_left"], params["slice_right"], params["start_allele"]
        r = self._run(self.start)
        return {"num_sites": len(r["sites"]["position"]), "num_mutations": len(r["mutations"]["site"]),
                "num_new_mutations": len(r["mutations"]["site"]), "next_allele": r["next_allele"]}

    def generate_end(self, site_base, mutation_base, allele_base):
        r = self._run(allele_base)
        m = r["mutations"]
        m["site"] = m["site"] + np.int32(site_base)
        m["parent"] = np.where(m["parent"] >= 0, m["parent"] + np.int32(mutation_base), -1).astype(np.int32)
        self.res = r
        return {"ancestral_state_bytes": len(r["sites"]["ancestral_state"]), "site_metadata_bytes": 0,
                "derived_state_bytes": len(m["derived_state"]), "mutation_metadata_bytes": 0}

    def shift_offsets(self, anc, smd, der, mmd):
        s, m = self.res["sites"], self.res["mutations"]
        s["ancestral_state_offset"] = s["ancestral_state_offset"] + np.uint64(anc)
        s["metadata_offset"] = s["metadata_offset"] + np.uint64(smd)
        m["derived_state_offset"] = m["derived_state_offset"] + np.uint64(der)
        m["metadata_offset"] = m["metadata_offset"] + np.uint64(mmd)

    def fetch(self):
        return {"sites": self.res["sites"], "mutations": self.res["mutations"]}


def rank_slice(rank, world, model_name):
    ts = cases.smc_ts(20, 2e4, 2)
    model = cases.MODELS[model_name]
    cuts = sharding.slice_cuts(ts["sequence_length"], world)
    engine = OracleSliceEngine(ts, model, 2e-6, seed=100 + rank)
    start = 7 if model_name == "infinite_alleles" else 0
    params = {"slice_left": cuts[rank], "slice_right": cuts[rank + 1], "start_allele": start}
    return ts, engine, params


def worker(rank, world, port, model_name, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ts, engine, params = rank_slice(rank, world, model_name)
        local, info = sharding.run_slice(engine, params, rank=rank, world=world,
                                         allgather=sharding.dist_allgather(),
                                         start_allele=params["start_allele"])
        merged = sharding.gather_results(local, rank=rank, world=world)
        if rank == 0:
            np.savez(out_path, info=np.array([info["num_sites"], info["num_mutations"], info["next_allele"]],
                                             dtype=np.uint64),
                     **{f"{t}/{c}": np.asarray(merged[t][c]) for t in merged for c in merged[t]})
    finally:
        dist.destroy_process_group()


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("model_name", ["infinite_alleles", "jc69"])
def test_two_ranks_over_gloo(tmp_path, model_name):
    world = 2
    out = str(tmp_path / "merged.npz")
    mp.spawn(worker, args=(world, free_port(), model_name, out), nprocs=world, join=True)
    z = np.load(out)
    merged = {"sites": {}, "mutations": {}}
    for key in z.files:
        if "/" in key:
            t, c = key.split("/")
            merged[t][c] = z[key]
    S, M, next_allele = (int(x) for x in z["info"])
    # the same two slices computed serially, bases applied by hand
    parts, counts = [], []
    for rank in range(world):
        ts, engine, params = rank_slice(rank, world, model_name)
        counts.append(engine.generate_begin(params))
    used = [c["next_allele"] - 7 if model_name == "infinite_alleles" else 0 for c in counts]
    sb = mb = ab = 0
    anc = der = 0
    for rank in range(world):
        ts, engine, params = rank_slice(rank, world, model_name)
        engine.generate_begin(params)
        b = engine.generate_end(sb, mb, params["start_allele"] + ab)
        engine.shift_offsets(anc, 0, der, 0)
        parts.append(engine.fetch())
        sb += counts[rank]["num_sites"]
        mb += counts[rank]["num_mutations"]
        ab += used[rank]
        anc += b["ancestral_state_bytes"]
        der += b["derived_state_bytes"]
    want = sharding.merge_slices(parts)
    for t in want:
        for c in want[t]:
            assert np.array_equal(np.asarray(want[t][c]), merged[t][c]), (t, c)
    assert (S, M) == (sb, mb)
    assert len(merged["sites"]["position"]) == S and len(merged["mutations"]["site"]) == M
    table_checks.check_result(ts, merged, discrete=True)
    if model_name == "infinite_alleles":
        assert next_allele == 7 + S + M
        nums = [int(x) for x in table_checks.ragged(merged["sites"]["ancestral_state"],
                                                    merged["sites"]["ancestral_state_offset"])]
        nums += [int(x) for x in table_checks.ragged(merged["mutations"]["derived_state"],
                                                     merged["mutations"]["derived_state_offset"])]
        assert sorted(nums) == list(range(7, 7 + S + M))  # globally consecutive, no gaps, no repeats


def test_slice_cuts():
    assert sharding.slice_cuts(10, 2) == [0.0, 5.0, 10.0]
    assert sharding.slice_cuts(2.5e8, 8)[3] == 93750000.0
    c = sharding.slice_cuts(7, 3)
    assert c[0] == 0 and c[-1] == 7 and all(x == int(x) for x in c[:-1])
    with pytest.raises(ValueError):
        sharding.slice_cuts(2, 4)
    assert sharding.exclusive_bases([[1, 10], [2, 20], [3, 30]], 2) == [3, 30]


def test_balanced_cuts_even_out_the_expected_mutations():
    """Equal-length slices of a genome whose rate map varies a lot hold very different numbers of mutations;
    the mass-balanced cuts hold the same number each, and are integers on a discrete
    genome.  The expectation per slice is the exact sum over clipped edges (msprime_b200/validate.py)."""
    import cases
    from msprime_b200 import validate

    ts = cases.smc_ts(100, 2e6, 7)
    L = ts["sequence_length"]
    rng = np.random.default_rng(3)
    pos = np.linspace(0, L, 41)
    rate = rng.lognormal(np.log(1e-8), 1.5, 40)

    def per_slice(cuts, **win):
        out = []
        e = ts["edges"]
        for a, b in zip(cuts[:-1], cuts[1:]):
            clipped = {"sequence_length": L, "nodes": ts["nodes"],
                       "edges": {"left": np.clip(e["left"], a, b), "right": np.clip(e["right"], a, b),
                                 "parent": e["parent"], "child": e["child"]}}
            out.append(validate.expected_new_mutations(clipped, pos, rate, discrete=True, **win))
        return np.array(out)

    for world in (2, 4, 8):
        for win in ({}, {"start_time": 1e3, "end_time": 2e4}):
            even = per_slice(sharding.slice_cuts(L, world), **win)
            cuts = sharding.balanced_cuts(ts, pos, rate, world, discrete_genome=True, **win)
            assert len(cuts) == world + 1 and cuts[0] == 0 and cuts[-1] == L
            assert all(c == np.floor(c) for c in cuts) and all(a < b for a, b in zip(cuts[:-1], cuts[1:]))
            bal = per_slice(cuts, **win)
            assert abs(bal.sum() - even.sum()) < 1e-6 * even.sum()
            assert bal.max() / bal.mean() < 1.02, (world, bal)
            assert even.max() / even.mean() > 1.2 or world == 2
    assert sharding.balanced_cuts(ts, pos, rate, 1) == [0.0, L]


# ---------------------------------------------------------------- run_slice_device: order of calls and the retry

class ScriptedContext:
    """Stands in for ``_native.Context`` in ``sharding.run_slice_device``: records the calls and lets the first
    ``generate_finish_device`` of a call begun without its wait ask for the host (``RetryWithHost``)."""

    def __init__(self, ask_for_host):
        import ctypes

        self.ask_for_host = ask_for_host
        self.calls = []
        self.counts = {"next_allele": 7}
        self.began_on_device = False

        class _P(ctypes.Structure):
            _fields_ = [("model_kind", ctypes.c_int32)]

        self.Params = _P

    def generate_begin_device(self, params):
        self.calls.append("begin_device")
        self.began_on_device = True

    def generate_begin(self, params):
        self.calls.append("begin")
        self.began_on_device = False

    def exchange_post_counts(self):
        self.calls.append("post")

    def generate_end_device(self, rank, world):
        self.calls.append("end_device")

    def generate_finish_device(self, rank, world):
        from msprime_b200 import _native

        self.calls.append("finish_device")
        if self.began_on_device and self.ask_for_host:
            raise _native.RetryWithHost("again, with the host", _native.ERR_RETRY_WITH_HOST)
        z = {k: 0 for k in ("num_sites", "num_mutations", "num_new_mutations", "ancestral_state_bytes",
                            "site_metadata_bytes", "derived_state_bytes", "mutation_metadata_bytes")}
        return dict(z), dict(z, num_sites=5), dict(z)

    def fetch(self, out=None):
        self.calls.append("fetch")
        return {"sites": {}, "mutations": {}}


class _ParamsBox:
    def __init__(self, kind):
        import ctypes

        class _P(ctypes.Structure):
            _fields_ = [("model_kind", ctypes.c_int32)]

        self.struct = _P(kind)


@pytest.mark.parametrize("ask_for_host,model_kind,host_flag", [(False, 1, False), (True, 1, False), (False, 2, False),
                                                               (False, 0, True)])
def test_run_slice_device_call_order(ask_for_host, model_kind, host_flag):
    """One wait per overlay: begin_device, counts, gather, end, gather, finish.  When finish asks for the host, the
    SAME sequence of collectives runs again behind the host-paced begin (every rank takes that turn together, so the
    all-gathers stay matched); the SLiM model (kind 2) and host_between_stages=True go there at once."""
    ctx = ScriptedContext(ask_for_host)
    gathers = []
    local, info = sharding.run_slice_device(ctx, _ParamsBox(model_kind), rank=0, world=2,
                                            gather=lambda: gathers.append(len(ctx.calls)), fetch=True,
                                            host_between_stages=host_flag)
    device = ["begin_device", "post", "end_device", "finish_device"]
    host = ["begin", "post", "end_device", "finish_device"]
    if model_kind == 2 or host_flag:
        assert ctx.calls == host + ["fetch"] and len(gathers) == 2
    elif ask_for_host:
        assert ctx.calls == device + host + ["fetch"] and len(gathers) == 4
    else:
        assert ctx.calls == device + ["fetch"] and len(gathers) == 2
    # each gather sits right after the counts / after the second stage
    for g, after in zip(gathers, ["post", "end_device"] * 2):
        assert ctx.calls[g - 1] == after
    assert info["num_sites"] == 5 and info["next_allele"] == 7 and local == {"sites": {}, "mutations": {}}
