#!/usr/bin/env python
"""Benchmark of the mutation-overlay hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]

One "step" is one overlay of the whole workload.  At N > 1 (torchrun, one rank per
GPU) the genome is cut into N position slices and the edges are replicated (strong
scaling of a fixed workload).  The collectives are NCCL calls on the library's stream:
two 8-integer all-gathers per overlay (row / allele / byte bases) and, end to end, the
in-place all-gather that replicates the tables after every rank has uploaded 1 / N of
them.  Rank 0 prints ONE JSON line:

* ``value``        new mutations per second with the tree sequence resident in HBM
                   (CUDA events on the library's stream, max over ranks);
* ``e2e``          the same metric through the host-buffer call sequence
                   load_tables -> generate -> fetch (H2D and D2H inside the timed region);
* ``roofline``     the dominant kernel's algorithmic bytes / its measured duration
                   against MEASURED_PEAKS.json;
* ``cpu_baseline`` the compiled reference mutgen (oracle/_ref) on the host cores, on a
                   bounded sample (a genome prefix) of the same workload.

``--impl reference`` times only the CPU reference (all host cores) and prints the
same line with ``"impl": "reference"``.

The oracle is used here only as the CPU baseline (and, for the keep=True workload, to prepare
the existing mutations of the INPUT with the reference itself, identically for both arms), never
as the measured product path.  Per-kernel times come from a second pass over the same steps.
"""
from __future__ import annotations

import argparse
import json
import multiprocessing as mp
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FMAX = float(np.finfo(np.float64).max)
METRIC = "sim_mutations_new_mutations_per_second"
UNIT = "mutations/s"

WORKLOADS = {
    # BASELINE.json configs[4]: the configuration the metric's target is quoted on
    "config5": dict(samples=1_000_000, L=2.5e8, recomb=1e-8, ne=1e4, ts_seed=1, rate=1e-7,
                    model="infinite_alleles", discrete=True,
                    text="config5: InfiniteAlleles rate=1e-7 discrete keep=False, 1e6 diploid samples x 2.5e8 bp"),
    # the other BASELINE.json configs: parity-test cases, timed once for BASELINE.md (not bench lines)
    "config1": dict(samples=1_000, L=1e6, recomb=1e-8, ne=1e4, ts_seed=1, rate=1e-8, model="jc69",
                    discrete=True, text="config1: JC69 rate=1e-8 discrete, 1e3 diploid samples x 1e6 bp"),
    "config2": dict(samples=10_000, L=1e7, recomb=1e-8, ne=1e4, ts_seed=1, rate=1e-8, model="binary",
                    discrete=False, text="config2: Binary rate=1e-8 continuous genome, 1e4 diploid samples x 1e7 bp"),
    "config3": dict(samples=100_000, L=1e8, recomb=1e-8, ne=1e4, ts_seed=1, rate=1e-8, model="gtr",
                    discrete=True, ratemap=1000,
                    text="config3: GTR, 1000-interval lognormal RateMap around 1e-8, 1e5 diploid samples x 1e8 bp"),
    "config4": dict(samples=100_000, L=1e8, recomb=1e-8, ne=1e4, ts_seed=1, rate=1e-8, model="hky",
                    discrete=True, keep_layer=dict(model="hky", rate=1e-8, seed=5), window=(1e3, 2e4),
                    text="config4: HKY(kappa=2) rate=1e-8 keep=True on an HKY layer (seed 5), time window [1e3, 2e4), "
                         "1e5 diploid samples x 1e8 bp"),
    "small": dict(samples=2_000, L=2e6, recomb=1e-8, ne=1e4, ts_seed=1, rate=1e-7, model="infinite_alleles",
                  discrete=True, text="small: 2e3 samples x 2e6 bp, InfiniteAlleles rate=1e-7"),
}


def log(*a):
    print(*a, file=sys.stderr, flush=True)


INPUT_TEXT = "synthetic SMC' tree sequence (msprime_b200/csrc/treegen.c)"


def config_dict(w, E, N, world):
    """The ``config`` object of the JSON line: the same for the B200 arm and the reference arm."""
    return {"workload": w["text"], "input": w.get("input_text", INPUT_TEXT), "edges": E, "nodes": N,
            "l2": "inputs larger than L2 (edge columns + records > 126 MB), no flush",
            "sharding": "whole genome on one GPU" if world == 1 else
            f"{world} genome position slices of equal expected mutation mass (edges cut to the slice, Philox stream per "
            "rank), edges replicated; "
            "two 8-integer NCCL all-gathers per overlay on the library's stream "
            "(row / allele / byte bases stay on the device), each rank copies its slice to the host"}


def overall_bytes(E, M, S, DB, AB, M0=0, kept_bytes=0):
    """SURVEY.md section 8(d) A_bytes of one overlay (compulsory traffic: every input read once, every
    output written once), with the offset columns as the survey counts them (two uint64 per row) and as
    this library writes them (one uint32 per row; metadata offsets of a result without metadata stay zero)."""
    common = E * (8 + 8 + 4 + 4) + E * (8 + 8) + M0 * (4 + 4 + 8) + kept_bytes + DB + AB
    return {"survey": common + M * (4 + 4 + 4 + 8 + 8 + 8) + S * (8 + 8 + 8),
            "as_written": common + M * (4 + 4 + 4 + 8 + 4) + S * (8 + 4)}


def load_prepared_input(w):
    """The node and edge columns scripts/make_reference_input.py left for this workload's ancestry arguments, or
    None."""
    key = "n%d_L%g_r%g_ne%g_seed%d" % (w["samples"], w["L"], w["recomb"], w["ne"], w["ts_seed"])
    path = os.path.join(ROOT, "oracle", "_ref", "inputs", key + ".npz")
    if not os.path.exists(path):
        return None
    from msprime_b200 import synth

    try:
        z = np.load(path)
        left, right = z["edge_left"], z["edge_right"]
        return {
            "sequence_length": float(z["sequence_length"]), "time_units": "generations",
            "nodes": {"time": z["node_time"], "flags": z["node_flags"]},
            "edges": {"left": left, "right": right, "parent": z["edge_parent"], "child": z["edge_child"]},
            "sites": synth.empty_sites(), "mutations": synth.empty_mutations(),
            "num_trees": int(len(np.unique(np.concatenate([left, right]))) - 1),
        }
    except Exception as e:  # noqa: BLE001  (a damaged file must not stop the bench: fall back to the generator)
        log(f"[bench] prepared input {path} unreadable ({type(e).__name__}: {e}); using the generator")
        return None


def make_input(w):
    from msprime_b200 import synth

    t0 = time.time()
    ts = None
    if w.get("input", "auto") in ("auto", "reference"):
        # The tree sequence msprime.sim_ancestry itself returns for the configuration, where it has been prepared:
        # scripts/make_reference_input.py runs the reference's own ancestry simulator (lib/msprime.c compiled
        # unmodified into oracle/_ref/libancestry_ref.so; seconds for config 2, five minutes for config 3, 2.6 hours
        # for config 5) and leaves the node and edge columns in oracle/_ref/inputs/ (git-ignored, travels with the
        # snapshot).  Plain numpy files: nothing of oracle/ is imported here.
        ts = load_prepared_input(w)
        if ts is None and w.get("input") == "reference":
            from oracle import ref_ancestry  # input preparation on the spot (small configurations)

            if not ref_ancestry.have():
                raise SystemExit("--input reference needs oracle/_ref/libancestry_ref.so (make -C oracle ancestry)")
            ts = ref_ancestry.sim_ancestry(w["samples"], sequence_length=w["L"], recombination_rate=w["recomb"],
                                           population_size=w["ne"], random_seed=w["ts_seed"])
        if ts is not None:
            w["input_text"] = ("tree sequence from the reference's own sim_ancestry (Hudson, lib/msprime.c compiled "
                               "unmodified; random_seed=%d)" % w["ts_seed"])
    if ts is None:
        ts = synth.native_treeseq(w["samples"], w["L"], w["recomb"], w["ne"], w["ts_seed"])
        w["input_text"] = INPUT_TEXT
    log(f"[bench] input: N={len(ts['nodes']['time'])} E={len(ts['edges']['left'])} trees={ts['num_trees']} "
        f"({time.time() - t0:.1f} s; {w['input_text']})")
    kl = w.get("keep_layer")
    if kl:
        # the existing mutations of a keep=True workload: one reference overlay (input preparation)
        from oracle import oracle

        _, which = cpu_kind()
        t0 = time.time()
        first = oracle.sim_mutations(ts, which=which, map_position=[0, ts["sequence_length"]], map_rate=[kl["rate"]],
                                     model=oracle_model(kl["model"]), discrete_genome=w["discrete"], keep=False,
                                     seed=kl["seed"])
        ts = dict(ts)
        ts["sites"], ts["mutations"] = first["sites"], first["mutations"]
        log(f"[bench] kept layer: {len(first['mutations']['site'])} mutations at {len(first['sites']['position'])} "
            f"sites ({time.time() - t0:.1f} s, reference overlay)")
    return ts


def time_window(w):
    lo, hi = w.get("window", (-FMAX, FMAX))
    return float(lo), float(hi)


def oracle_model(name):
    if name == "infinite_alleles":
        return {"kind": "infinite_alleles", "start_allele": 0}
    if name == "jc69":
        m = (np.full((4, 4), 1 / 3) - np.eye(4) / 3).tolist()
        return {"kind": "matrix", "alleles": ["A", "C", "G", "T"], "root_distribution": [0.25] * 4,
                "transition_matrix": m}
    if name in ("binary", "gtr", "hky"):
        # the host mirror's matrix builders (msprime/mutations.py:179-201, 763-850); BASELINE.md
        # fixes the GTR parameters that BASELINE.json leaves open
        from msprime_b200 import mutations

        mm = (mutations.BinaryMutationModel() if name == "binary" else mutations.HKY(kappa=2) if name == "hky" else
              mutations.GTR(relative_rates=[1, 2, 1, 1, 2, 1], equilibrium_frequencies=[0.3, 0.2, 0.2, 0.3]))
        return {"kind": "matrix", "alleles": list(mm.alleles), "root_distribution": [float(x) for x in mm.root_distribution],
                "transition_matrix": np.asarray(mm.transition_matrix).tolist()}
    raise ValueError(name)


def rate_map(w, L):
    """(positions, rates) of the workload's mutation rate map over [0, L)."""
    n = w.get("ratemap")
    if not n:
        return [0.0, float(L)], [w["rate"]]
    full = float(w["L"])
    pos = np.linspace(0, full, n + 1)
    rate = np.random.default_rng(3).lognormal(np.log(w["rate"]), 1.0, n)
    if L < full:  # a genome prefix of the workload (CPU sample)
        k = int(np.searchsorted(pos, L, side="left"))
        pos = np.concatenate([pos[:k], [L]])
        rate = rate[:k]
    return pos.tolist(), rate.tolist()


def native_model(name):
    m = oracle_model(name)
    if m["kind"] == "matrix":
        return {"kind": 0, "alleles": m["alleles"], "root_distribution": m["root_distribution"],
                "transition_matrix": m["transition_matrix"]}
    return {"kind": 1, "start_allele": 0}


def genome_prefix(ts, frac):
    """The bounded CPU sample: the same tree sequence restricted to [0, frac * L)."""
    L = float(ts["sequence_length"])
    X = float(max(1.0, np.floor(L * frac)))
    if X >= L:
        return ts, 1.0
    e = ts["edges"]
    keep = e["left"] < X
    out = dict(ts)
    out["sequence_length"] = X
    out["edges"] = {"left": e["left"][keep], "right": np.minimum(e["right"][keep], X),
                    "parent": e["parent"][keep], "child": e["child"][keep]}
    if "sites" in ts and len(ts["sites"]["position"]) > 0:
        # kept rows of the prefix (sites are sorted by position, mutations grouped by site)
        sp = np.asarray(ts["sites"]["position"])
        ns = int(np.searchsorted(sp, X, side="left"))
        ms = np.asarray(ts["mutations"]["site"])
        nm = int(np.searchsorted(ms, ns, side="left"))

        def ragged(cols, name, n):
            off = np.asarray(cols[name + "_offset"])[: n + 1]
            return {name: np.asarray(cols[name])[: int(off[-1])], name + "_offset": off}

        st, mt = ts["sites"], ts["mutations"]
        out["sites"] = {"position": sp[:ns], **ragged(st, "ancestral_state", ns), **ragged(st, "metadata", ns)}
        out["mutations"] = {"site": ms[:nm], "node": np.asarray(mt["node"])[:nm],
                            "parent": np.asarray(mt["parent"])[:nm], "time": np.asarray(mt["time"])[:nm],
                            **ragged(mt, "derived_state", nm), **ragged(mt, "metadata", nm)}
    return out, X / L


# ------------------------------------------------------------------ CPU reference
_CPU_JOB = {}


def _cpu_worker(seed):
    from oracle import oracle

    j = _CPU_JOB
    t0 = time.perf_counter()
    mp_, mr_ = rate_map(j["w"], j["ts"]["sequence_length"])
    lo, hi = time_window(j["w"])
    keep = bool(j["w"].get("keep_layer"))
    r = oracle.sim_mutations(j["ts"], which=j["which"], map_position=mp_, map_rate=mr_,
                             model=oracle_model(j["model"]), discrete_genome=j["discrete"], keep=keep, seed=seed,
                             start_time=lo, end_time=hi)
    kept = len(j["ts"]["mutations"]["site"]) if keep else 0
    return len(r["mutations"]["site"]) - kept, r["seconds"], time.perf_counter() - t0


def cpu_reference_pass(ts, w, cores, seed0, which):
    """One pass: ``cores`` independent processes, each overlaying the sample with its own
    seed (the reference is single-threaded, so this is how it uses a whole host)."""
    _CPU_JOB.update(ts=ts, which=which, w=w, model=w["model"], discrete=w["discrete"])
    t0 = time.perf_counter()
    if cores == 1:
        res = [_cpu_worker(seed0)]
    else:
        ctx = mp.get_context("fork")
        with ctx.Pool(cores) as pool:
            res = pool.map(_cpu_worker, [seed0 + i for i in range(cores)])
    wall = time.perf_counter() - t0
    muts = sum(r[0] for r in res)
    inner = max(r[1] for r in res)
    return muts, inner, wall


def cpu_kind():
    from oracle import oracle

    if oracle.have("ref"):
        return "reference", "ref"
    oracle.build(("port",))
    return "port", "port"


def reference_sample(ts, w, expected, seconds):
    """The bounded CPU sample both arms use: the genome prefix of the workload that takes one core about
    ``seconds`` at the reference's measured ~1.7e5 mutations/s/core (config 5)."""
    frac = min(1.0, max(1e-3, seconds * 1.7e5 / max(expected, 1.0)))
    return genome_prefix(ts, frac)


def expected_mutations(ts, w):
    from msprime_b200 import validate

    mp_, mr_ = rate_map(w, ts["sequence_length"])
    lo, hi = time_window(w)
    return validate.expected_new_mutations(ts, mp_, mr_, discrete=w["discrete"],
                                           start_time=None if lo == -FMAX else lo, end_time=None if hi == FMAX else hi)


CPU_SAMPLE_SECONDS = 9.0  # one-core seconds of reference work per process and pass, both arms


def run_reference_arm(args, rank, world):
    """``--impl reference``: the reference's own CPU implementation on the host cores."""
    if rank != 0:
        return
    w = dict(WORKLOADS[args.workload], input=args.input)
    kind, which = cpu_kind()
    from oracle import oracle

    oracle._load(which)  # in the parent, before any fork: the pool's workers inherit the mapped library
    ts = make_input(w)
    E, N = len(ts["edges"]["left"]), len(ts["nodes"]["time"])
    cores = os.cpu_count() or 1
    expected = expected_mutations(ts, w)
    sample, frac = reference_sample(ts, w, expected, CPU_SAMPLE_SECONDS)
    for i in range(args.warmup):
        cpu_reference_pass(sample, w, cores, 1000 + 100 * i, which)
    muts = 0
    wall = 0.0
    for i in range(args.steps):
        m, _, t = cpu_reference_pass(sample, w, cores, 5000 + 100 * i, which)
        muts += m
        wall += t
    value = muts / wall
    # the WHOLE workload once, one process (the reference is single-threaded): no extrapolation
    full = None
    if not args.no_full_config:
        m, inner, t = cpu_reference_pass(ts, w, 1, 9000, which)
        full = {"mutations": m, "seconds_mutgen_generate": inner, "seconds_wall": t, "value": m / t, "unit": UNIT,
                "cores": 1}
    sample_txt = (f"genome prefix [0, {sample['sequence_length']:.0f}) = {frac:.4f} of the workload, "
                  f"{cores} processes x 1 overlay each per step, MT19937 seeds differ per process")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall / max(1, args.steps),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(w, E, N, max(1, args.gpus)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample_txt},
        "full_config_one_core": full,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # noqa: BLE001
            log(f"[bench] NVML unavailable: {e}")

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            "hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
            "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
            "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
            "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap,
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:  # noqa: BLE001
                pass
            self._stop_evt.wait(0.02)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------ roofline bookkeeping
def algorithmic_bytes(kernel, n):
    """Compulsory bytes of ONE launch of ``kernel`` (DESIGN.md, section "Kernels").
    n: E edges, R records entering the sort, M mutation rows, S site rows, DB/AB
    derived/ancestral state bytes."""
    E, R, M, S = n["E"], n["R"], n["M"], n["S"]
    table = {
        # left,right f64 + parent,child i32 + two node-time gathers in, goff u32 out
        "k1_count": E * (8 + 8 + 4 + 4 + 16) + E * 4,
        # goff in, edge columns of edges that mutate (<= E), Rec 16 B + key 4 B out
        "k2_place": E * 4 + min(E, M) * (8 + 8 + 4 + 4 + 16) + M * (16 + 4),
        "rs_global_hist": R * n["KB"],
        # key + u32 payload read and written once per pass (first pass has no payload read)
        "rs_onesweep": R * 2 * (n["KB"] + 4),
        # key in, site row out
        "p1_sites": R * (n["KB"] + 4),
        # key in (records of multi-mutation sites are a small fraction)
        "k4_resolve": R * n["KB"],
        # key, slot, site row and the 16-byte record in (the record is gathered through the sorted slot: the
        # 32-byte DRAM sector it sits in is traffic, not algorithm); table rows + ragged bytes out
        # (offset columns are uint32; all-zero metadata offsets are not rewritten)
        "k5_emit": R * (n["KB"] + 4 + 4 + 16) + M * (4 + 4 + 4 + 8 + 4) + S * (8 + 4) + n["DB"] + n["AB"],
    }
    return table.get(kernel)


# ------------------------------------------------------------------ the B200 arm
def pinned_like(a):
    """A pinned-host copy of numpy array ``a`` (torch owns the buffer)."""
    import torch

    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return t, t.numpy()


def pinned_empty(n, dtype):
    import torch

    tdt = {np.float64: torch.float64, np.int32: torch.int32, np.uint8: torch.uint8, np.uint64: torch.int64,
           np.uint32: torch.int32}[dtype]
    t = torch.empty(int(n), dtype=tdt, pin_memory=True)
    a = t.numpy()
    return t, (a.view(dtype) if dtype in (np.uint64, np.uint32) else a)


def bind_near_gpu(index):
    """Best effort: run this process on the CPUs of the NUMA node its GPU hangs off, so that the
    pinned host buffers (first touched below) sit on that node and the DMA does not cross sockets.
    Returns (original affinity, description); changes nothing when NVML or the cgroup says no."""
    original = None
    try:
        original = os.sched_getaffinity(0)
        import pynvml

        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, word in enumerate(words) for b in range(64) if (word >> b) & 1}
        cpus &= original
        if cpus and cpus != original:
            os.sched_setaffinity(0, cpus)
            return original, f"{len(cpus)} of {len(original)} cpus (NVML affinity of GPU {index})"
        return original, "unchanged (the GPU's CPUs are all of this process's CPUs)"
    except Exception as e:  # noqa: BLE001
        return original, f"unchanged ({type(e).__name__})"


def python_call_timing(ts, w, map_pos, map_rate, device):
    """Wall time of ``msprime_b200.sim_mutations`` on the bench workload (TreeSequence in, TreeSequence out), and of
    its steps: table dict, upload + index, overlay, download into fresh numpy arrays, result object.  tskit is
    not installed on the GPU box, so the tree sequence is the package's own stand-in for it (same dict protocol;
    with real tskit the call adds ``TableCollection.fromdict`` + ``tree_sequence()``, msprime/mutations.py:1699-1702)."""
    import msprime_b200 as msp
    from msprime_b200 import intervals, mutations, tables as mtables

    name = w["model"]
    model = (mutations.InfiniteAlleles() if name == "infinite_alleles" else mutations.JC69() if name == "jc69" else
             mutations.BinaryMutationModel() if name == "binary" else mutations.HKY(kappa=2) if name == "hky" else
             mutations.GTR(relative_rates=[1, 2, 1, 1, 2, 1], equilibrium_frequencies=[0.3, 0.2, 0.2, 0.3]))
    rate = w["rate"] if len(map_rate) == 1 else intervals.RateMap(position=map_pos, rate=map_rate)
    lo, hi = time_window(w)
    kw = dict(rate=rate, model=model, discrete_genome=w["discrete"], keep=bool(w.get("keep_layer")),
              start_time=None if lo == -FMAX else lo, end_time=None if hi == FMAX else hi, device=device)
    tsobj = mtables.TableCollection.fromdict(ts).tree_sequence()
    from msprime_b200 import _native

    walls, pool = [], []
    out = None
    for i in range(14):  # the first calls create the context, size its buffers and pin the result blocks
        t0 = time.perf_counter()
        out = msp.sim_mutations(tsobj, random_seed=11 + i, **kw)
        walls.append(time.perf_counter() - t0)
        pool.append(_native.pinned_stats())
    new = out.num_mutations - (len(ts["mutations"]["site"]) if w.get("keep_layer") else 0)
    # the same steps by hand on the call's own context
    ctx = mutations._context(device)
    from msprime_b200 import _native

    steps = {}
    res = None
    for _ in range(2):  # (the second pass finds the pinned blocks of the first in the pool, as a steady caller does)
        res = None
        t0 = time.perf_counter()
        d, kind = mtables.from_tree_sequence(tsobj)
        steps["tables_dict"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        ctx.load_tables(d)
        steps["load_tables_from_pageable"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        prm = _native.Params(map_position=map_pos, map_rate=map_rate, model=model._native(),
                             discrete_genome=w["discrete"], keep=bool(w.get("keep_layer")), start_time=lo, end_time=hi,
                             seed=99)
        steps["params"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        ctx.generate(prm)
        steps["generate"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        res = ctx.fetch()
        steps["fetch_into_pooled_pinned"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        mtables.to_tree_sequence(mutations._merge_result(d, res), kind, None)
        steps["result_object"] = time.perf_counter() - t0
    return {"wall_ms_first_call": round(1e3 * walls[0], 2), "wall_ms": round(1e3 * min(walls[1:]), 2),
            "wall_ms_median_of_last_four": round(1e3 * float(np.median(walls[-4:])), 2),
            "wall_ms_all": [round(1e3 * x, 2) for x in walls],
            "pinned_pool_after_each_call": [list(p) for p in pool],
            "value": new / min(walls[1:]), "unit": UNIT,
            "steps_ms": {k: round(1e3 * v, 2) for k, v in steps.items()},
            "tree_sequence": "msprime_b200.tables stand-in (tskit is not installed here)"}


def oneshot_call_timing(ts, w, map_pos, map_rate, device):
    """Wall time of `mspb_sim_mutations`, the ONE call INTEGRATION.md's stub binds where the reference calls
    mutgen_alloc .. mutgen_generate (msprime/_msprimemodule.c:3047-3120): tables in PAGEABLE host memory in,
    library-owned host columns out, released with mspb_columns_free."""
    import ctypes as C

    from msprime_b200 import _native

    lib = _native.lib()
    t, hold = _native.Context._tables_struct(ts)
    lo, hi = time_window(w)
    walls, new = [], 0
    for i in range(8):
        p = _native.Params(map_position=map_pos, map_rate=map_rate, model=native_model(w["model"]),
                           discrete_genome=w["discrete"], keep=bool(w.get("keep_layer")), start_time=lo, end_time=hi,
                           seed=2000 + i, offsets32=True)
        c, cols = _native._Counts(), _native._Columns()
        t0 = time.perf_counter()
        _native.check(lib.mspb_sim_mutations(int(device), C.byref(t), C.byref(p.struct), C.byref(c), C.byref(cols)))
        lib.mspb_columns_free(C.byref(cols))
        walls.append(time.perf_counter() - t0)
        new = int(c.num_new_mutations)
    del hold
    lib.mspb_oneshot_release(int(device))
    steady = float(np.median(walls[-4:]))
    return {"wall_ms_first_call": round(1e3 * walls[0], 2), "wall_ms_median_of_last_four": round(1e3 * steady, 2),
            "value": new / steady, "unit": UNIT,
            "path": "mspb_sim_mutations + mspb_columns_free, pageable host tables in, pooled pinned columns out"}


def run_b200_arm(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist

    from msprime_b200 import _native, build, sharding

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    if rank == 0:
        build.build()
    stdout_fd = None
    if world > 1:
        # NCCL prints its version banner on stdout: park stdout on stderr until the JSON line
        sys.stdout.flush()
        stdout_fd = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist.barrier()
    torch.cuda.set_device(local_rank)
    w = dict(WORKLOADS[args.workload], input=args.input)
    ts = make_input(w)
    affinity0, binding = bind_near_gpu(local_rank)
    log(f"[bench] rank {rank}: cpu binding {binding}")
    L = float(ts["sequence_length"])
    map_pos, map_rate = rate_map(w, L)
    # slices of equal expected mutation mass (branch length x rate), not of equal length
    lo_t, hi_t = time_window(w)
    cuts = sharding.balanced_cuts(ts, map_pos, map_rate, world, discrete_genome=w["discrete"],
                                  start_time=None if lo_t == -FMAX else lo_t, end_time=None if hi_t == FMAX else hi_t)
    sl, sr = (0.0, 0.0) if world == 1 else (cuts[rank], cuts[rank + 1])

    # host inputs in pinned memory (what a caller hands to the C ABI)
    hold = []
    host_ts = {"sequence_length": L, "time_units": ts["time_units"], "nodes": {}, "edges": {},
               "sites": ts["sites"], "mutations": ts["mutations"]}
    t, host_ts["nodes"]["time"] = pinned_like(ts["nodes"]["time"])
    hold.append(t)
    for k in ("left", "right", "parent", "child"):
        t, host_ts["edges"][k] = pinned_like(ts["edges"][k])
        hold.append(t)
    E, N = len(ts["edges"]["left"]), len(ts["nodes"]["time"])
    # node and edge columns: the whole of them at one GPU, this rank's 1 / world share otherwise
    # (sharded upload; the shares are padded to equal sizes); kept rows are copied whole
    h2d_bytes = sum(-(-n // world) * w for n, w in ((E, 8), (E, 8), (E, 4), (E, 4), (N, 8)))
    for tbl in ("sites", "mutations"):  # kept rows of a keep=True workload
        h2d_bytes += sum(int(np.asarray(v).nbytes) for v in ts[tbl].values())

    ctx = _native.Context(local_rank)
    ctx.force_k2({"auto": 0, "dense": 1, "sparse": 2}[args.k2])
    stream = torch.cuda.ExternalStream(ctx.stream(), device=torch.device("cuda", local_rank))

    pcache = {}

    def params(step):
        # argument marshalling (numpy -> C struct) is host preparation: the timed loops below
        # find their per-step parameter blocks (one seed each) ready
        if step not in pcache:
            pcache[step] = make_params(step)
        return pcache[step]

    def make_params(step):
        return _native.Params(map_position=map_pos, map_rate=map_rate, model=native_model(w["model"]),
                              discrete_genome=w["discrete"], keep=bool(w.get("keep_layer")),
                              start_time=time_window(w)[0], end_time=time_window(w)[1],
                              seed=1000 + step, slice_left=sl, slice_right=sr, offsets32=True,
                              slice_stream=0 if world == 1 else rank + 1)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def reduce(x, op):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=op)
        return float(t.item())

    # counts are exchanged on the device: NCCL all-gathers enqueued on the context's stream
    gather = sharding.nccl_device_allgather(ctx, world) if world > 1 else None

    def overlay(step, fetch_into=None):
        """One overlay of this rank's slice; returns the number of new mutations it holds."""
        if world == 1:
            c = ctx.generate(params(step))
            if fetch_into is not None:
                ctx.fetch(fetch_into)
            return c["num_new_mutations"]
        sharding.run_slice_device(ctx, params(step), rank=rank, world=world, gather=gather,
                                  fetch=fetch_into is not None, out=fetch_into)
        return ctx.counts["num_new_mutations"]

    def load():
        """This step's input tables, host -> device (+ index build).  With several GPUs the tables
        are replicated: every rank uploads 1 / world of them and NCCL all-gathers the shares."""
        if world == 1:
            ctx.load_tables(host_ts)
        else:
            sharding.load_tables_sharded(ctx, host_ts, rank=rank, world=world)

    load()
    # ---- device-resident throughput
    for i in range(args.warmup):
        overlay(i)
    c = ctx.counts
    sizes = ctx.result_sizes()
    out_bufs = {}
    for name, dtype in _native._COLUMNS:
        if name.endswith("_offset"):
            dtype = np.uint32  # offsets32=True
        t, out_bufs[name] = pinned_empty(int(sizes[name] * 1.3) + 4096, dtype)
        hold.append(t)
    for i in range(args.steps):
        params(100 + i)
        params(300 + i)
    launches0 = ctx.kernel_launches()
    sampler = ClockSampler(local_rank)
    sync_all()
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    muts = 0
    lib_ms = 0.0
    last_timed_new = 0
    for i in range(args.steps):
        last_timed_new = overlay(100 + i)
        muts += last_timed_new
        lib_ms += ctx.counts["device_ms"]
    ev1.record(stream)
    sync_all()
    clocks = sampler.stop()
    dev_ms = reduce(ev0.elapsed_time(ev1), dist.ReduceOp.MAX if world > 1 else None)
    total_muts = reduce(float(muts), dist.ReduceOp.SUM if world > 1 else None)
    launches = ctx.kernel_launches() - launches0
    # per-kernel durations: the same K steps once more with an event pair around every launch
    # (kept out of the loop above: the extra events sit between back-to-back kernels)
    ctx.kernel_timing(True)
    for i in range(args.steps):
        overlay(100 + i)
    sync_all()
    ktimes = ctx.kernel_times()
    phases = ctx.phase_ms()
    ctx.kernel_timing(False)
    value = total_muts / (dev_ms * 1e-3)

    # ---- end to end through the host-buffer calls
    d2h_bytes = 0
    for i in range(min(2, args.warmup)):
        load()
        overlay(200 + i, out_bufs)
    sync_all()
    ev2, ev3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    wall0 = time.perf_counter()
    ev2.record(stream)
    e2e_muts = 0
    e2e_phase = {"load_tables": 0.0, "overlay": 0.0, "fetch": 0.0}  # host wall seconds (every phase ends in a wait)
    for i in range(args.steps):
        t0 = time.perf_counter()
        load()
        t1 = time.perf_counter()
        e2e_muts += overlay(300 + i)
        t2 = time.perf_counter()
        ctx.fetch(out_bufs)
        t3 = time.perf_counter()
        e2e_phase["load_tables"] += t1 - t0
        e2e_phase["overlay"] += t2 - t1
        e2e_phase["fetch"] += t3 - t2
        sizes = ctx.result_sizes()
        d2h_bytes = ctx.last_fetch_bytes()  # what crossed the bus (all-zero metadata offsets are not copied)
    ev3.record(stream)
    sync_all()
    wall = time.perf_counter() - wall0
    # per phase, the slowest rank's average (the phases of different ranks overlap, so they need not add up)
    e2e_phase_ms = {k: reduce(1e3 * v / args.steps, dist.ReduceOp.MAX if world > 1 else None)
                    for k, v in e2e_phase.items()}
    e2e_ms = reduce(max(ev2.elapsed_time(ev3), wall * 1e3), dist.ReduceOp.MAX if world > 1 else None)
    e2e_total = reduce(float(e2e_muts), dist.ReduceOp.SUM if world > 1 else None)
    e2e_value = e2e_total / (e2e_ms * 1e-3)
    h2d_total = reduce(float(h2d_bytes), dist.ReduceOp.SUM if world > 1 else None)
    d2h_total = reduce(float(d2h_bytes), dist.ReduceOp.SUM if world > 1 else None)
    e2e_serial = {"value": e2e_value, "ms_per_step": e2e_ms / args.steps, "phases_ms_slowest_rank": e2e_phase_ms}

    # ---- the same calls with several overlays in flight (N = 1): PCIe is full duplex and the copy engines
    # run beside the kernels, so step i + 1 uploads while step i computes and step i - 1 downloads.  Every step
    # still makes the whole call sequence (mspb_ctx_load_tables + mspb_generate + mspb_fetch, pinned host buffers in
    # and out); the steps are dealt round-robin to `depth` contexts, each driven by its own host thread
    # (the C calls release the GIL), and K steps are timed from the first call to the last return.
    e2e_pipe = None
    depth = max(1, args.e2e_depth)
    if world == 1 and depth > 1:
        ctxs = [ctx] + [_native.Context(local_rank) for _ in range(depth - 1)]
        bufs = [out_bufs]
        for _ in range(depth - 1):
            b = {}
            for name, dtype in _native._COLUMNS:
                if name.endswith("_offset"):
                    dtype = np.uint32
                t, b[name] = pinned_empty(len(out_bufs[name]), dtype)
                hold.append(t)
            bufs.append(b)
        for i in range(600, 600 + args.steps + depth):
            params(i)
        done = [0] * depth

        def pipe_worker(k, first, count, barrier):
            torch.cuda.set_device(local_rank)
            barrier.wait()
            for i in range(first + k, first + count, depth):
                ctxs[k].load_tables(host_ts)
                c = ctxs[k].generate(params(i))
                ctxs[k].fetch(bufs[k])
                done[k] += c["num_new_mutations"]

        def run_pipe(first, count):
            barrier = threading.Barrier(depth + 1)
            ths = [threading.Thread(target=pipe_worker, args=(k, first, count, barrier)) for k in range(depth)]
            for t in ths:
                t.start()
            torch.cuda.synchronize()
            barrier.wait()
            t0 = time.perf_counter()
            for t in ths:
                t.join()
            torch.cuda.synchronize()
            return time.perf_counter() - t0

        run_pipe(600, depth)  # warm-up: every context sizes its buffers
        for k in range(depth):
            done[k] = 0
        wall_p = run_pipe(600 + depth, args.steps)
        e2e_pipe = {"value": sum(done) / wall_p, "ms_per_step": 1e3 * wall_p / args.steps, "depth": depth}
        for c in ctxs[1:]:
            c.close()
        e2e_ms = 1e3 * wall_p
        e2e_value = e2e_pipe["value"]

    # ---- N > 1: every rank checks its own slice of one more overlay (outside every timed region): the slice as a
    # table of its own (ids and offsets minus the slice's bases) obeys the table invariants, its mutations lie inside
    # the slice and on branches of their nodes, and their number fits the Poisson mean of the slice
    sliced_validation = None
    if world > 1 and not args.no_validate and not w.get("keep_layer"):
        from msprime_b200 import validate

        t0 = time.perf_counter()
        ok, err_text, z = 1.0, "", 0.0
        local, info = sharding.run_slice_device(ctx, params(100 + args.steps - 1), rank=rank, world=world,
                                                gather=gather, fetch=True)
        try:
            res = {"sites": dict(local["sites"]), "mutations": dict(local["mutations"])}
            mm = res["mutations"]
            mm["site"] = np.asarray(mm["site"]).astype(np.int64) - info["site_base"]
            par = np.asarray(mm["parent"]).astype(np.int64)
            mm["parent"] = np.where(par >= 0, par - info["mutation_base"], -1)
            for tbl, cols in (("sites", ("ancestral_state", "metadata")), ("mutations", ("derived_state", "metadata"))):
                for col in cols:
                    off = np.asarray(res[tbl][col + "_offset"]).astype(np.int64)
                    res[tbl][col + "_offset"] = off - (off[0] if len(off) else 0)
            sp = np.asarray(res["sites"]["position"])
            assert len(sp) == 0 or (sp[0] >= sl and sp[-1] < sr), "a site outside the slice"
            validate.check_result(ts, res, discrete=w["discrete"])
            e = ts["edges"]
            clipped = {"sequence_length": ts["sequence_length"], "nodes": ts["nodes"],
                       "edges": {"left": np.clip(e["left"], sl, sr), "right": np.clip(e["right"], sl, sr),
                                 "parent": e["parent"], "child": e["child"]}}
            _, z = validate.check_count(clipped, len(mm["site"]), map_pos, map_rate, discrete=w["discrete"])
        except AssertionError as ex:
            ok, err_text = 0.0, f"rank {rank}: {type(ex).__name__}: {ex}"[:200]
        all_ok = reduce(ok, dist.ReduceOp.MIN)
        worst_z = reduce(abs(float(z)), dist.ReduceOp.MAX)
        total_new = reduce(float(len(local["mutations"]["site"])), dist.ReduceOp.SUM)
        sliced_validation = {"ok": bool(all_ok == 1.0), "slices_checked": world, "num_new_mutations": int(total_new),
                             "largest_count_z": round(worst_z, 3), "seconds": round(time.perf_counter() - t0, 1)}
        if err_text:
            log(f"[bench] slice validation: {err_text}")

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (rank 0's launches)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:  # noqa: BLE001
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    kt = sorted(ktimes.items(), key=lambda kv: -kv[1][1])
    total_kernel_ms = sum(v[1] for v in ktimes.values()) or 1.0
    last = ctx.counts
    nn = {"E": E, "R": last["num_mutations"], "M": last["num_mutations"],
          "S": last["num_sites"], "DB": last["derived_state_bytes"], "AB": last["ancestral_state_bytes"],
          "KB": 4 if (w["discrete"] and L <= 2**32) else 8}
    roofline = None
    # dram bytes per launch from the committed `ncu --set full` capture of this workload at N = 1
    traffic = {}
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if tj.get("workload") == args.workload and world == 1:
            traffic = tj.get("kernels", {})
    except Exception:  # noqa: BLE001
        pass
    for name, (n_launch, ms) in kt:
        ab = algorithmic_bytes(name, nn)
        if ab is None:
            continue
        avg_ms = ms / n_launch
        achieved = ab / (avg_ms * 1e-3) / 1e9
        roofline = {"bound": "hbm", "kernel": name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": traffic.get(name), "peak_source": peak_src,
                    "avg_launch_ms": avg_ms, "launches_per_step": n_launch / args.steps,
                    "share_of_kernel_time": ms / total_kernel_ms, "algorithmic_bytes_per_launch": ab}
        break
    kernels = {k: {"launches": v[0], "ms_total": round(v[1], 4)} for k, v in kt[:12]}
    if roofline is not None and roofline["kernel"] == "k5_emit":
        # SURVEY.md section 8(d) counts K5's OUTPUT terms only; reported beside the figure above, which
        # also charges the kernel's compulsory inputs (key, slot, site row, 16-byte record)
        out_only = nn["M"] * (4 + 4 + 4 + 8 + 4) + nn["S"] * (8 + 4) + nn["DB"] + nn["AB"]
        roofline["frac_output_terms_only"] = out_only / (roofline["avg_launch_ms"] * 1e-3) / 1e9 / peak
    # whole overlay against the same peak: SURVEY.md section 8(d) A_bytes / device time per overlay (N = 1)
    roofline_e2e = None
    if world == 1:
        kept_bytes = sum(int(np.asarray(ts[t][c]).nbytes) for t, c in (
            ("sites", "ancestral_state"), ("mutations", "derived_state"))) if w.get("keep_layer") else 0
        ob = overall_bytes(E, nn["M"], nn["S"], nn["DB"], nn["AB"],
                           M0=len(ts["mutations"]["site"]) if w.get("keep_layer") else 0, kept_bytes=kept_bytes)
        step_s = dev_ms / args.steps * 1e-3
        roofline_e2e = {"bound": "hbm", "unit": "GB/s", "peak": peak,
                        "a_bytes_survey": ob["survey"], "a_bytes_as_written": ob["as_written"],
                        "achieved_survey": ob["survey"] / step_s / 1e9, "frac_survey": ob["survey"] / step_s / 1e9 / peak,
                        "achieved_as_written": ob["as_written"] / step_s / 1e9,
                        "frac_as_written": ob["as_written"] / step_s / 1e9 / peak}

    # ---- the last timed overlay, checked OUTSIDE the timed region: valid tables, right count, right numbering
    validated = None
    if not args.no_validate:
        from msprime_b200 import validate

        t0 = time.perf_counter()
        step = 100 + args.steps - 1  # seed of the last timed overlay
        validated = {"step_seed": 1000 + step, "ok": False}
        try:
            if world == 1:
                cnt = ctx.generate(params(step))
                res = ctx.fetch()
                lo, hi = time_window(w)
                win = dict(start_time=None if lo == -FMAX else lo, end_time=None if hi == FMAX else hi)
                M0 = len(ts["mutations"]["site"]) if w.get("keep_layer") else 0
                validated["num_new_mutations"] = int(cnt["num_new_mutations"])
                validated["same_count_as_timed_step"] = bool(cnt["num_new_mutations"] == last_timed_new)
                validate.check_result(ts, res, discrete=w["discrete"], kept_mutations=M0, **win)
                validated["check_result"] = True
                mean, z = validate.check_count(ts, cnt["num_new_mutations"], map_pos, map_rate, discrete=w["discrete"],
                                               **win)
                validated["expected_new_mutations"] = mean
                validated["count_z"] = z
                if w["model"] == "infinite_alleles" and M0 == 0:
                    validate.check_infinite_alleles_numbering(res, 0)
                    validated["infinite_alleles_numbering"] = True
                validated["ok"] = bool(validated["same_count_as_timed_step"])
            elif sliced_validation is not None:
                validated.update(sliced_validation)
                validated["how"] = ("every rank checked its own slice (table invariants, branches, count against the "
                                    "slice's Poisson mean); tests/test_gpu_multi.py checks merged tables bit for bit")
            else:
                validated["skipped"] = "N > 1 with kept rows: tests/test_gpu_multi.py checks the merged tables"
        except AssertionError as e:
            validated["error"] = f"{type(e).__name__}: {e}"[:300]
        if world == 1:
            validated["seconds"] = round(time.perf_counter() - t0, 1)

    # ---- the consumers of the resident tables, outside every timed region: the sample-count
    # index is built once per loaded tree sequence, the site AFS is one query per overlay
    consumers_info = None
    if world == 1 and not args.no_consumers:
        try:
            first_build = ctx.build_sample_index(ts["nodes"]["flags"])["build_ms"]  # sizes its buffers
            info = ctx.build_sample_index(ts["nodes"]["flags"])
            info["build_ms_first"] = first_build
            cnt = ctx.generate(params(100 + args.steps - 1))
            ctx.site_afs()
            first_afs = ctx.afs_ms  # (allocates the per-row scratch)
            afs = ctx.site_afs()
            consumers_info = {"sample_index": info, "site_afs_ms": round(ctx.afs_ms, 4),
                              "site_afs_ms_first": round(first_afs, 4),
                              "mutations": int(cnt["num_mutations"]), "afs_total": int(afs.sum()),
                              "afs_head": [int(v) for v in afs[:6]]}
        except Exception as e:  # noqa: BLE001
            consumers_info = {"error": f"{type(e).__name__}: {e}"[:200]}

    # ---- the call the north star names, msprime_b200.sim_mutations(tree_sequence, ...) -> tree sequence, with the
    # caller's plain (pageable) numpy columns in and fresh numpy columns out; wall time, and the same steps one by one
    e2e_python = None
    if world == 1 and not args.no_e2e_python:
        try:
            e2e_python = python_call_timing(ts, w, map_pos, map_rate, local_rank)
        except Exception as e:  # noqa: BLE001
            e2e_python = {"error": f"{type(e).__name__}: {e}"[:200]}
    e2e_oneshot = None
    if world == 1 and not args.no_e2e_python:
        try:
            e2e_oneshot = oneshot_call_timing(ts, w, map_pos, map_rate, local_rank)
        except Exception as e:  # noqa: BLE001
            e2e_oneshot = {"error": f"{type(e).__name__}: {e}"[:200]}

    # ---- CPU baseline (N = 1 only): the compiled reference on all host cores
    cpu = None
    if affinity0 is not None:
        try:
            os.sched_setaffinity(0, affinity0)  # the CPU baseline below uses every core
        except OSError:
            pass
    if world == 1 and not args.no_cpu_baseline:
        kind, which = cpu_kind()
        from oracle import oracle

        oracle._load(which)
        cores = os.cpu_count() or 1
        sample, frac = reference_sample(ts, w, expected_mutations(ts, w), CPU_SAMPLE_SECONDS)
        m, inner, wall_cpu = cpu_reference_pass(sample, w, cores, 7000, which)
        cpu = {"value": m / wall_cpu, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": f"genome prefix [0, {sample['sequence_length']:.0f}) = {frac:.4f} of the workload; "
                         f"{cores} processes x 1 overlay each ({m} mutations in {wall_cpu:.1f} s wall, "
                         f"slowest mutgen_generate {inner:.1f} s)"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(w, E, N, world),
        "mutations_per_step": total_muts / args.steps,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_total, "d2h_bytes_per_step": d2h_total,
                "ms_per_step": e2e_ms / args.steps,
                "one_at_a_time": e2e_serial,
                "in_flight": e2e_pipe["depth"] if e2e_pipe else 1,
                "path": ("mspb_ctx_load_tables + mspb_generate + mspb_fetch, pinned host buffers"
                         + (f"; {e2e_pipe['depth']} overlays in flight (one context and host thread each): upload of "
                            "step i + 1, kernels of step i and download of step i - 1 overlap" if e2e_pipe else ""))
                if world == 1 else
                "per rank: mspb_ctx_load_tables_begin (1 / world of the replicated tables over PCIe) / in-place NCCL "
                "all-gather of the shares over NVLink / mspb_ctx_load_tables_end, then mspb_generate_begin / count "
                "all-gather (NCCL) / mspb_generate_end / byte all-gather / mspb_shift_offsets + mspb_fetch of the "
                "rank's slice into pinned host buffers"},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": roofline,
        "roofline_e2e": roofline_e2e,
        "validated": validated,
        "consumers": consumers_info,
        "e2e_python": e2e_python,
        "e2e_oneshot": e2e_oneshot,
        "cpu_baseline": cpu,
        "phases_ms_last_step": {k: round(v, 4) for k, v in phases.items()},
        "kernels": kernels,
        "library_device_ms_per_step": lib_ms / args.steps,
    }
    if stdout_fd is not None:
        sys.stdout.flush()
        os.dup2(stdout_fd, 1)
        os.close(stdout_fd)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="config5", choices=sorted(WORKLOADS))
    ap.add_argument("--input", default="auto", choices=["auto", "treegen", "reference"],
                    help="where the tree sequence comes from: the reference's own ancestry simulator where its output "
                         "has been prepared (auto, the default; scripts/make_reference_input.py), else this project's "
                         "SMC' generator; 'reference' also runs the simulator on the spot, 'treegen' never uses it")
    ap.add_argument("--k2", default="auto", choices=["auto", "dense", "sparse"],
                    help="placement kernel: by density (default), one block per 256 edges, one thread per mutation")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-validate", action="store_true", help="skip the checks of the last timed overlay")
    ap.add_argument("--no-consumers", action="store_true", help="skip the sample-count index / site AFS figures")
    ap.add_argument("--no-e2e-python", action="store_true", help="skip the timing of the Python-level call")
    ap.add_argument("--e2e-depth", type=int, default=3,
                    help="end-to-end overlays in flight at N = 1 (1 = one at a time)")
    ap.add_argument("--no-full-config", action="store_true",
                    help="reference arm: skip the single-process overlay of the whole workload")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    args.warmup = max(args.warmup, 3)
    run_b200_arm(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
