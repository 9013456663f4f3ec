"""What the two-phase (sharded) call costs beyond the one-call overlay, on ONE GPU with world = 1: the same
kernels, plus the protocol (generate_begin / exchange / generate_end_device / exchange / generate_finish_device), with
the all-gather replaced by a device copy on the context's stream.  Prints per-phase host times and the device time
per overlay for both paths (config 5 unless --workload says otherwise)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from msprime_b200 import _native, sharding  # noqa: E402


def main():
    import torch

    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="config5")
    ap.add_argument("--steps", type=int, default=20)
    args = ap.parse_args()
    w = dict(bench.WORKLOADS[args.workload])
    ts = bench.make_input(w)
    L = ts["sequence_length"]
    pos, rate = bench.rate_map(w, L)
    ctx = _native.Context(0)
    ctx.load_tables(ts)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=torch.device("cuda", 0))
    lo, hi = bench.time_window(w)

    def params(seed, sliced):
        return _native.Params(map_position=pos, map_rate=rate, model=bench.native_model(w["model"]),
                              discrete_genome=w["discrete"], keep=False, start_time=lo, end_time=hi, seed=seed,
                              slice_left=0.0, slice_right=L if sliced else 0.0, offsets32=True,
                              slice_stream=1 if sliced else 0)

    send, recv = ctx.exchange_tensors(1)

    streams = {}

    def gather():
        h = ctx.exchange_stream()  # the context's stream, or its second one while the exchanges run beside K4
        if h not in streams:
            streams[h] = torch.cuda.ExternalStream(h, device=torch.device("cuda", 0))
        with torch.cuda.stream(streams[h]):
            recv[:send.numel()].copy_(send, non_blocking=True)

    def timed(fn, n):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        t0 = time.perf_counter()
        for i in range(n):
            fn(i)
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n, (time.perf_counter() - t0) * 1e3 / n

    pc = {s: params(1000 + s, False) for s in range(args.steps + 5)}
    ps = {s: params(1000 + s, True) for s in range(args.steps + 5)}
    for i in range(5):
        ctx.generate(pc[i])
    one = timed(lambda i: ctx.generate(pc[i]), args.steps)
    phases = {k: 0.0 for k in ("begin", "post", "gather1", "end", "gather2", "finish")}

    begin = [ctx.generate_begin]

    def two_phase(i):
        t = time.perf_counter()
        begin[0](ps[i]); t1 = time.perf_counter(); phases["begin"] += t1 - t
        ctx.exchange_post_counts(); t2 = time.perf_counter(); phases["post"] += t2 - t1
        gather(); t3 = time.perf_counter(); phases["gather1"] += t3 - t2
        ctx.generate_end_device(0, 1); t4 = time.perf_counter(); phases["end"] += t4 - t3
        gather(); t5 = time.perf_counter(); phases["gather2"] += t5 - t4
        ctx.generate_finish_device(0, 1); t6 = time.perf_counter(); phases["finish"] += t6 - t5

    for i in range(5):
        two_phase(i)
    for k in phases:
        phases[k] = 0.0
    two = timed(two_phase, args.steps)
    host_phases = {k: v * 1e3 / args.steps for k, v in phases.items()}
    begin[0] = ctx.generate_begin_device
    for i in range(5):
        two_phase(i)
    for k in phases:
        phases[k] = 0.0
    dev = timed(two_phase, args.steps)
    out = {"workload": args.workload, "one_call": {"device_ms": one[0], "wall_ms": one[1]},
           "two_phase_world1": {"device_ms": two[0], "wall_ms": two[1], "host_ms_per_phase": host_phases},
           "two_phase_world1_begin_device": {"device_ms": dev[0], "wall_ms": dev[1],
                                             "host_ms_per_phase": {k: v * 1e3 / args.steps for k, v in phases.items()}},
           "library_device_ms_last": ctx.counts["device_ms"]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
