import sys, time, cProfile, pstats, io
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
import bench
from msprime_b200 import synth, tables as mtables, mutations, _native
import msprime_b200 as msp
w = bench.WORKLOADS["config5"]
ts = synth.native_treeseq(w["samples"], w["L"], w["recomb"], w["ne"], w["ts_seed"])
tsobj = mtables.TableCollection.fromdict(ts).tree_sequence()
out = None
for i in range(4):
    t0 = time.perf_counter(); out = msp.sim_mutations(tsobj, rate=1e-7, model=mutations.InfiniteAlleles(), random_seed=5 + i); print("call", i, round(1e3 * (time.perf_counter() - t0), 1), _native.pinned_stats(), flush=True)
pr = cProfile.Profile(); pr.enable()
out = msp.sim_mutations(tsobj, rate=1e-7, model=mutations.InfiniteAlleles(), random_seed=77)
pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(18); print(s.getvalue()[:3500])
