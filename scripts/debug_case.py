import sys
import numpy as np
sys.path.insert(0, "tests")
import cases
import test_gpu_replay as T
from msprime_b200 import _native
sys.path.insert(0, ".")
from oracle import oracle as oracle_mod  # noqa
oracle_mod.build(("port",))

ctx = _native.Context(0)
ts = cases.smc_ts(50, 1e5, 1)
L = ts["sequence_length"]
model = cases.MODELS["jc69"]
rng = np.random.default_rng(1)
stream = rng.random(3_000_000)
okw = dict(map_position=[0, L], map_rate=[2e-5], model=model, discrete_genome=True, keep=False,
           start_time=-cases.FMAX, end_time=cases.FMAX)
want = oracle_mod.sim_mutations(ts, which="port", stream=stream, **okw)
got = T.gpu_run(ctx, ts, model=model, stream=stream, map_position=[0, L], map_rate=[2e-5])
print("layer1", T.describe_diff(want, got))
def analyse(want, got):
    a = np.asarray(want["sites"]["ancestral_state"]); b = np.asarray(got["sites"]["ancestral_state"])
    bad = np.flatnonzero(a != b)
    print("bad sites", bad[:50], len(bad))
    ms = np.asarray(want["mutations"]["site"])
    cnt = np.bincount(ms, minlength=len(a))
    print("mutations per bad site", cnt[bad][:50])
    first = np.searchsorted(ms, bad)
    print("first mutation row of bad sites", first[:50], "mod 1024", first[:50] % 1024)
analyse(want, got)
ts2 = cases.with_kept(ts, want)
rng = np.random.default_rng(2)
stream = rng.random(3_000_000)
okw["keep"] = True
want2 = oracle_mod.sim_mutations(ts2, which="port", stream=stream, **okw)
got2 = T.gpu_run(ctx, ts2, model=model, stream=stream, map_position=[0, L], map_rate=[2e-5], keep=True)
print("layer2", T.describe_diff(want2, got2))
analyse(want2, got2)
