import os, sys, time
sys.path.insert(0, os.getcwd())
import numpy as np, torch, torch.distributed as dist
from msprime_b200 import _native, sharding, synth
rank=int(os.environ["RANK"]); world=int(os.environ["WORLD_SIZE"])
dist.init_process_group("nccl", device_id=torch.device("cuda", rank)); torch.cuda.set_device(rank)
ts=synth.native_treeseq(1000000,2.5e8,1e-8,1e4,1); L=ts["sequence_length"]
ctx=_native.Context(rank); ctx.load_tables(ts)
cuts=sharding.slice_cuts(L,world)
ag=sharding.dist_allgather(device=f"cuda:{rank}")
def P(step): return _native.Params(map_position=[0,L],map_rate=[1e-7],model={"kind":1,"start_allele":0},discrete_genome=True,keep=False,start_time=-1e308,end_time=1e308,seed=step,slice_left=cuts[rank],slice_right=cuts[rank+1],offsets32=True,slice_stream=rank+1)
T={}
def tick(name,t0):
    T[name]=T.get(name,0)+time.perf_counter()-t0
for step in range(13):
    if step==3: T.clear()
    t=time.perf_counter(); p=P(step); tick("params",t)
    t=time.perf_counter(); c1=ctx.generate_begin(p); tick("begin",t)
    t=time.perf_counter(); a1=ag([c1["num_sites"],c1["num_mutations"],c1["next_allele"],c1["num_new_mutations"]]); tick("ag1",t)
    sb,mb,ab,_=sharding.exclusive_bases(a1,rank)
    t=time.perf_counter(); c2=ctx.generate_end(sb,mb,ab); tick("end",t)
    t=time.perf_counter(); a2=ag([c2["ancestral_state_bytes"],0,c2["derived_state_bytes"],0]); tick("ag2",t)
    t=time.perf_counter(); ctx.shift_offsets(*sharding.exclusive_bases(a2,rank)); tick("shift",t)
if rank==0:
    print({k:round(1e3*v/10,3) for k,v in T.items()}, "ms per step; begin device_ms", c1["device_ms"], ctx.phase_ms())
dist.destroy_process_group()
