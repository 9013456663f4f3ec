"""What the host link of this box can do: one-direction and full-duplex copy rates between pinned host
memory and the GPU, at the byte counts of one end-to-end overlay of the bench workload (272 MB up, 736 MB
down).  The end-to-end floor of bench.py's `e2e` is the slower of the two directions WITH the other one
running.  Usage (under gpurun): python scripts/pcie_duplex.py > gpurun_out/pcie_duplex.json"""
import json
import time

import torch

UP, DOWN = 272_000_000, 736_000_000
dev = torch.device("cuda", 0)
h_up = torch.empty(UP, dtype=torch.uint8, pin_memory=True)
h_dn = torch.empty(DOWN, dtype=torch.uint8, pin_memory=True)
d_up = torch.empty(UP, dtype=torch.uint8, device=dev)
d_dn = torch.empty(DOWN, dtype=torch.uint8, device=dev)
s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)


def run(up, down, reps=10):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if up:
            with torch.cuda.stream(s_up):
                d_up.copy_(h_up, non_blocking=True)
        if down:
            with torch.cuda.stream(s_dn):
                h_dn.copy_(d_dn, non_blocking=True)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


for _ in range(2):
    run(True, True, 2)
out = {"up_bytes": UP, "down_bytes": DOWN, "up_alone_ms": run(True, False), "down_alone_ms": run(False, True),
       "both_ms": run(True, True)}
out["up_alone_gbs"] = UP / out["up_alone_ms"] / 1e6
out["down_alone_gbs"] = DOWN / out["down_alone_ms"] / 1e6
out["both_gbs_sum"] = (UP + DOWN) / out["both_ms"] / 1e6
print(json.dumps(out))
