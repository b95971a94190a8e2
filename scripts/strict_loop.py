"""Strict alternation (prefetch group 1: draw k+1 sees write-back k) device loop, Atari-u8 rows, batch 512:
microseconds per iteration with and without the frame gather.  RETH_PDL=0/1 switches programmatic dependent
launch of the tree chain (read once per process, so run the script once per setting)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from reth_b200 import ReplayShard

ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
cap, B, steps = int(os.environ.get("CAP", 1 << 18)), 512, 4000
sh = ReplayShard(cap, ATARI, max_batch=B, ring_depth=4)
g = torch.Generator(device="cuda").manual_seed(0)
for st in range(0, cap, 1 << 14):
    n = min(1 << 14, cap - st)
    s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
    z = torch.zeros(n, device="cuda")
    sh.append([s0, z.long(), z, s0, z], torch.rand(n, device="cuda", generator=g) + 0.01)
prio = torch.rand(steps * B, device="cuda", generator=g) + 0.01
out = {"pdl": os.environ.get("RETH_PDL", "1")}
for label, kw in (("tree_only_us", {"gather": False}), ("with_gather_us", {})):
    sh.step_loop(200, B, prio, **kw)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sh.step_loop(steps, B, prio, **kw)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / steps)
    out[label] = best
out["M_transitions_per_s"] = B / out["with_gather_us"]
print(json.dumps(out))
sh.close()
