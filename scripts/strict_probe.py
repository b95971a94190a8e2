"""Strict alternation (prefetch group 1): device loop with / without the gather, fused repair+draw kernel on / off,
programmatic dependent launch on / off.  Env: RETH_FUSE_DRAW, RETH_PDL.  One JSON line."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from reth_b200 import ReplayShard
ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
cap, B, iters = int(os.environ.get("CAP", 1 << 18)), int(os.environ.get("B", 512)), 2048
sh = ReplayShard(cap, ATARI, max_batch=B, ring_depth=8)
g = torch.Generator(device="cuda").manual_seed(0)
for st in range(0, cap, 1 << 14):
    n = 1 << 14
    s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
    z = torch.zeros(n, device="cuda")
    sh.append([s0, z.long(), z, s0, z], torch.rand(n, device="cuda", generator=g) + 0.01)
prio = torch.rand(iters * B, device="cuda") + 0.01
out = {"fuse": os.environ.get("RETH_FUSE_DRAW", "1"), "pdl": os.environ.get("RETH_PDL", "1"), "batch": B}
for name, kw in (("with_gather", {}), ("tree_only", {"gather": False})):
    sh.step_loop(iters, B, prio, seed=1, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        sh.step_loop(iters, B, prio, seed=1, **kw)
    e1.record(); torch.cuda.synchronize()
    out[name + "_us"] = e0.elapsed_time(e1) * 1e3 / (3 * iters)
print(json.dumps(out))
