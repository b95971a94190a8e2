"""Gather kernel variants at config-2 shapes: gather-only (rotating slots) and the pipelined loop.
Env: RETH_GATHER_MODE, RETH_BULK_UNITS, RETH_BULK_CTAS, RETH_BULK_EF_LOADS, RETH_BULK_EF_STORES."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from reth_b200 import ReplayShard

ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
R = 56464
cap = int(os.environ.get("CAP", 1 << 20)); B = int(os.environ.get("B", 512)); depth = int(os.environ.get("DEPTH", 16))
sh = ReplayShard(cap, ATARI, max_batch=B, ring_depth=depth, update_cluster=int(os.environ.get("RETH_CLUSTER", 0)))
g = torch.Generator(device="cuda").manual_seed(0)
chunk = 1 << 14
for st in range(0, cap, chunk):
    n = min(chunk, cap - st)
    s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
    a = torch.zeros(n, dtype=torch.int64, device="cuda"); r = torch.zeros(n, device="cuda")
    w = torch.randn(n, device="cuda", generator=g).abs() * (1 + 9 * (torch.rand(n, device="cuda", generator=g) < 0.05))
    sh.append([s0, a, r, s0, r], w.float(), alpha=0.6)
torch.cuda.synchronize()
prio = torch.randn(4096 * B, device="cuda").abs().float()
tag = " ".join(f"{k[5:]}={v}" for k, v in sorted(os.environ.items()) if k.startswith("RETH_"))
out = []
for label, kw, steps in [("gather-only g1", {"sample": False, "update": False, "serial": True}, 960),
                         ("gather-only g4", {"sample": False, "update": False, "serial": True, "group": 4}, 960),
                         ("gather-only g8", {"sample": False, "update": False, "serial": True, "group": 8}, 960),
                         ("pipelined g1", {}, 960), ("pipelined g4", {"group": 4}, 960), ("pipelined g8", {"group": 8}, 960),
                         ("tree-only g4", {"gather": False, "group": 4}, 960)]:
    sh.step_loop(96, B, prio, **kw); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    sh.step_loop(steps, B, prio, **kw)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / steps
    out.append(f"{label}: {us:.2f} us/step {B/us:.1f} M/s {2*B*R/us/1e3:.0f} GB/s")
print(f"[{tag}] " + " | ".join(out))
