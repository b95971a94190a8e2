"""Small fixed workload for ncu: config-2 shard, N serial steps (sample, gather, update)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from reth_b200 import ReplayShard
ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
cap = int(os.environ.get("CAP", 1 << 20)); B = int(os.environ.get("B", 512)); steps = int(os.environ.get("STEPS", 20))
G = int(os.environ.get("GROUP", 1))
sh = ReplayShard(cap, ATARI, max_batch=B, ring_depth=max(4, 2 * G), update_cluster=int(os.environ.get("CLUSTER", 0)))
g = torch.Generator(device="cuda").manual_seed(0)
chunk = 1 << 14
for st in range(0, cap, chunk):
    n = min(chunk, cap - st)
    s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
    a = torch.zeros(n, dtype=torch.int64, device="cuda"); r = torch.zeros(n, device="cuda")
    w = torch.randn(n, device="cuda", generator=g).abs() * (1 + 9 * (torch.rand(n, device="cuda", generator=g) < 0.05))
    sh.append([s0, a, r, s0, r], w.float(), alpha=0.6)
prio = torch.randn(steps * B, device="cuda").abs().float()
torch.cuda.synchronize()
sh.step_loop(steps, B, prio, serial=True, group=G)
torch.cuda.synchronize()
print("done", sh.stats()["kernel_launches"])
