import os, sys
os.environ["RETH_DEBUG_TIMING"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from reth_b200 import ReplayShard
cap, B = 1 << 20, 512
for cluster in (8,):
    sh = ReplayShard(cap, [], max_batch=B, store_rows=False, update_cluster=cluster, ring_depth=2)
    w = torch.randn(cap, device="cuda").abs().float() + 0.01
    for st in range(0, cap, 16384): sh.append([], w[st:st+16384])
    prio = torch.rand(64 * B, device="cuda")
    sh.step_loop(20, B, prio, gather=False, serial=True)
    torch.cuda.synchronize()
    acc = None
    for rep in range(10):
        sh.step_loop(1, B, prio, gather=False, serial=True)
        s = sh.debug_stamps()
        d = {k: s[k] - s[0] for k in (1, 2, 11, 12, 13, 3, 4, 15, 17, 7, 8, 14, 16, 5, 6, 20, 23, 24, 22) if s[k]}
        acc = d if acc is None else {k: acc[k] + d.get(k, 0) for k in acc}
    print("cluster", cluster, {k: v // 10 for k, v in acc.items()})
    sh.close()
