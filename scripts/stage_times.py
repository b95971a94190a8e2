import os, sys
os.environ["RETH_DEBUG_TIMING"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from reth_b200 import ReplayShard
cap, B = int(os.environ.get("CAP", 1 << 20)), 512
KEYS = tuple(range(1, 32))
for cluster in (16,):
    for G in (1, 4, 8):
        sh = ReplayShard(cap, [], max_batch=B, store_rows=False, update_cluster=cluster, ring_depth=2 * G)
        w = torch.randn(cap, device="cuda").abs().float() + 0.01
        for st in range(0, cap, 16384): sh.append([], w[st:st+16384])
        prio = torch.rand(256 * B, device="cuda")
        sh.step_loop(20 * G, B, prio, gather=False, serial=True, group=G)
        torch.cuda.synchronize()
        acc = {}
        for rep in range(10):
            sh.step_loop(G, B, prio, gather=False, serial=True, group=G)
            s = sh.debug_stamps()
            for k in KEYS:
                if s[k]: acc[k] = acc.get(k, 0) + s[k] - s[0]
        order = sorted(acc.items(), key=lambda kv: kv[1])
        print(f"cluster {cluster} G {G}:", " ".join(f"{k}:{v // 10}" for k, v in order))
        sh.close()
