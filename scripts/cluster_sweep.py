"""Update-kernel latency vs cluster size and batch (tree-only, capacity 2^16 and 2^20)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from reth_b200 import ReplayShard
for cap in (1 << 16, 1 << 20):
    w = torch.rand(cap, device="cuda") + 0.01
    for B in (64, 256, 512, 2048):
        row = []
        for cl in (1, 2, 4, 8, 16):
            sh = ReplayShard(cap, [], max_batch=B, store_rows=False, ring_depth=2, update_cluster=cl)
            for st in range(0, cap, 16384):
                sh.append([], w[st:st + 16384])
            prio = torch.rand(200 * B, device="cuda")
            res = []
            for kw in ({"sample": False}, {}):
                sh.step_loop(20, B, prio, gather=False, serial=True, **kw)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); sh.step_loop(200, B, prio, gather=False, serial=True, **kw); e1.record()
                torch.cuda.synchronize()
                res.append(e0.elapsed_time(e1) * 1e3 / 200)
            row.append(f"c{cl}: {res[0]:.1f}/{res[1]:.1f}")
            sh.close()
        print(f"cap 2^{int(np.log2(cap))} B {B}: update / sample+update us  " + "  ".join(row), flush=True)
