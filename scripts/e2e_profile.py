"""Where the host time of the e2e loop goes: per-call wall time of loader.sample() and
client.update_priorities(), split into calls that cross a group boundary (pipe_next / commit)
and calls that do not.  Run on the GPU box:  python scripts/e2e_profile.py [group] [ahead]"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import bench  # noqa: E402
from reth_b200 import service  # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else 8
A = int(sys.argv[2]) if len(sys.argv) > 2 else 1
B, CAP = 512, 1 << 20
proc, addr = service.start_per(CAP, B, alpha=bench.ALPHA, beta=bench.BETA, sample_start=B,
                               max_ring=max(16, (A + 1) * G))
eng = service._ENGINES[addr]
eng._ensure_shard(None, struct=bench.ATARI)
bench.fill_shard(torch, eng.shard, CAP, seed=0)
eng.cnt = CAP
client = service.Client(addr)
loader = service.TorchCudaLoader(addr, buffer_size=(A + 1) * G, prefetch=G, ahead=A)
newp = np.abs(np.random.default_rng(3).standard_normal((64, B))).astype(np.float32)
N = int(os.environ.get("ITERS", 4000))
ts = np.zeros((N, 3))
for i in range(200):
    d, idx, w = loader.sample()
    client.update_priorities(idx, newp[i % 64])
eng.drain()
pc = time.perf_counter
t00 = pc()
for i in range(N):
    t0 = pc()
    d, idx, w = loader.sample()
    t1 = pc()
    client.update_priorities(idx, newp[i % 64])
    t2 = pc()
    ts[i] = (t0, t1, t2)
eng.drain()
tot = pc() - t00
s = (ts[:, 1] - ts[:, 0]) * 1e6
u = (ts[:, 2] - ts[:, 1]) * 1e6
k = np.arange(N) % G
out = {"group": G, "ahead": A, "rate_M": N * B / tot / 1e6, "us_per_batch": tot / N * 1e6,
       "sample_inner_us": float(np.median(s[k != 0])), "sample_boundary_us": float(np.median(s[k == 0])),
       "sample_boundary_p90_us": float(np.percentile(s[k == 0], 90)),
       "update_inner_us": float(np.median(u[k != G - 1])), "update_commit_us": float(np.median(u[k == G - 1])),
       "sum_us_per_group": float(np.median(s[k != 0]) * (G - 1) + np.median(s[k == 0]) + np.median(u[k != G - 1]) * (G - 1)
                                 + np.median(u[k == G - 1]))}
b = s[k == 0]
out.update({"boundary_mean_us": float(b.mean()), "boundary_max_us": float(b.max()), "boundary_p99_us": float(np.percentile(b, 99)),
            "boundary_over_100us": int((b > 100).sum()), "sample_inner_mean_us": float(s[k != 0].mean()), "sample_inner_max_us": float(s[k != 0].max()),
            "update_mean_us": float(u.mean()), "update_max_us": float(u.max()), "n_groups": int(len(b))})
print(json.dumps(out))
loader.close()
proc.terminate()
