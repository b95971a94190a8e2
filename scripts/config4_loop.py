"""BASELINE config 4: Ape-X learner loop on one shard -- per learner step, actor messages are
ingested (device-resident producer) and one batch of 512 is sampled, gathered and written back.
Two ingest ratios: Ape-X-like (64 rows per learner step) and stress (256 streams x 64 rows per
step).  alpha 0.5, beta 0.4 (test/apex-dqn/config.yaml:26-27).  Prints one JSON line per case."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from reth_b200 import ReplayShard

ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
R = 56464
cap, B, G = 1 << 20, 512, 8
sh = ReplayShard(cap, ATARI, max_batch=B, ring_depth=2 * G)
g = torch.Generator(device="cuda").manual_seed(0)
for st in range(0, cap, 1 << 14):
    n = 1 << 14
    s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
    z = torch.zeros(n, device="cuda")
    sh.append([s0, z.long(), z, s0, z], torch.rand(n, device="cuda", generator=g) + 0.01, alpha=0.5)
prio = torch.rand(256 * B, device="cuda")
for label, rows_per_step in (("apex-like: 64 rows appended per learner step", 64),
                             ("stress: 256 actor messages x 64 rows per learner step", 256 * 64)):
    n = rows_per_step * G                      # rows ingested per group of G learner steps
    s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
    z = torch.zeros(n, device="cuda")
    cols, w = [s0, z.long(), z, s0, z], torch.rand(n, device="cuda", generator=g)
    groups = 40 if rows_per_step > 64 else 200

    def loop(k):
        for _ in range(k):
            sh.append(cols, w, alpha=0.5)
            sh.step_loop(G, B, prio, alpha=0.5, beta=0.4, group=G)
    loop(5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loop(groups)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (groups * G)
    print(json.dumps({"case": label, "us_per_learner_step": us, "sampled_transitions_per_s": B / us * 1e6,
                      "appended_rows_per_s": rows_per_step / us * 1e6,
                      "hbm_GBps": 2 * R * (B + rows_per_step) / us / 1e3, "prefetch_group": G}))
