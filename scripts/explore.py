"""GPU exploration: per-kernel durations and step-loop throughput at config 2 shapes."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from reth_b200 import ReplayShard

ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
R = 56464
cap = int(os.environ.get("CAP", 1 << 20)); B = int(os.environ.get("B", 512))
depth = int(os.environ.get("DEPTH", 8)); cluster = int(os.environ.get("CLUSTER", 8))
sh = ReplayShard(cap, ATARI, max_batch=B, ring_depth=depth, update_cluster=cluster)
g = torch.Generator(device="cuda").manual_seed(0)
chunk = 1 << 14
t0 = time.time()
for st in range(0, cap, chunk):
    n = min(chunk, cap - st)
    s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
    s1 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
    a = torch.randint(0, 18, (n,), device="cuda", generator=g)
    r = torch.randint(-1, 2, (n,), device="cuda", generator=g).float()
    d = (torch.rand(n, device="cuda", generator=g) < 0.01).float()
    w = torch.randn(n, device="cuda", generator=g).abs() * (1 + 9 * (torch.rand(n, device="cuda", generator=g) < 0.05))
    sh.append([s0, a, r, s1, d], w.float(), alpha=0.6)
torch.cuda.synchronize()
print(f"fill {cap} rows in {time.time()-t0:.1f}s, hbm {sh.stats()['hbm_bytes']/1e9:.1f} GB")

def timeit(fn, n, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t = time.perf_counter(); e0.record()
    for _ in range(n): fn()
    e1.record(); th = time.perf_counter() - t
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n, th * 1e6 / n   # us gpu, us host

N = 200
prio = torch.randn(4096 * B, device="cuda").abs().float()
# append throughput (device source)
na = 16384
cols = [torch.randint(0, 256, (na, 4, 84, 84), dtype=torch.uint8, device="cuda"), torch.zeros(na, dtype=torch.int64, device="cuda"),
        torch.zeros(na, device="cuda"), torch.randint(0, 256, (na, 4, 84, 84), dtype=torch.uint8, device="cuda"), torch.zeros(na, device="cuda")]
wa = torch.rand(na, device="cuda")
g_us, h_us = timeit(lambda: sh.append(cols, wa), 20)
print(f"append {na} rows: {g_us:.1f} us gpu -> {2*na*R/g_us/1e3:.0f} GB/s ({na/g_us:.2f} Mrows/s) host {h_us:.1f} us")
for name, fn in [("sample", lambda: sh.sample(B, slot=0)), ("gather", lambda: sh.gather(0)),
                 ("update_slot", lambda: sh.update_slot(0, prio[:B])),]:
    sh.sample(B, slot=0)
    g_us, h_us = timeit(fn, N)
    extra = f" -> {2*B*R/g_us/1e3:.0f} GB/s" if name == "gather" else ""
    print(f"{name:12s} back-to-back: {g_us:.2f} us/launch gpu, host {h_us:.2f} us{extra}")
# gather over distinct slots (no L2 reuse of output)
for s in range(depth): sh.sample(B, slot=s)
k = [0]
def gat():
    sh.gather(k[0] % depth); k[0] += 1
g_us, h_us = timeit(gat, N)
print(f"gather rotating slots: {g_us:.2f} us -> {2*B*R/g_us/1e3:.0f} GB/s")
for label, kw in [("sample-only", {"gather": False, "update": False, "serial": True}),
                  ("update-only", {"gather": False, "sample": False, "serial": True}),
                  ("gather-only", {"sample": False, "update": False, "serial": True}),
                  ("pipelined", {}), ("pipelined g2", {"group": 2}), ("pipelined g4", {"group": 4}),
                  ("tree-only g4", {"gather": False, "group": 4}), ("gather-only g4", {"sample": False, "update": False, "serial": True, "group": 4}), ("serial", {"serial": True}), ("tree-only", {"gather": False}), ("pipelined+timed", {"time_gather": True})]:
    steps = 1000
    sh.step_loop(48, B, prio, **kw); torch.cuda.synchronize(); sh.gather_time()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t = time.perf_counter(); e0.record()
    sh.step_loop(steps, B, prio, **kw)
    e1.record(); th = time.perf_counter() - t
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / steps
    gt = sh.gather_time()
    print(f"step_loop {label:16s}: {us:.2f} us/step -> {B/us:.2f} M transitions/s, {2*B*R/us/1e3:.0f} GB/s; host {th*1e6/steps:.2f} us/step; gather_time {gt}")
print(sh.stats())
