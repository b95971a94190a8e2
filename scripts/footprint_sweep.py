"""Repair-kernel footprint sweep: the headline loop (config 2, prefetch group 8) with
per_update_kernel built lean (<= 64 registers) or full, 256 / 512 threads per CTA, cluster 8 / 16, plus the
gather-only loops that bound what the tree kernels can cost.  One JSON line per variant."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from reth_b200 import ReplayShard

ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
RB = 56464
cap = int(os.environ.get("CAP", 1 << 20)); B = 512; G = int(os.environ.get("GROUP", 8)); iters = int(os.environ.get("ITERS", 2048))
steps = int(os.environ.get("STEPS", 5))
VARIANTS = os.environ.get("VARIANTS", "0:512:16,1:512:16,1:512:8,-1:512:0").split(",")   # lean:threads:cluster; -1 / 0 = by launch size


def fill(sh):
    g = torch.Generator(device="cuda").manual_seed(0)
    for st in range(0, cap, 1 << 14):
        n = min(1 << 14, cap - st)
        s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
        z = torch.zeros(n, device="cuda")
        w = torch.randn(n, device="cuda", generator=g).abs() * (1 + 9 * (torch.rand(n, device="cuda", generator=g) < 0.05))
        sh.append([s0, z.long(), z, s0, z], w.float(), alpha=0.6)
    torch.cuda.synchronize()


def timed(sh, prio, group, **kw):
    sh.step_loop(iters, B, prio, seed=1, group=group, **kw)
    torch.cuda.synchronize()
    sh.gather_time()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        sh.step_loop(iters, B, prio, seed=1, group=group, **kw)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    gt, gn = sh.gather_time()
    return steps * iters * B / (ms / 1e3), (gt * 1e3 / gn if gn else None)


for v in VARIANTS:
    lean, thr, cl = v.split(":")
    os.environ["RETH_UPDATE_LEAN"], os.environ["RETH_UPDATE_THREADS"], os.environ["RETH_UPDATE_CLUSTER"] = lean, thr, cl
    sh = ReplayShard(cap, ATARI, max_batch=B, ring_depth=24, device=0)
    fill(sh)
    prio = torch.randn(iters * B, device="cuda").abs().float()
    val, _ = timed(sh, prio, G)
    _, gus = timed(sh, prio, G, time_gather=True, one_gather_stream=True)
    strict, _ = timed(sh, prio, 1)
    tree_only, _ = timed(sh, prio, G, gather=False)
    out = {"gather_units": int(os.environ.get("RETH_GATHER_UNITS", 1024)), "lean": int(lean), "threads": int(thr), "cluster": int(cl), "group": G, "value_Mps": val / 1e6,
           "hbm_TBs": 2 * RB * val / 1e12, "gather_us_one_stream": gus,
           "gather_frac": (2 * G * B * RB / (gus * 1e-6) / 1e9 / 6522.4) if gus else None,
           "strict_Mps": strict / 1e6, "tree_only_us_per_group": G * B / tree_only * 1e6}
    if v == VARIANTS[0]:
        # upper bounds: the same loop without the write-back / without any tree kernel in the steady state
        nu, _ = timed(sh, prio, G, update=False)
        ns, _ = timed(sh, prio, G, update=False, sample=False)
        out["no_update_Mps"], out["gather_only_Mps"] = nu / 1e6, ns / 1e6
        out["gather_only_TBs"] = 2 * RB * ns / 1e12
    print(json.dumps(out), flush=True)
    sh.close()
    del sh, prio
    torch.cuda.empty_cache()
