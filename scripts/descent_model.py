"""Sampling against a latency-bound descent model (BASELINE.json north_star: "sampling throughput against a
latency-bound tree-descent model").

per_sample_kernel walks the tiled tree with one warp per stratified target: ceil(depth/5) DEPENDENT 512 B tile
loads, 5 shuffle levels each.  Model, with L = tile levels of the capacity and W = warps of the launch:

  latency regime  (W <= resident warps = 148 SMs x 64):   t(launch) = t0 + L * t_tile
  throughput regime (W >> resident warps):                samples/s = resident_warps / (t_w + L * t_tile')

This script measures sample-only loops (reth_replay_step_loop, serial, no gather, no update; CUDA events around
`steps` back-to-back launches) for capacities with 1..5 tile levels and batches from 512 to 2^19, fits
(t0, t_tile) by least squares on the latency regime and prints model vs measured.
Writes gpurun_out/descent_model.json + .md."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from reth_b200 import ReplayShard

CAPS = [1 << 4, 1 << 9, 1 << 14, 1 << 19, 1 << 20, 1 << 22, 1 << 24]
BATCHES = [512, 2048, 8192, 65536, 524288]
RESIDENT = 148 * 64


def tile_levels(cap):
    d = int(np.floor(np.log2(cap)))
    return d // 5 + 1


rows = []
for cap in CAPS:
    g = torch.Generator(device="cuda").manual_seed(0)
    w = (torch.randn(cap, device="cuda", generator=g).abs() *
         (1 + 9 * (torch.rand(cap, device="cuda", generator=g) < 0.05))).float() + 1e-3
    for B in BATCHES:
        sh = ReplayShard(cap, [], max_batch=B, store_rows=False, ring_depth=2)
        chunk = 16384
        for st in range(0, cap, chunk):
            sh.append([], w[st:st + chunk], alpha=0.6)
        steps = 200 if B <= 8192 else 40
        prio = torch.rand(steps * B, device="cuda")
        sh.step_loop(10, B, prio, gather=False, serial=True, update=False)
        torch.cuda.synchronize()
        best = None
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            sh.step_loop(steps, B, prio, gather=False, serial=True, update=False)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / steps
            best = us if best is None else min(best, us)
        sh.close()
        del prio
        row = {"capacity": cap, "tile_levels": tile_levels(cap), "batch": B, "us_per_launch": best,
               "Msamples_per_s": B / best, "ns_per_sample": best * 1e3 / B,
               "waves": B / RESIDENT}
        rows.append(row)
        print(row, flush=True)

# latency regime fit: t = t0 + L * t_tile over B = 512 and capacities that stay in L2 (<= 2^20)
lat = [r for r in rows if r["batch"] == 512 and r["capacity"] <= (1 << 20)]
A = np.array([[1.0, r["tile_levels"]] for r in lat])
y = np.array([r["us_per_launch"] for r in lat])
(t0, t_tile), *_ = np.linalg.lstsq(A, y, rcond=None)
# throughput regime fit: per-warp time = t_w + L * t_tile2 from the largest batch
thr = [r for r in rows if r["batch"] == BATCHES[-1] and r["capacity"] <= (1 << 20)]
A2 = np.array([[1.0, r["tile_levels"]] for r in thr])
y2 = np.array([RESIDENT * r["ns_per_sample"] * 1e-3 for r in thr])     # us of residency per warp
(tw, t_tile2), *_ = np.linalg.lstsq(A2, y2, rcond=None)
fit = {"latency_regime": {"t0_us": float(t0), "t_tile_us": float(t_tile)},
       "throughput_regime": {"t_warp_us": float(tw), "t_tile_us": float(t_tile2), "resident_warps": RESIDENT}}
for r in rows:
    L = r["tile_levels"]
    r["model_latency_us"] = float(t0 + L * t_tile)
    r["model_throughput_us"] = float(r["batch"] / RESIDENT * (tw + L * t_tile2))
    r["model_us"] = max(r["model_latency_us"], r["model_throughput_us"])
print(fit)
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"fit": fit, "rows": rows}, open("gpurun_out/descent_model.json", "w"), indent=1)
with open("gpurun_out/descent_model.md", "w") as f:
    f.write(f"fit: latency regime t = {t0:.2f} us + L x {t_tile:.3f} us (B = 512, capacity <= 2^20); throughput regime "
            f"residency per warp = {tw:.2f} us + L x {t_tile2:.3f} us over {RESIDENT} resident warps (B = {BATCHES[-1]})\n\n")
    f.write("| capacity | tile levels L | batch | waves | measured us/launch | model us | M samples/s | ns/sample |\n|---|---|---|---|---|---|---|---|\n")
    for r in rows:
        f.write(f"| 2^{int(np.log2(r['capacity']))} | {r['tile_levels']} | {r['batch']} | {r['waves']:.2f} | {r['us_per_launch']:.2f} | "
                f"{r['model_us']:.2f} | {r['Msamples_per_s']:.1f} | {r['ns_per_sample']:.2f} |\n")
