"""BASELINE config 5: priority-update sweep, batch 64..8192 x capacity 2^16..2^24, tree-only.
GPU: device-resident loops (reth_replay_step_loop, serial, no gather): sample-only, update-only and
sample+update, microseconds per call.  CPU: the oracle port (C restatement of the reference's njit
tree) on one host core, same shapes.  Writes gpurun_out/sweep_config5.json + .md."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from reth_b200 import ReplayShard
from oracle import per_oracle as po

po.build()
caps = [1 << 16, 1 << 18, 1 << 20, 1 << 22, 1 << 24]
batches = [64, 256, 512, 2048, 8192]
if os.environ.get("QUICK"):
    caps, batches = caps[:3], batches[:3]
rows = []
for cap in caps:
    rng = np.random.default_rng(0)
    w = (np.abs(rng.standard_normal(cap)) * (1 + 9 * (rng.random(cap) < 0.05))).astype(np.float32)
    tree = po.OracleSumTree(cap)
    tree.update(np.arange(cap), po.normalize_weights(w, 0.6))
    wd = torch.from_numpy(w).cuda()
    for B in batches:
        sh = ReplayShard(cap, [], max_batch=B, store_rows=False, ring_depth=2)
        for st in range(0, cap, 16384):
            sh.append([], wd[st:st + 16384], alpha=0.6)
        steps = 200
        prio = torch.rand(steps * B, device="cuda")
        res = {}
        for label, kw in (("sample", {"update": False}), ("update", {"sample": False}), ("both", {})):
            sh.step_loop(20, B, prio, gather=False, serial=True, **kw)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            sh.step_loop(steps, B, prio, gather=False, serial=True, **kw)
            e1.record()
            torch.cuda.synchronize()
            res[label] = e0.elapsed_time(e1) * 1e3 / steps
        sh.close()
        # CPU oracle, one core: sample + update per step
        csteps = max(5, min(200, 200000 // B))
        u = np.random.default_rng(1).random(csteps * B)
        newp = np.abs(np.random.default_rng(2).standard_normal(csteps * B)).astype(np.float32)
        po.step_loop(tree, None, 0, 2, B, u, newp, 0.6, 0.4)
        t0 = time.perf_counter()
        po.step_loop(tree, None, 0, csteps, B, u, newp, 0.6, 0.4)
        cpu_us = (time.perf_counter() - t0) * 1e6 / csteps
        row = {"capacity": cap, "batch": B, "gpu_sample_us": res["sample"], "gpu_update_us": res["update"],
               "gpu_both_us": res["both"], "cpu_both_us_1core": cpu_us,
               "gpu_Midx_per_s": B / res["both"], "cpu_Midx_per_s_1core": B / cpu_us}
        rows.append(row)
        print(row, flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/sweep_config5.json", "w"), indent=1)
with open("gpurun_out/sweep_config5.md", "w") as f:
    f.write("| capacity | batch | GPU sample us | GPU update us | GPU sample+update us | CPU (1 core) sample+update us | GPU M idx/s | CPU M idx/s |\n|---|---|---|---|---|---|---|---|\n")
    for r in rows:
        f.write(f"| 2^{int(np.log2(r['capacity']))} | {r['batch']} | {r['gpu_sample_us']:.1f} | {r['gpu_update_us']:.1f} | "
                f"{r['gpu_both_us']:.1f} | {r['cpu_both_us_1core']:.1f} | {r['gpu_Midx_per_s']:.1f} | {r['cpu_Midx_per_s_1core']:.1f} |\n")
