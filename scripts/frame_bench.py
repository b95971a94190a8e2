"""Frame-de-duplicated layout vs full-stack rows: HBM per transition and gather time for the same batch.
One long synthetic episode: transition i has s0 = frames i..i+3, s1 = frames i+1..i+4."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from reth_b200 import ReplayShard

K, FRAME, B, G = 4, (84, 84), 512, 8
cap = int(os.environ.get("CAP", 1 << 18))
STRUCT = [(np.uint8, (K,) + FRAME), (np.int64, ()), (np.float32, ()), (np.uint8, (K,) + FRAME), (np.float32, ())]
g = torch.Generator(device="cuda").manual_seed(0)
out = {"capacity": cap, "batch_rows_per_launch": B * G}


def timed_gather(sh, reps=30):
    u = torch.rand(G * B, device="cuda", dtype=torch.float64, generator=g)
    for _ in range(3):
        sh.sample_multi(B, 0, G, uniforms_dev=u)
        sh.gather_multi(0, G)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0.0
    for _ in range(reps):
        sh.sample_multi(B, 0, G, uniforms_dev=torch.rand(G * B, device="cuda", dtype=torch.float64, generator=g))
        e0.record()
        sh.gather_multi(0, G)
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / reps * 1e3


for cast in (None, "f32/255"):
    oc = [cast, None, None, cast, None] if cast else None
    tag = "u8" if cast is None else "f32"
    # full-stack rows (what the reference stores)
    sh = ReplayShard(cap, STRUCT, max_batch=B, ring_depth=G, out_cast=oc)
    for st in range(0, cap, 8192):
        n = min(8192, cap - st)
        s0 = torch.randint(0, 256, (n, K) + FRAME, dtype=torch.uint8, device="cuda", generator=g)
        z = torch.zeros(n, device="cuda")
        sh.append([s0, z.long(), z, s0, z], torch.rand(n, device="cuda", generator=g) + 0.01)
    out[f"full_{tag}_hbm_bytes_per_transition"] = sh.stats()["hbm_bytes"] / cap
    out[f"full_{tag}_gather_us"] = timed_gather(sh)
    sh.close()
    del sh
    torch.cuda.empty_cache()
    # frame ring + references
    fcap = cap + 16
    sh = ReplayShard(cap, STRUCT, max_batch=B, ring_depth=G, out_cast=oc,
                     frames={"shape": FRAME, "capacity": fcap, "columns": [0, 3]})
    for st in range(0, fcap, 16384):
        n = min(16384, fcap - st)
        sh.append_frames(torch.randint(0, 256, (n,) + FRAME, dtype=torch.uint8, device="cuda", generator=g))
    base = torch.arange(cap, device="cuda", dtype=torch.int32)[:, None]
    r0 = (base + torch.arange(K, device="cuda", dtype=torch.int32)[None, :]).contiguous()
    r1 = (r0 + 1).contiguous()
    z = torch.zeros(cap, device="cuda")
    for st in range(0, cap, 16384):
        n = min(16384, cap - st)
        sh.append([r0[st:st + n].contiguous(), z[st:st + n].long(), z[st:st + n], r1[st:st + n].contiguous(), z[st:st + n]],
                  torch.rand(n, device="cuda", generator=g) + 0.01)
    out[f"frames_{tag}_hbm_bytes_per_transition"] = sh.stats()["hbm_bytes"] / cap
    out[f"frames_{tag}_gather_us"] = timed_gather(sh)
    sh.close()
    del sh
    torch.cuda.empty_cache()
rowb = 2 * K * 7056
out["gather_GBps_full_u8"] = 2 * B * G * rowb / out["full_u8_gather_us"] / 1e3
out["gather_GBps_frames_u8"] = 2 * B * G * rowb / out["frames_u8_gather_us"] / 1e3
print(json.dumps(out))
