"""Gather of a Mode G share: a slot sampled with sample_global(GB, n_shards = 8) holds about GB/8 rows of this
shard; its gather launch is sized for GB rows.  Times that gather against a plain gather of the same number of
rows (the ideal).  One GPU: the seven peer roots are copies of the shard's own root."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from reth_b200 import ReplayShard

ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
cap = int(os.environ.get("CAP", 1 << 18))
out = []
for GB in (4096, 16384, 65536):
    sh = ReplayShard(cap, ATARI, max_batch=GB, ring_depth=2)
    g = torch.Generator(device="cuda").manual_seed(0)
    for st in range(0, cap, 1 << 14):
        n = min(1 << 14, cap - st)
        s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
        z = torch.zeros(n, device="cuda")
        sh.append([s0, z.long(), z, s0, z], torch.rand(n, device="cuda", generator=g) + 0.01)
    roots = torch.as_tensor(sh.root_dev(), device="cuda").repeat(8).clone()
    sh.sample_global(GB, roots, 8, 3, beta=0.4, seed=1, slot=0)
    first, count = sh.slot_meta(0)
    sh.sample(count, seed=2, slot=1)

    def timed(slot, reps=50):
        for _ in range(5):
            sh.gather(slot)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            sh.gather(slot)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / reps
    row = {"global_batch": GB, "share_rows": int(count), "share_gather_us": timed(0), "plain_gather_us": timed(1)}
    row["share_GBps"] = count * 2 * 56464 / row["share_gather_us"] / 1e3
    row["plain_GBps"] = count * 2 * 56464 / row["plain_gather_us"] / 1e3
    out.append(row)
    print(json.dumps(row), flush=True)
    sh.close()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/sparse_gather.json", "w"), indent=1)
