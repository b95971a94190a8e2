"""Device-source append rate: the append kernel vs the driver's D2D copies (RETH_APPEND_MEMCPY=1)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from reth_b200 import ReplayShard
sh = ReplayShard(int(os.environ.get("CAP", 1 << 18)), bench.ATARI, max_batch=512, ring_depth=2)
out = {"memcpy": os.environ.get("RETH_APPEND_MEMCPY", "0")}
for n in (4096, 16384, 65536):
    torch.cuda.synchronize()
    rows, gbs = bench.append_rate(torch, sh, n=n, reps=10)
    out[f"n{n}_gbs"] = round(gbs, 1)
print(json.dumps(out))
