"""Widening gather (uint8 ring -> float32 batch) vs the two alternatives a float32 learner has:
(a) raw uint8 gather + a torch .float() pass over the batch, (b) float32 rows stored in the ring
(what test/apex-dqn/worker.py:47-51 sends) + raw gather.  Bytes per sampled row: widening R + 4R,
(a) 2R + R + 4R, (b) 8R (frames only)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from reth_b200 import ReplayShard
B, G, cap = 512, 8, 1 << 17
U8 = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
F32 = [(np.float32, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.float32, (4, 84, 84)), (np.float32, ())]
FR = 28224
def fill(sh, f32):
    g = torch.Generator(device="cuda").manual_seed(0)
    for st in range(0, cap, 1 << 13):
        n = 1 << 13
        s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
        if f32: s0 = s0.float()
        z = torch.zeros(n, device="cuda")
        sh.append([s0, z.long(), z, s0, z], torch.rand(n, device="cuda", generator=g) + 0.01)
def timeit(fn, n=60):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n
out = {}
for name, struct, cast, post in (("widen", U8, ["f32", None, None, "f32", None], False),
                                 ("u8_gather_then_torch_float", U8, None, True),
                                 ("f32_rows_in_ring", F32, None, False)):
    sh = ReplayShard(cap, struct, max_batch=B, ring_depth=2 * G, out_cast=cast)
    fill(sh, struct is F32)
    k = [0]
    def step():
        first = (k[0] % 2) * G; k[0] += 1
        sh.sample_multi(B, first, G, seed=1)
        sh.gather_multi(first, G)
        if post:
            for s in range(first, first + G):
                d, _, _, _ = sh.slot_tensors(s)
                d[0].float(); d[3].float()
    us = timeit(step)
    out[name] = {"us_per_batch_of_512": us / G, "float32_frame_rows_per_s": G * B / us * 1e6}
    sh.close(); del sh
    torch.cuda.empty_cache()
print(json.dumps(out))
