import sys; sys.path.insert(0, "/root/repo")
import numpy as np
from reth_b200.buffer import PrioritizedBuffer
rng = np.random.default_rng(5)
for mode in ("device", "host"):
    vals = []
    for trial in range(15):
        cap = 100
        buf = PrioritizedBuffer(cap)
        data = [np.arange(cap, dtype="i8"), rng.random((cap, 8)).astype("f4")]
        w = rng.random(cap)
        buf.append_batch(data, weights=w)
        cnt = np.zeros(cap)
        for _ in range(100):
            _, idx, _ = buf.sample(64, uniforms=None if mode == "device" else rng.random(64))
            np.add.at(cnt, idx, 1)
        p = (w + 1e-6) ** 0.6
        vals.append(np.sum(np.abs(cnt / cnt.sum() - p / p.sum()) / (p / p.sum())) / cap)
    print(mode, np.mean(vals), np.min(vals), np.max(vals))
