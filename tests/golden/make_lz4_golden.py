"""Golden LZ4 frames written by liblz4 itself (through pyarrow's bundled copy, codec "lz4" = the
LZ4 frame format), for tests/test_lz4f_cpu.py.  Run in the build container:
    python tests/golden/make_lz4_golden.py
The python `lz4` package the reference uses (utils/pack.py:69,155) is not installed here; pyarrow's
liblz4 writes the same frame format (linked 64 KB blocks, no content size)."""
import os

import numpy as np
import pyarrow as pa

rng = np.random.default_rng(7)
codec = pa.Codec("lz4")
plain = {
    "empty": np.zeros(0, np.uint8),
    "short": np.frombuffer(b"abcabcabcabcabcabcabcabc-the end", np.uint8),
    "runs": np.repeat(rng.integers(0, 6, 40000, dtype=np.uint8), rng.integers(1, 9, 40000)),     # > 64 KB: linked blocks
    "frame": np.repeat(np.repeat(rng.integers(0, 256, (21, 21), dtype=np.uint8), 4, 0), 4, 1).reshape(-1),
    "noise": rng.integers(0, 256, 70000, dtype=np.uint8),                                        # stored blocks
    "period5": np.tile(np.arange(5, dtype=np.uint8), 20001),
}
out = {}
for k, a in plain.items():
    out["plain_" + k] = a
    out["frame_" + k] = np.frombuffer(codec.compress(a.tobytes(), asbytes=True), np.uint8)
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lz4_frames.npz")
np.savez_compressed(path, **out)
print(path, os.path.getsize(path), {k: (plain[k].size, out["frame_" + k].size) for k in plain})
