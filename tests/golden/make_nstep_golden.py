#!/usr/bin/env python
"""Golden tape for the n-step assembly (device kernel nstep_emit_kernel, oracle nstep_fold) from the UNMODIFIED reference class
(reth/reth/utils/nstep_adder.py).  Build container only:  python tests/golden/make_nstep_golden.py"""
import importlib.util
import os

import numpy as np

spec = importlib.util.spec_from_file_location("ref_nstep", "/root/reference/reth/reth/utils/nstep_adder.py")
mod = importlib.util.module_from_spec(spec)
spec.loader.exec_module(mod)

rng = np.random.default_rng(12)
out = {}
for name, (gamma, step, n) in {"g99_s3": (0.99, 3, 200), "g9_s5": (0.9, 5, 300), "g1_s1": (1.0, 1, 50)}.items():
    adder = mod.NStepAdder(gamma, step)
    s0 = rng.integers(0, 1000, n)
    s1 = rng.integers(0, 1000, n)
    a = rng.integers(0, 6, n)
    r = rng.standard_normal(n)
    done = rng.random(n) < 0.08
    emitted = []                       # (push index, s0, a, r, s1, done) or push index alone when None
    for i in range(n):
        res = adder.push(int(s0[i]), int(a[i]), float(r[i]), int(s1[i]), bool(done[i]))
        emitted.append([i, -1, -1, 0.0, -1, -1] if res is None else [i, res[0], res[1], res[2], res[3], int(res[4])])
    out[name + "_in"] = np.stack([s0, a, s1, done.astype(np.int64)], 1)
    out[name + "_r"] = r
    out[name + "_out"] = np.array(emitted, dtype=np.float64)
    out[name + "_cfg"] = np.array([gamma, step, n], dtype=np.float64)
# the way test/apex-dqn/worker.py:47-53 feeds it: 0-d numpy arrays, reward / done as float32.  `item[2] += t_gamma * r`
# then runs in numpy (in place on the 0-d float32 array); recorded under this container's numpy (see _numpy below).
for name, (gamma, step, n) in {"f32_g99_s3": (0.99, 3, 300), "f32_g97_s4": (0.97, 4, 200)}.items():
    adder = mod.NStepAdder(gamma, step)
    s0 = rng.integers(0, 1000, n)
    s1 = rng.integers(0, 1000, n)
    a = rng.integers(0, 6, n)
    r = rng.standard_normal(n).astype(np.float32)
    done = (rng.random(n) < 0.08).astype(np.float32)
    emitted = []
    for i in range(n):
        res = adder.push(np.asarray(s0[i], dtype="i8"), np.asarray(a[i], dtype="i8"), np.asarray(r[i], dtype="f4"),
                         np.asarray(s1[i], dtype="i8"), np.asarray(done[i], dtype="f4"))
        if res is None:
            emitted.append([i, -1, -1, 0.0, -1, -1])
        else:
            assert res[2].dtype == np.float32
            emitted.append([i, int(res[0]), int(res[1]), float(res[2]), int(res[3]), int(res[4])])
    out[name + "_in"] = np.stack([s0, a, s1, done.astype(np.int64)], 1)
    out[name + "_r"] = r
    out[name + "_out"] = np.array(emitted, dtype=np.float64)
    out[name + "_cfg"] = np.array([gamma, step, n], dtype=np.float64)
out["_numpy"] = np.array(np.__version__)
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "nstep_adder.npz")
np.savez_compressed(path, **out)
print(path, os.path.getsize(path))
