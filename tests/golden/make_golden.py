#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference code.

Run in the build container only (needs /root/reference and numba):
    python tests/golden/make_golden.py
The reference has no seeded known-answer tests of its own, so
these vectors are the pin: its njit sum-tree (reth_buffer/reth_buffer/utils/
sumtree.py), PERSampler, FIFOPolicy, Schedule, pack and NumpyBuffer are imported
from /root/reference and driven with seeded inputs and *injected* uniforms
(_numba_find_index is called on explicit targets, as _numba_sample would).
The fixtures travel to the GPU box; /root/reference does not.
"""
import hashlib
import importlib.util
import os
import sys
import types

import numpy as np

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def load_reference():
    # non-arithmetic modules missing in this image are stubbed
    for name in ("lmdb", "netifaces", "portpicker", "lz4", "lz4.frame"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["lz4"].frame = sys.modules["lz4.frame"]
    sys.path.insert(0, os.path.join(REF, "reth_buffer"))
    import reth_buffer  # noqa: F401  (the reference package itself)
    from reth_buffer.utils import sumtree as st
    from reth_buffer.sampler.per_sampler import PERSampler
    from reth_buffer.cache_policy import FIFOPolicy
    from reth_buffer.utils.schedule import Schedule
    from reth_buffer.utils import pack
    nb = _load(os.path.join(REF, "reth/reth/buffer/buffer.py"), "ref_numpy_buffer")
    return st, PERSampler, FIFOPolicy, Schedule, pack, nb


def ref_sample(st, tree, batch, u):
    """sumtree.py:70-79 with u injected."""
    segment = tree._sum[0] / batch
    targets = (np.arange(batch) + u) * segment
    idx = np.array([st._numba_find_index(tree._tree, t) for t in targets], dtype=np.int64)
    return idx, tree._val[idx].copy()


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


def tape_case(st, cap, batch, steps, seed, fill_idx, prio_fn, store_arrays=True, min_mode_service=True):
    """Generic raw-priority tape: fill via update() in chunks of 64, then
    steps x (sample with injected u -> update sampled indices with new prios)."""
    rng = np.random.default_rng(seed)
    tree = st.NumbaSumTree(cap)
    fill_p = prio_fn(rng, len(fill_idx))
    for s in range(0, len(fill_idx), 64):
        tree.update(fill_idx[s:s + 64], fill_p[s:s + 64])
    out = {"cap": np.int64(cap), "batch": np.int64(batch), "fill_idx": fill_idx, "fill_p": fill_p}
    U, IDX, VAL, NEWP, ROOT = [], [], [], [], []
    for _ in range(steps):
        u = rng.random(batch)
        idx, val = ref_sample(st, tree, batch, u)
        ROOT.append([tree.sum(), tree.min(), tree._min[0]])
        newp = prio_fn(rng, batch)
        tree.update(idx, newp)
        U.append(u); IDX.append(idx); VAL.append(val); NEWP.append(newp)
    out.update(u=np.array(U), idx=np.array(IDX), val=np.array(VAL), newp=np.array(NEWP),
               root=np.array(ROOT), final_root=np.array([tree.sum(), tree.min(), tree._min[0]]))
    if store_arrays:
        out.update(sum=tree._sum.copy(), min=tree._min.copy(), valarr=tree._val.copy())
    else:
        out.update(sum_sha=sha(tree._sum), min_sha=sha(tree._min), val_sha=sha(tree._val),
                   sum_head=tree._sum[:2048].copy(), min_head=tree._min[:2048].copy())
    return out


def main():
    st, PERSampler, FIFOPolicy, Schedule, pack, nb = load_reference()

    # 1. cap 1000 (non power of two), full FIFO fill, duplicates arise naturally
    c = tape_case(st, 1000, 64, 30, 11, np.arange(1000, dtype=np.int64),
                  lambda r, n: np.abs(r.standard_normal(n)) + 1e-3)
    np.savez_compressed(os.path.join(OUT, "tree_cap1000_full.npz"), **c)

    # 2. cap 4096, 300 sparse arbitrary slots, all priorities > 1 (exposes the
    #    `_min` seeding quirk of sumtree.py:11 and untouched-node skipping :14,18)
    rng = np.random.default_rng(5)
    sparse = rng.choice(4096, 300, replace=False).astype(np.int64)
    c = tape_case(st, 4096, 32, 20, 12, sparse, lambda r, n: 1.5 + 8.5 * r.random(n))
    np.savez_compressed(os.path.join(OUT, "tree_cap4096_sparse.npz"), **c)

    # 3. tile/level boundary capacities (1..; 31/32/33 = one 5-level tile +-1;
    #    1023..1025 and 1056..1058 = two tile levels +-1)
    edge = {}
    for cap in (1, 2, 3, 4, 5, 7, 8, 31, 32, 33, 62, 63, 64, 65, 1023, 1024, 1025, 1056, 1057, 1058):
        b = min(16, cap)
        c = tape_case(st, cap, b, 3, 100 + cap, np.arange(cap, dtype=np.int64),
                      lambda r, n: 0.05 + r.random(n))
        for k, v in c.items():
            edge[f"c{cap}_{k}"] = v
    edge["caps"] = np.array([1, 2, 3, 4, 5, 7, 8, 31, 32, 33, 62, 63, 64, 65,
                             1023, 1024, 1025, 1056, 1057, 1058], np.int64)
    np.savez_compressed(os.path.join(OUT, "tree_edges.npz"), **edge)

    # 4. cap 100000 (non power of two), heavy-tailed priorities, B=512
    def heavy(r, n):
        return np.abs(r.standard_normal(n)) * (1 + 9 * (r.random(n) < 0.05)) + 1e-4
    c = tape_case(st, 100000, 512, 10, 13, np.arange(100000, dtype=np.int64), heavy, store_arrays=False)
    c.pop("fill_p")  # regenerated from the seed by the test (same rng stream)
    c["seed"] = np.int64(13)
    np.savez_compressed(os.path.join(OUT, "tree_cap100000_heavy.npz"), **c)

    # 5. config 1 through the reference PERSampler (float32 |TD| priorities,
    #    alpha 0.6, beta 0.4), cap 2^16, B=64, FIFO fill in chunks of 64
    cap, B, steps = 1 << 16, 64, 100
    rng = np.random.default_rng(0)
    urng = np.random.default_rng(1)
    sampler = PERSampler(cap, alpha=0.6, beta=0.4)
    policy = FIFOPolicy(cap)
    fill_w = np.abs(rng.standard_normal(cap)).astype(np.float32)
    for s in range(0, cap, 64):
        sampler.update(policy.get_indices(64), fill_w[s:s + 64])
    fill_p32 = sampler._normalize_weights(fill_w)
    assert fill_p32.dtype == np.float32
    assert np.array_equal(sampler.sumtree._val, fill_p32.astype(np.float64))
    U, IDX, W, NEWW, NEWP32, ROOT = [], [], [], [], [], []
    for _ in range(steps):
        u = urng.random(B)
        idx, val = ref_sample(st, sampler.sumtree, B, u)
        w = (val / sampler.sumtree.min()) ** (-sampler.beta.value())   # per_sampler.py:26
        ROOT.append([sampler.sumtree.sum(), sampler.sumtree.min()])
        neww = np.abs(rng.standard_normal(B)).astype(np.float32)
        sampler.on_step()
        sampler.update(idx, neww)
        U.append(u); IDX.append(idx); W.append(w); NEWW.append(neww)
        NEWP32.append(sampler._normalize_weights(neww))
    np.savez_compressed(
        os.path.join(OUT, "per_cfg1_cap65536.npz"), cap=np.int64(cap), batch=np.int64(B),
        alpha=0.6, beta=0.4, fill_w=fill_w, fill_p32=fill_p32, u=np.array(U), idx=np.array(IDX),
        w=np.array(W), neww=np.array(NEWW), newp32=np.array(NEWP32), root=np.array(ROOT),
        sum_sha=sha(sampler.sumtree._sum), min_sha=sha(sampler.sumtree._min),
        val_sha=sha(sampler.sumtree._val),
        final_root=np.array([sampler.sumtree.sum(), sampler.sumtree.min()]))

    # 6. schedules (utils/schedule.py) + PERSampler.on_step with apex config
    sched = {}
    for name, s in (("const", 0.6), ("lin", "0.4,1,2000"), ("exp", "exp,0.4,1,500")):
        sc = Schedule.from_str(s)
        vals = [sc.value()]
        for _ in range(2500):
            sc.step()
            vals.append(sc.value())
        sched[name] = np.array(vals)
    sampler = PERSampler(2048, alpha=0.5, beta="0.4,1,2000")
    rng = np.random.default_rng(21)
    sampler.update(np.arange(2048), rng.random(2048).astype(np.float32) + 0.01)
    U, IDX, W = [], [], []
    for _ in range(40):
        u = rng.random(32)
        idx, val = ref_sample(st, sampler.sumtree, 32, u)
        W.append((val / sampler.sumtree.min()) ** (-sampler.beta.value()))
        U.append(u); IDX.append(idx)
        for _ in range(60):
            sampler.on_step()
    sched.update(per_fill=sampler.sumtree._val.copy(), per_u=np.array(U), per_idx=np.array(IDX),
                 per_w=np.array(W))
    np.savez_compressed(os.path.join(OUT, "schedule_per.npz"), **sched)

    # 7. FIFO slots + NumpyBuffer ring append/select with wrap
    pol = FIFOPolicy(10)
    fifo = [pol.get_indices(n) for n in (4, 7, 3, 10, 1)]
    buf = nb.NumpyBuffer(10)
    rng = np.random.default_rng(3)
    ring = {}
    for i, n in enumerate((4, 7, 3, 10, 1)):
        cols = [rng.integers(0, 256, (n, 2, 3), dtype=np.uint8), rng.integers(0, 18, n),
                rng.standard_normal(n).astype(np.float32)]
        idx = buf.append_batch(cols)
        ring[f"in{i}_0"], ring[f"in{i}_1"], ring[f"in{i}_2"] = cols
        ring[f"idx{i}"] = np.asarray(idx, np.int64)
        ring[f"fifo{i}"] = fifo[i].astype(np.int64)
    sel = np.array([9, 0, 3, 3, 7], np.int64)
    s0, s1, s2 = buf.select(sel)
    ring.update(sel=sel, sel_0=s0, sel_1=s1, sel_2=s2)
    np.savez_compressed(os.path.join(OUT, "fifo_ring.npz"), **ring)

    # 8. wire format (utils/pack.py) with a fixed clock: per-row + outer message
    #    exactly as Client.append builds them (client/client.py:26-35)
    import time as _time
    real = _time.time
    pack.time.time = lambda: 1234.5
    try:
        rng = np.random.default_rng(4)
        cols = [rng.integers(0, 256, (3, 4, 6, 6), dtype=np.uint8), np.arange(3, dtype=np.int64),
                rng.standard_normal(3).astype(np.float32)]
        weights = rng.random(3).astype(np.float32)
        rows = [pack.serialize([col[i, ...] for col in cols], compress=False) for i in range(3)]
        msg = pack.serialize([rows, weights], compress=False)
        upd = pack.serialize([np.array([5, 1, 5], np.int64), weights, True])
    finally:
        pack.time.time = real
    np.savez_compressed(os.path.join(OUT, "pack_wire.npz"), c0=cols[0], c1=cols[1], c2=cols[2],
                        weights=weights, msg=np.frombuffer(bytes(msg), np.uint8),
                        upd=np.frombuffer(bytes(upd), np.uint8),
                        row0=np.frombuffer(bytes(rows[0]), np.uint8))
    # 9. the other sampler plugins (sampler/uniform_sampler.py, sampler/fifo_sampler.py), seeded global RNG
    from reth_buffer.sampler.uniform_sampler import UniformSampler
    from reth_buffer.sampler.fifo_sampler import FIFOSampler
    rng = np.random.default_rng(6)
    out = {}
    us = UniformSampler(300)
    fs = FIFOSampler(50)
    msgs = [rng.integers(0, 1000, n).astype(np.int64) for n in (64, 64, 100, 40, 32)]   # 300 in total: fills the log exactly
    wts = [rng.random(len(m)) for m in msgs]
    for k, (m, w) in enumerate(zip(msgs, wts)):
        us.update(m, w)
        fs.update(m, w)
        out[f"msg{k}"], out[f"w{k}"] = m, w
        np.random.seed(100 + k)
        draws = np.random.choice(us.tail, 16)                     # what sample() will draw next
        np.random.seed(100 + k)
        ui, uw = us.sample(16)
        out[f"u_draws{k}"], out[f"u_idx{k}"], out[f"u_w{k}"] = draws, np.asarray(ui), np.asarray(uw)
        out[f"u_tail{k}"] = np.int64(us.tail)
        if fs.ready_sample(8):
            fi, fw = fs.sample(8)
            out[f"f_idx{k}"], out[f"f_w{k}"] = np.asarray(fi, np.int64), np.asarray(fw)
        out[f"f_len{k}"] = np.int64(len(fs.buffer))
    us.update(np.arange(5), np.ones(5))                           # a full log ignores further messages
    out["u_tail_final"] = np.int64(us.tail)
    np.savez_compressed(os.path.join(OUT, "sampler_plugins.npz"), **out)
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
