"""GPU parity: the CUDA shard (through the C ABI) against the oracle and the
golden vectors of the reference.  Bit-exact for indices, leaf values, sums and
mins (float64, same operation order); IS weights within 1e-5 relative
(north_star tolerance; CUDA pow vs libm pow); transform within 1e-5 relative."""
import hashlib

import numpy as np
import pytest

from oracle import per_oracle as po

pytestmark = pytest.mark.gpu

RTOL = 1e-5


def _sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


@pytest.fixture(scope="module")
def Shard():
    import __graft_entry__ as entry
    entry.build()
    from reth_b200 import ReplayShard
    return ReplayShard


def _replay_tape(Shard, g, prefix="", fill_p=None, cluster=0):
    cap, batch = int(g[prefix + "cap"]), int(g[prefix + "batch"])
    sh = Shard(cap, [], max_batch=max(batch, 64), store_rows=False, update_cluster=cluster)
    fill_idx = g[prefix + "fill_idx"]
    fill_p = g[prefix + "fill_p"] if fill_p is None else fill_p
    for s in range(0, len(fill_idx), 64):
        sh.update(fill_idx[s:s + 64], fill_p[s:s + 64], transform=False)
    for t in range(len(g[prefix + "u"])):
        sh.sample(batch, beta=0.4, uniforms=g[prefix + "u"][t], slot=t % 2)
        _, idx, w = sh.slot_read(t % 2)
        assert np.array_equal(idx, g[prefix + "idx"][t]), f"{prefix} step {t}"
        total, mn = sh.root()
        root = g[prefix + "root"][t]
        assert total == root[0] and mn == root[1]
        wref = (g[prefix + "val"][t] / root[1]) ** (-0.4)
        assert np.allclose(w, wref, rtol=RTOL, atol=0)
        sh.update(idx, g[prefix + "newp"][t], transform=False)
    return sh


@pytest.mark.parametrize("name", ["tree_cap1000_full", "tree_cap4096_sparse"])
@pytest.mark.parametrize("cluster", [1, 8])
def test_golden_tape_full_arrays(Shard, golden, name, cluster):
    g = golden(name)
    sh = _replay_tape(Shard, g, cluster=cluster)
    s, m, v = sh.tree_arrays()
    assert np.array_equal(v, g["valarr"])
    assert np.array_equal(s, g["sum"])
    assert np.array_equal(m, g["min"])     # includes the untouched-node / seed-1 quirk
    sh.close()


def test_golden_edge_capacities(Shard, golden):
    g = golden("tree_edges")
    for cap in g["caps"]:
        pre = f"c{cap}_"
        sh = _replay_tape(Shard, g, pre)
        s, m, v = sh.tree_arrays()
        assert np.array_equal(s, g[pre + "sum"]), cap
        assert np.array_equal(m, g[pre + "min"]), cap
        sh.close()


def test_golden_heavy_tail_cap100000(Shard, golden):
    g = golden("tree_cap100000_heavy")
    rng = np.random.default_rng(int(g["seed"]))
    n = 100000
    fill_p = np.abs(rng.standard_normal(n)) * (1 + 9 * (rng.random(n) < 0.05)) + 1e-4
    sh = _replay_tape(Shard, g, fill_p=fill_p)
    s, m, v = sh.tree_arrays()
    assert np.array_equal(_sha(s), g["sum_sha"])
    assert np.array_equal(_sha(m), g["min_sha"])
    assert np.array_equal(_sha(v), g["val_sha"])
    sh.close()


def test_golden_cfg1_per_sampler(Shard, golden):
    """BASELINE config 1 shape through the PERSampler path (float32 |TD| in)."""
    g = golden("per_cfg1_cap65536")
    cap, B = int(g["cap"]), int(g["batch"])
    struct = [(np.float32, (4,)), (np.int64, ()), (np.float32, ()), (np.float32, (4,)), (np.float32, ())]
    sh = Shard(cap, struct, max_batch=B)
    rng = np.random.default_rng(0)
    cols = [rng.standard_normal((cap, 4)).astype(np.float32), rng.integers(0, 2, cap),
            rng.standard_normal(cap).astype(np.float32), rng.standard_normal((cap, 4)).astype(np.float32),
            np.zeros(cap, np.float32)]
    for s in range(0, cap, 64):
        first = sh.append([c[s:s + 64] for c in cols], g["fill_w"][s:s + 64], alpha=0.6)
        assert first == s                     # slot == arrival order (test_rb.py:27-29)
    _, _, v = sh.tree_arrays()
    p32 = g["fill_p32"].astype(np.float64)
    assert np.allclose(v, p32, rtol=RTOL, atol=0)          # transform gate
    # NumPy's SIMD float32 pow is itself <=1 ulp off the correctly rounded value in ~21% of
    # cases (measured in tests/golden/make_golden.py's container), so only the gate is asserted;
    # the device computes the correctly rounded float32 power: error vs NumPy <= 1 float32 ulp
    assert np.max(np.abs(v - p32) / p32) < 2.4e-7
    # index parity is defined for identical leaves: inject the reference's
    sh.update(np.arange(cap), p32, transform=False)
    for t in range(len(g["u"])):
        sh.sample(B, beta=0.4, uniforms=g["u"][t])
        sh.gather(0)
        data, idx, w = sh.slot_read(0, with_rows=True)
        assert np.array_equal(idx, g["idx"][t])
        assert np.allclose(w, g["w"][t], rtol=RTOL, atol=0)
        for c in range(5):
            assert np.array_equal(data[c], cols[c][idx].astype(data[c].dtype))
        total, mn = sh.root()
        assert total == g["root"][t][0] and mn == g["root"][t][1]
        sh.update(idx, g["newp32"][t].astype(np.float64), transform=False)
    s, m, v = sh.tree_arrays()
    assert np.array_equal(_sha(s), g["sum_sha"])
    assert np.array_equal(_sha(m), g["min_sha"])
    assert np.array_equal(_sha(v), g["val_sha"])
    sh.close()


def test_schedule_golden_weights(Shard, golden):
    """beta schedule 0.4 -> 1 (test/apex-dqn/config.yaml:27) feeding the fused IS weights."""
    from reth_b200.sampler import Schedule
    g = golden("schedule_per")
    sh = Shard(2048, [], max_batch=32, store_rows=False)
    sh.update(np.arange(2048), g["per_fill"], transform=False)
    beta = Schedule.from_str("0.4,1,2000")
    for t in range(len(g["per_u"])):
        sh.sample(32, beta=beta.value(), uniforms=g["per_u"][t])
        _, idx, w = sh.slot_read(0)
        assert np.array_equal(idx, g["per_idx"][t])
        assert np.allclose(w, g["per_w"][t], rtol=RTOL, atol=0)
        for _ in range(60):
            beta.step()
    sh.close()


@pytest.mark.parametrize("cap,B", [(1 << 16, 64), (1 << 20, 512), ((1 << 20) + 12345, 8192), (1 << 22, 2048)])
def test_random_tape_vs_oracle(Shard, cap, B):
    """Seeded tapes at BASELINE sizes: duplicates inside batches, sparse then full trees."""
    rng = np.random.default_rng(cap % 1000 + B)
    sh = Shard(cap, [], max_batch=B, store_rows=False)
    ref = po.OracleSumTree(cap)
    heavy = lambda n: np.abs(rng.standard_normal(n)) * (1 + 9 * (rng.random(n) < 0.05)) + 1e-3  # noqa: E731
    # sparse arbitrary slots first (untouched-node min semantics), with explicit duplicates
    idx = rng.integers(0, cap, 5000)
    idx[100:200] = idx[0:100]
    p = heavy(5000) + 1.0
    sh.update(idx, p, transform=False)
    ref.update(idx, p)
    s, m, v = sh.tree_arrays()
    assert np.array_equal(v, ref._val) and np.array_equal(s, ref._sum) and np.array_equal(m, ref._min)
    # FIFO fill of everything (tree-only append: iota slots)
    for st in range(0, cap, 1 << 18):
        n = min(1 << 18, cap - st)
        pw = heavy(n)
        first = sh.append([], pw, transform=False)
        assert first == st
        ref.update(np.arange(st, st + n), pw)
    for t in range(12):
        u = rng.random(B)
        sh.sample(B, beta=0.7, uniforms=u, slot=t % 2)
        _, gi, gw = sh.slot_read(t % 2)
        ri, rv = ref.sample(B, u)
        assert np.array_equal(gi, ri), t
        assert np.allclose(gw, po.is_weights(rv, ref.min(0), 0.7), rtol=RTOL, atol=0)
        newp = heavy(B)
        if t % 3 == 0:
            sh.update_slot(t % 2, newp, transform=False)
        else:
            sh.update(gi if t % 3 == 1 else gi.astype(np.int32), newp, transform=False)
        ref.update(ri, newp)
        assert sh.root() == (ref.sum(), ref.min(0))
    s, m, v = sh.tree_arrays()
    assert np.array_equal(v, ref._val) and np.array_equal(s, ref._sum) and np.array_equal(m, ref._min)
    sh.close()


def test_transform_dtypes(Shard):
    rng = np.random.default_rng(9)
    sh = Shard(4096, [], max_batch=64, store_rows=False)
    w32 = np.abs(rng.standard_normal(4096)).astype(np.float32)
    sh.update(np.arange(4096), w32, alpha=0.5)
    _, _, v = sh.tree_arrays()
    assert np.allclose(v, po.normalize_weights(w32, 0.5), rtol=RTOL, atol=0)
    w64 = np.abs(rng.standard_normal(4096))
    sh.update(np.arange(4096), w64, alpha=0.6)
    _, _, v = sh.tree_arrays()
    assert np.allclose(v, po.normalize_weights(w64, 0.6), rtol=1e-12, atol=0)
    sh.close()


def test_errors(Shard):
    from reth_b200 import ReplayError
    sh = Shard(100, [(np.float32, (3,))], max_batch=16)
    with pytest.raises(AssertionError):
        sh.append([np.zeros((4, 3), np.float32)], np.ones(5, np.float32))
    with pytest.raises(ReplayError):
        sh.append([np.zeros((101, 3), np.float32)], np.ones(101, np.float32))   # > capacity
    with pytest.raises(ReplayError):
        sh.sample(17)
    sh.update(np.array([5, 100]), np.ones(2), transform=False)                  # 100 is out of range
    with pytest.raises(ReplayError, match="outside"):
        sh.tree_arrays()
    sh.close()


@pytest.mark.parametrize("G", [2, 4])
def test_prefetch_group_vs_oracle(Shard, G):
    """sample_multi / update_slots: G independent stratified draws from one tree state, then one
    sequential-semantics write-back of the concatenation (duplicates across draws included)."""
    import torch
    cap, B = 200_000, 256
    rng = np.random.default_rng(40 + G)
    sh = Shard(cap, [], max_batch=B, ring_depth=2 * G, store_rows=False)
    ref = po.OracleSumTree(cap)
    p = np.abs(rng.standard_normal(cap)) + 1e-3
    p[rng.choice(cap, 40, replace=False)] = 2500.0          # leaves wider than a stratum: duplicates
    sh.append([], p, transform=False)
    ref.update(np.arange(cap), p)
    for t in range(6):
        first = (t % 2) * G
        u = rng.random(G * B)
        sh.sample_multi(B, first, G, beta=0.5, uniforms_dev=torch.from_numpy(u).cuda())
        ridx = []
        for g in range(G):
            _, gi, gw = sh.slot_read(first + g)
            ri, rv = ref.sample(B, u[g * B:(g + 1) * B])
            assert np.array_equal(gi, ri), (t, g)
            assert np.allclose(gw, po.is_weights(rv, ref.min(0), 0.5), rtol=RTOL, atol=0)
            ridx.append(ri)
        ridx = np.concatenate(ridx)
        if t == 0:
            assert len(np.unique(ridx)) < len(ridx) - B // 4   # duplicates within and across draws
        newp = np.abs(rng.standard_normal(G * B)) + 1e-3
        newp[rng.random(G * B) < 0.05] = 2500.0            # keep some leaves wider than a stratum
        sh.update_slots(first, G, newp, transform=False)
        ref.update(ridx, newp)
        assert sh.root() == (ref.sum(), ref.min(0))
    s, m, v = sh.tree_arrays()
    assert np.array_equal(v, ref._val) and np.array_equal(s, ref._sum) and np.array_equal(m, ref._min)
    sh.close()


def test_step_loop_matches_stepwise_calls(Shard):
    """The measured loop (two streams, events) equals the same calls issued one by one."""
    import torch
    cap, B, steps = 1 << 16, 128, 24
    struct = [(np.uint8, (512,)), (np.float32, ())]
    rng = np.random.default_rng(77)
    rows = [rng.integers(0, 256, (cap, 512), dtype=np.uint8), rng.standard_normal(cap).astype(np.float32)]
    w0 = (np.abs(rng.standard_normal(cap)) + 0.01).astype(np.float32)
    u = torch.from_numpy(rng.random(steps * B)).cuda()
    prio = torch.from_numpy(np.abs(rng.standard_normal(steps * B)).astype(np.float32)).cuda()
    out = []
    for mode in ("loop", "loop_g4", "calls", "calls_g4"):
        sh = Shard(cap, struct, max_batch=B, ring_depth=8)
        sh.append(rows, w0, alpha=0.6)
        G = 4 if mode.endswith("g4") else 1
        if mode.startswith("loop"):
            sh.step_loop(steps, B, prio, uniforms_dev=u, alpha=0.6, beta=0.4, group=G)
        else:
            for k in range(0, steps, G):
                slot = k % 8
                if G == 1:
                    sh.sample(B, beta=0.4, uniforms=u[k * B:(k + 1) * B], slot=slot)
                    sh.gather(slot)
                    sh.update_slot(slot, prio[k * B:(k + 1) * B], alpha=0.6)
                else:
                    sh.sample_multi(B, slot, G, beta=0.4, uniforms_dev=u[k * B:(k + G) * B])
                    sh.gather_multi(slot, G)
                    sh.update_slots(slot, G, prio[k * B:(k + G) * B], alpha=0.6)
        torch.cuda.synchronize()
        last = [sh.slot_read(s, with_rows=True) for s in range(8)]
        out.append((sh.tree_arrays(), last))
        for data, idx, _ in last:                               # gathered rows are the ring rows
            assert np.array_equal(data[0], rows[0][idx]) and np.array_equal(data[1], rows[1][idx])
        sh.close()
    for a, b in ((0, 2), (1, 3)):
        for x, y in zip(out[a][0], out[b][0]):
            assert np.array_equal(x, y)
        for (da, ia, wa), (db, ib, wb) in zip(out[a][1], out[b][1]):
            assert np.array_equal(ia, ib) and np.array_equal(wa, wb) and np.array_equal(da[0], db[0])


@pytest.mark.parametrize("n_shards", [2, 4])
def test_global_mode_vs_oracle(Shard, n_shards):
    """Mode G: n shards (emulated on one GPU), stratified targets over the global
    cdf, each shard serves its contiguous range with the global minimum -- against n reference
    trees + the same range rule (oracle.global_sample).  One shard is left empty."""
    import torch
    cap, B = 5000, 256
    rng = np.random.default_rng(21)
    shards, trees = [], []
    for g in range(n_shards):
        sh = Shard(cap, [], max_batch=B, store_rows=False)
        t = po.OracleSumTree(cap)
        if g != 1 or n_shards == 2:                      # with 4 shards, shard 1 stays empty
            n = cap if g % 2 == 0 else 1234
            p = rng.random(n) * (1 + 3 * g) + 0.01
            sh.update(np.arange(n), p, transform=False)
            t.update(np.arange(n), p)
        shards.append(sh)
        trees.append(t)
    torch.cuda.synchronize()
    roots = torch.cat([torch.as_tensor(sh.root_dev(), device="cuda").clone() for sh in shards])
    for step in range(4):
        u = rng.random(B)
        covered = 0
        for g, sh in enumerate(shards):
            sh.sample_global(B, roots, n_shards, g, beta=0.5, uniforms=u, slot=step % 2)
            first, count = sh.slot_meta(step % 2)
            _, idx, w = sh.slot_read(step % 2)
            ofirst, oidx, ow, _ = po.global_sample(trees, g, B, u, 0.5)
            assert count == len(oidx) and first == covered
            assert np.array_equal(idx[:count], oidx)
            assert np.allclose(w[:count], ow, rtol=RTOL, atol=0)
            covered += count
        assert covered == B
    # device generator: same seed and draw counter on every shard -> the slices still partition the batch
    covered = 0
    for g, sh in enumerate(shards):
        sh.sample_global(B, roots, n_shards, g, beta=0.5, uniforms=None, seed=99, slot=0)
        first, count = sh.slot_meta(0)
        assert first == covered
        covered += count
    assert covered == B
    for sh in shards:
        sh.close()


def test_global_mode_learner_loop_update_slot(Shard):
    """Mode G learner loop: every step the shard roots are re-read, each shard draws its share of the
    global batch and writes new priorities back with update_slot -- which must touch only the
    `count` entries of the slot that belong to the shard (the rest of the slot is stale)."""
    import torch
    cap, B, n_shards = 3000, 512, 3
    rng = np.random.default_rng(5)
    shards, trees = [], []
    for g in range(n_shards):
        sh = Shard(cap, [], max_batch=B, store_rows=False)
        t = po.OracleSumTree(cap)
        p = rng.random(cap) * (1 + 5 * g) + 0.01
        sh.update(np.arange(cap), p, transform=False)
        t.update(np.arange(cap), p)
        shards.append(sh)
        trees.append(t)
    for sh in shards:                                    # a stale full-size local draw in both slots
        sh.sample(B, uniforms=rng.random(B), slot=0)
        sh.sample(B, uniforms=rng.random(B), slot=1)
    for step in range(6):
        torch.cuda.synchronize()
        roots = torch.cat([torch.as_tensor(sh.root_dev(), device="cuda").clone() for sh in shards])
        u = rng.random(B)
        newp = rng.random(B) * 4 + 0.01
        for g, sh in enumerate(shards):
            sh.sample_global(B, roots, n_shards, g, beta=0.5, uniforms=u, slot=step % 2)
            ofirst, oidx, ow, _ = po.global_sample(trees, g, B, u, 0.5)
            first, count = sh.slot_meta(step % 2)
            assert (first, count) == (ofirst, len(oidx))
        for g, sh in enumerate(shards):                  # write-back after every shard has drawn
            ofirst, oidx, _, _ = po.global_sample(trees, g, B, u, 0.5)
            mine = np.zeros(B)
            mine[:len(oidx)] = newp[ofirst:ofirst + len(oidx)]
            sh.update_slot(step % 2, mine, transform=False)
        # oracle write-back (indices recomputed before any oracle tree changes)
        draws = [po.global_sample(trees, g, B, u, 0.5) for g in range(n_shards)]
        for g, (ofirst, oidx, _, _) in enumerate(draws):
            trees[g].update(oidx, newp[ofirst:ofirst + len(oidx)])
        for g, sh in enumerate(shards):
            s_, m_, v_ = sh.tree_arrays()
            assert np.array_equal(v_, trees[g]._val), (step, g)
            assert np.array_equal(s_, trees[g]._sum), (step, g)
            assert np.array_equal(m_, trees[g]._min), (step, g)
    for sh in shards:
        sh.close()


def test_config4_learner_loop_with_actor_streams(Shard):
    """BASELINE config 4 shape (scaled): per learner step 256 actor messages of 64 rows are appended
    (the ring wraps every 4 steps), then one batch of 512 is drawn and written back; alpha 0.5,
    beta "0.4,1,2000000" (test/apex-dqn/config.yaml:26-27).  Leaves are copied from the engine into
    the oracle after each write (CUDA pow vs libm pow differ in the last ulp) and
    checked to 1e-5; everything downstream -- indices, sums, mins, gathered rows -- is exact."""
    from reth_b200.sampler import Schedule
    cap, B, streams, rows = 1 << 16, 512, 256, 64
    struct = [(np.float32, (4,)), (np.int64, ()), (np.float32, ()), (np.float32, (4,)), (np.float32, ())]
    rng = np.random.default_rng(44)
    sh = Shard(cap, struct, max_batch=B, ring_depth=2)
    ring = po.OracleRing(cap, struct)
    ref = po.OracleSumTree(cap)
    beta = Schedule.from_str("0.4,1,2000000")
    alpha = 0.5
    for step in range(10):
        n = streams * rows
        cols = [rng.standard_normal((n, 4)).astype(np.float32), rng.integers(0, 2, n),
                rng.standard_normal(n).astype(np.float32), rng.standard_normal((n, 4)).astype(np.float32),
                (rng.random(n) < 0.01).astype(np.float32)]
        w = np.abs(rng.standard_normal(n)).astype(np.float32)
        for m in range(0, n, 4096):                           # several actor messages per call
            first = sh.append([c[m:m + 4096] for c in cols], w[m:m + 4096], alpha=alpha)
            slots = ring.append_batch([c[m:m + 4096] for c in cols])
            assert first == slots[0]
        _, _, v = sh.tree_arrays()
        slots = (ring.tail - n + np.arange(n)) % cap
        assert np.allclose(v[slots], po.normalize_weights(w, alpha), rtol=RTOL, atol=0)
        ref.update(slots, v[slots])
        u = rng.random(B)
        sh.sample(B, beta=beta.value(), uniforms=u, slot=step % 2)
        sh.gather(step % 2)
        data, gi, gw = sh.slot_read(step % 2, with_rows=True)
        ri, rv = ref.sample(B, u)
        assert np.array_equal(gi, ri), step
        assert np.allclose(gw, po.is_weights(rv, ref.min(0), beta.value()), rtol=RTOL, atol=0)
        for got, want in zip(data, ring.select(ri)):
            assert np.array_equal(got, want)
        td = np.abs(rng.standard_normal(B)).astype(np.float32)
        sh.update(gi, td, alpha=alpha)
        beta.step()
        _, _, v = sh.tree_arrays()
        ref.update(ri, v[ri])
        assert sh.root() == (ref.sum(), ref.min(0))
    s, m, v = sh.tree_arrays()
    assert np.array_equal(s, ref._sum) and np.array_equal(m, ref._min)
    sh.close()


def test_capacity_2p24_tree_only(Shard):
    """BASELINE config 5 upper end: capacity 2^24 (five tile levels, three sparse ones), batch 8192."""
    cap, B = 1 << 24, 8192
    rng = np.random.default_rng(24)
    sh = Shard(cap, [], max_batch=B, store_rows=False)
    ref = po.OracleSumTree(cap)
    for st in range(0, cap, 1 << 20):
        pw = np.abs(rng.standard_normal(1 << 20)) + 1e-3
        sh.append([], pw, transform=False)
        ref.update(np.arange(st, st + (1 << 20)), pw)
    for t in range(4):
        u = rng.random(B)
        sh.sample(B, beta=0.4, uniforms=u, slot=t % 2)
        _, gi, gw = sh.slot_read(t % 2)
        ri, rv = ref.sample(B, u)
        assert np.array_equal(gi, ri)
        newp = np.abs(rng.standard_normal(B)) + 1e-3
        sh.update_slot(t % 2, newp, transform=False)
        ref.update(ri, newp)
        assert sh.root() == (ref.sum(), ref.min(0))
    s, m, v = sh.tree_arrays()
    assert np.array_equal(s, ref._sum) and np.array_equal(m, ref._min)
    sh.close()


@pytest.mark.parametrize("cap,B,G,groups,rows,frame", [
    (3000, 64, 4, 10, 1000, (4, 84, 84)),     # a third of the ring recycled per group: sampled slots collide with new rows
    (20000, 64, 4, 6, 9000, (64,)),           # write-back + new leaves exceed one repair launch: the cut falls among the new rows
    (20000, 1100, 8, 4, 500, (64,)),          # ... and among the write-back entries
    (3000, 64, 1, 12, 300, (64,)),            # strict alternation (programmatic dependent launches on the tree stream)
])
def test_step_loop_ingest_matches_call_by_call(Shard, cap, B, G, groups, rows, frame):
    """reth_replay_step_loop_ingest (config 4 without draining the loop: the new rows' leaves ride in the write-back's
    repair launch, row copy behind the gathers in flight, later gathers behind the copy) against the same operations
    issued one by one on a twin shard: per group  sample_multi -> gather_multi -> update_slots -> append.  Trees and
    rings bit-identical, the batches of the last group row for row; the ring (which wraps) against the oracle's."""
    import torch
    struct = [(np.uint8, frame), (np.int64, ()), (np.float32, ())]
    rng = np.random.default_rng(61)
    fill = [rng.integers(0, 256, (cap,) + frame, dtype=np.uint8), np.arange(cap), rng.standard_normal(cap).astype(np.float32)]
    fill_w = (rng.random(cap) + 0.05).astype(np.float32)
    new = [rng.integers(0, 256, (rows,) + frame, dtype=np.uint8), 10 ** 6 + np.arange(rows), rng.standard_normal(rows).astype(np.float32)]
    new_w = (rng.random(rows) + 0.05).astype(np.float32)
    a, b = (Shard(cap, struct, max_batch=B, ring_depth=2 * G) for _ in range(2))
    ring = po.OracleRing(cap, struct)
    ring.append_batch(fill)
    for sh in (a, b):
        sh.append(fill, fill_w, alpha=0.5)
    prio = torch.rand(groups * G * B, device="cuda", generator=torch.Generator(device="cuda").manual_seed(3)) + 0.01
    dev_new = [torch.from_numpy(c).cuda() for c in new]
    dev_w = torch.from_numpy(new_w).cuda()
    a.step_loop(groups * G, B, prio, seed=77, alpha=0.5, beta=0.4, group=G, ingest=(dev_new, dev_w))
    for k in range(groups):
        first = (k % 2) * G
        b.sample_multi(B, first, G, beta=0.4, seed=77)
        b.gather_multi(first, G)
        b.update_slots(first, G, prio[k * G * B:(k + 1) * G * B], alpha=0.5)
        b.append(dev_new, dev_w, alpha=0.5)
        ring.append_batch(new)
    torch.cuda.synchronize()
    for x, y in zip(a.tree_arrays(), b.tree_arrays()):
        assert np.array_equal(x, y)
    assert a.stats()["tail"] == b.stats()["tail"] == ring.tail and a.stats()["appended_rows"] == cap + groups * rows
    for c in range(3):
        got = a.ring_read(c, 0, cap)
        assert np.array_equal(got, b.ring_read(c, 0, cap)) and np.array_equal(got, ring.column(c)), c
    last = ((groups - 1) % 2) * G
    for g in range(G):
        da, ia, wa = a.slot_read(last + g, with_rows=True)
        db, ib, wb = b.slot_read(last + g, with_rows=True)
        assert np.array_equal(ia, ib) and np.array_equal(wa, wb)
        for x, y in zip(da, db):
            assert np.array_equal(x, y)
