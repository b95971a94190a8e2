"""Shard roots over NVLink (reth_replay_board_* / roots_exchange): the device-side all-gather of the
16-byte shard roots that Mode G needs before every global stratified draw, and the
Mode-G device loop built on it.  Here n shards live on ONE GPU in one process (boards reached
through plain pointers); tests/test_multi_gpu.py runs the same exchange between two GPUs in two
processes (CUDA IPC + NVLink)."""
import threading

import numpy as np
import pytest

from oracle import per_oracle as po

pytestmark = pytest.mark.gpu
RTOL = 1e-5


@pytest.fixture(scope="module")
def Shard():
    import __graft_entry__ as entry
    entry.build()
    from reth_b200 import ReplayShard
    return ReplayShard


def device_uniforms(seed, base, n):
    """Host restatement of the engine's counter-based generator (tree.cuh uniform53)."""
    with np.errstate(over="ignore"):
        ctr = np.arange(base, base + n, dtype=np.uint64)
        z = np.uint64(seed) + (ctr + np.uint64(1)) * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z ^= z >> np.uint64(31)
    return (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def _connect(shards):
    blobs = [sh.board_export() for sh in shards]
    for g, sh in enumerate(shards):
        sh.board_connect(blobs, g)


def _exchange_all(shards):
    # one host thread, one stream: every shard publishes before any shard waits
    for sh in shards:
        sh.roots_exchange(1)
    for sh in shards:
        sh.roots_exchange(2)


def test_board_exchange_matches_roots_and_oracle(Shard):
    import torch
    cap, B, n_shards = 5000, 256, 3
    rng = np.random.default_rng(77)
    shards, trees = [], []
    for g in range(n_shards):
        sh = Shard(cap, [], max_batch=B, store_rows=False)
        t = po.OracleSumTree(cap)
        n = cap if g != 1 else 1500
        p = rng.random(n) * (1 + 2 * g) + 0.01
        sh.update(np.arange(n), p, transform=False)
        t.update(np.arange(n), p)
        shards.append(sh)
        trees.append(t)
    _connect(shards)
    views = [torch.as_tensor(sh.all_roots_dev(n_shards), device="cuda") for sh in shards]
    for step in range(5):                                  # odd and even exchange numbers
        _exchange_all(shards)
        torch.cuda.synchronize()
        want = np.concatenate([torch.as_tensor(sh.root_dev(), device="cuda").cpu().numpy() for sh in shards])
        assert np.array_equal(want[0::2], [t.sum() for t in trees])
        for v in views:
            assert np.array_equal(v.cpu().numpy(), want)
        u = rng.random(B)
        covered = 0
        for g, sh in enumerate(shards):
            sh.sample_global(B, views[g], n_shards, g, beta=0.5, uniforms=u, slot=0)
            first, count = sh.slot_meta(0)
            _, idx, w = sh.slot_read(0)
            ofirst, oidx, ow, _ = po.global_sample(trees, g, B, u, 0.5)
            assert count == len(oidx) and first == covered
            assert np.array_equal(idx[:count], oidx)
            assert np.allclose(w[:count], ow, rtol=RTOL, atol=0)
            covered += count
        assert covered == B
        for g, sh in enumerate(shards):                    # the roots change before the next exchange
            k = rng.integers(0, 1500, 64)
            p = rng.random(64) * 3 + 0.01
            sh.update(k, p, transform=False)
            trees[g].update(k, p)
    for sh in shards:
        sh.close()


def test_board_errors_and_timeout(Shard, monkeypatch):
    from reth_b200._cabi import ReplayError
    monkeypatch.setenv("RETH_BOARD_TIMEOUT_MS", "30")
    a = Shard(64, [], max_batch=8, store_rows=False)
    b = Shard(64, [], max_batch=8, store_rows=False)
    for sh in (a, b):
        sh.update(np.arange(64), np.ones(64), transform=False)
    with pytest.raises(ReplayError):                       # not connected yet
        a.roots_exchange(3)
    blobs = [a.board_export(), b.board_export()]
    with pytest.raises(ReplayError):
        a.board_connect(blobs, 2)
    a.board_connect(blobs, 0)
    b.board_connect(blobs, 1)
    with pytest.raises(ReplayError):
        a.board_connect(blobs, 0)                          # already connected
    with pytest.raises(ReplayError):
        b.roots_exchange(2)                                # wait before any publish
    a.roots_exchange(3)                                    # b never publishes: the wait gives up
    with pytest.raises(ReplayError) as ei:
        a.sample(8, uniforms=np.full(8, 0.5), slot=0)
        a.slot_read(0)
    assert ei.value.code == -5 and "roots exchange" in str(ei.value)
    # the group recovers once the late shard catches up with the same exchange number
    b.roots_exchange(1)
    a.roots_exchange(2)
    b.roots_exchange(2)
    a.sample(8, uniforms=np.full(8, 0.5), slot=0)
    a.slot_read(0)
    import torch
    assert np.array_equal(torch.as_tensor(a.all_roots_dev(2), device="cuda").cpu().numpy(), [64.0, 1.0, 64.0, 1.0])
    a.close()
    b.close()


@pytest.mark.parametrize("cap,GB,n_shards,G,heavy", [(3000, 128, 3, 4, 1.0), (20000, 8192, 4, 4, 40.0)])
def test_mode_g_prefetch_group_packed(Shard, cap, GB, n_shards, G, heavy):
    """sample_global_multi: G global stratified draws in one launch, this shard's shares packed back to back;
    gather_multi / update_slots treat the slots as one batch of sum(count) rows.  Against n reference trees +
    the range rule per draw, rows byte for byte, trees bit for bit after the write-back.
    Second case: shard 0 holds most of the priority mass, so its share (about 0.8 x 32768 entries) is far beyond what
    the write-back's ordinary launches are sized for (1.25 x an equal share) and the catch-all launch runs two rounds;
    the draw kernel's warps stride over a share three times their number."""
    import torch
    struct = [(np.uint8, (48,)), (np.int64, ())]
    rng = np.random.default_rng(11)
    shards, trees, datas = [], [], []
    for g in range(n_shards):
        sh = Shard(cap, struct, max_batch=GB, ring_depth=2 * G)
        cols = [rng.integers(0, 256, (cap, 48), dtype=np.uint8), np.arange(cap) + 100000 * g]
        p = (rng.random(cap) * (1 + 3 * g) + 0.01) * (heavy if g == 0 else 1.0)
        sh.append(cols, None)
        sh.update(np.arange(cap), p, transform=False)
        t = po.OracleSumTree(cap)
        t.update(np.arange(cap), p)
        shards.append(sh); trees.append(t); datas.append(cols)
    for step in range(3):
        torch.cuda.synchronize()
        roots = torch.cat([torch.as_tensor(sh.root_dev(), device="cuda").clone() for sh in shards])
        u = rng.random(G * GB)
        u_dev = torch.from_numpy(u).cuda()
        first_slot = (step % 2) * G
        newp = rng.random(G * GB) * 2 + 0.01
        draws = [[po.global_sample(trees, g, GB, u[d * GB:(d + 1) * GB], 0.5) for d in range(G)] for g in range(n_shards)]
        for d in range(G):                                   # the shares of every draw partition its global batch
            assert sum(len(draws[g][d][1]) for g in range(n_shards)) == GB
        for g, sh in enumerate(shards):
            sh.sample_global_multi(GB, roots, n_shards, g, first_slot, G, beta=0.5, uniforms_dev=u_dev)
            q = sh.slot_quota(first_slot, G)
            oidx = np.concatenate([draws[g][d][1] for d in range(G)])
            ow = np.concatenate([draws[g][d][2] for d in range(G)])
            off = 0
            for d in range(G):
                n = len(draws[g][d][1])
                assert q[d, 1] == n and q[d, 2] == off
                if n:
                    assert q[d, 0] == draws[g][d][0]
                off += n
            assert tuple(sh.slot_meta(first_slot)) == (0, off)
            idx, w = sh.group_read(first_slot, G, with_weights=True)
            assert np.array_equal(idx[:off], oidx)
            assert np.allclose(w[:off], ow, rtol=RTOL, atol=0)
            sh.gather_multi(first_slot, G)
            torch.cuda.synchronize()
            got = [torch.cat([sh.slot_tensors(first_slot + d)[0][c] for d in range(G)]).cpu().numpy() for c in range(2)]
            for c in range(2):
                assert np.array_equal(got[c][:off], datas[g][c][oidx]), (step, g, c)
            mine = np.zeros(G * GB)
            mine[:off] = newp[:off]
            sh.update_slots(first_slot, G, torch.from_numpy(mine).cuda() if G * GB > 16384 else mine, transform=False)
        for g in range(n_shards):                            # oracle write-back after every shard has drawn
            oidx = np.concatenate([draws[g][d][1] for d in range(G)])
            trees[g].update(oidx, newp[:len(oidx)])
        for g, sh in enumerate(shards):
            for x, y in zip(sh.tree_arrays(), (trees[g]._sum, trees[g]._min, trees[g]._val)):
                assert np.array_equal(x, y), (step, g)
    for sh in shards:
        sh.close()


def _loop_body(G=1):
    import torch
    from reth_b200 import ReplayShard as Shard
    cap, GB, n_shards, steps, seed = 4096, 256, 2, 8, 1234
    struct = [(np.uint8, (64,)), (np.int64, ())]
    rng = np.random.default_rng(3)
    fills = []
    for g in range(n_shards):
        cols = [rng.integers(0, 256, (cap, 64), dtype=np.uint8), np.arange(cap) + 10000 * g]
        p = (rng.random(cap) * (1 + 4 * g) + 0.01).astype(np.float32)
        fills.append((cols, p))
    prio = [torch.rand(steps * GB, device="cuda", generator=torch.Generator(device="cuda").manual_seed(50 + g)) + 0.01
            for g in range(n_shards)]

    def make():
        out = []
        for g in range(n_shards):
            sh = Shard(cap, struct, max_batch=GB, ring_depth=4)
            sh.append(fills[g][0], fills[g][1], alpha=0.6)
            out.append(sh)
        _connect(out)
        return out

    loop, step = make(), make()
    torch.cuda.synchronize()
    # call by call, with the oracle beside it
    trees = []
    for g in range(n_shards):
        s_, m_, v_ = step[g].tree_arrays()
        t = po.OracleSumTree(cap)
        t.update(np.arange(cap), v_)
        trees.append(t)
    views = [torch.as_tensor(sh.all_roots_dev(n_shards), device="cuda") for sh in step]
    for k in range(0, steps, G):
        _exchange_all(step)
        us = [device_uniforms(seed, (k + d) * GB, GB) for d in range(G)]
        draws = [np.concatenate([po.global_sample(trees, g, GB, us[d], 0.4)[1] for d in range(G)]) for g in range(n_shards)]
        assert sum(len(x) for x in draws) == G * GB
        for g, sh in enumerate(step):
            if G == 1:
                sh.sample_global(GB, views[g], n_shards, g, beta=0.4, seed=seed, slot=k % 4)
                first, count = sh.slot_meta(k % 4)
                _, idx, w = sh.slot_read(k % 4)
                assert first == (0 if g == 0 else len(draws[0])), (k, g)
                sh.gather(k % 4)
                sh.update_slot(k % 4, prio[g][k * GB:(k + 1) * GB], alpha=0.6)
            else:
                sh.sample_global_multi(GB, views[g], n_shards, g, k % 4, G, beta=0.4, seed=seed)
                first, count = sh.slot_meta(k % 4)
                idx, w = sh.group_read(k % 4, G, with_weights=True)
                sh.gather_multi(k % 4, G)
                sh.update_slots(k % 4, G, prio[g][k * GB:(k + G) * GB], alpha=0.6)
            assert count == len(draws[g]), (k, g)
            assert np.array_equal(idx[:count], draws[g])
        for g, sh in enumerate(step):                      # oracle leaves follow the engine's float32 pow
            _, _, v_ = sh.tree_arrays()
            trees[g].update(draws[g], v_[draws[g]])
    # device loop, AFTER the call-by-call twin has loaded every kernel (a lazy module load while a peer's
    # exchange kernel spins would stall on it): one host thread per shard, each with its own caller
    # stream (on a shared stream the second loop would queue behind the first one's join)
    errs = []
    streams = [torch.cuda.Stream() for _ in range(n_shards)]

    def run(g):
        try:
            loop[g].step_loop(steps, GB, prio[g], seed=seed, alpha=0.6, beta=0.4, global_mode=True, group=G,
                              stream=streams[g])
        except Exception as e:                              # pragma: no cover
            errs.append(e)
    th = [threading.Thread(target=run, args=(g,)) for g in range(n_shards)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    torch.cuda.synchronize()
    torch.cuda.synchronize()
    for g in range(n_shards):
        for x, y in zip(loop[g].tree_arrays(), step[g].tree_arrays()):
            assert np.array_equal(x, y), g
        last = (steps - G) % 4
        first, count = loop[g].slot_meta(last)
        assert (first, count) == tuple(step[g].slot_meta(last))

        def packed(sh):
            parts = [sh.slot_tensors(last + d) for d in range(G)]
            cols = [torch.cat([p[0][c] for p in parts]) for c in range(len(struct))]
            return cols, torch.cat([p[1] for p in parts]), torch.cat([p[2] for p in parts])
        (da, ia, wa), (db, ib, wb) = packed(loop[g]), packed(step[g])
        assert torch.equal(ia[:count], ib[:count]) and torch.equal(wa[:count], wb[:count])
        for x, y in zip(da, db):
            assert torch.equal(x[:count], y[:count])
    for sh in loop + step:
        sh.close()


def _loop_worker(q):
    import traceback
    try:
        _loop_body(1)
        _loop_body(2)                                       # prefetch group: two global draws per exchange, packed shares
        q.put("ok")
    except BaseException:                                   # report, never hang the parent
        q.put(traceback.format_exc())


def test_mode_g_device_loop_matches_stepwise_calls(Shard, monkeypatch):
    """reth_replay_step_loop with the Mode-G flag (exchange -> global draw -> gather || write-back of the
    share, all enqueued without a host round trip) against the same steps issued call by call on twin
    shards, and the draws against the oracle's range rule with the engine's own uniforms.
    Two shards on ONE device wait for each other's kernels, so their streams must not share a hardware
    work queue (the default 8 connections alias the 18 streams of this test): the body runs in a child
    process started with CUDA_DEVICE_MAX_CONNECTIONS=32 and eager module loading.  One shard per GPU
    needs none of this."""
    import torch.multiprocessing as mp
    monkeypatch.setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32")
    monkeypatch.setenv("CUDA_MODULE_LOADING", "EAGER")
    monkeypatch.setenv("RETH_BOARD_TIMEOUT_MS", "20000")     # co-scheduling hiccups on one device must not fail the wait
    ctx = mp.get_context("spawn")
    res = None
    for attempt in range(2):                                 # two shards sharing one GPU: allow one retry
        q = ctx.Queue()
        p = ctx.Process(target=_loop_worker, args=(q,), daemon=True)
        p.start()
        try:
            res = q.get(timeout=300)
            p.join(30)
        finally:
            if p.is_alive():
                p.kill()
        if res == "ok":
            break
    assert res == "ok", res
