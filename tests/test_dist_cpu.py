"""Multi-shard host logic on CPU: world-size-2 gloo process group, one oracle tree per rank,
all-gather of the shard roots and the Mode-G range rule (reth_b200/sharded.py).  Every rank must
serve a contiguous slice of the global batch, the slices must partition it, and the per-rank
indices must equal the single-process Mode-G oracle.  Mode R (round-robin, trainer.py:33-36) is
checked alongside."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import per_oracle as po
    from reth_b200 import sharded
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        cap, B = 3000, 64
        rng = np.random.default_rng(100 + rank)
        tree = po.OracleSumTree(cap)
        tree.update(np.arange(cap), rng.random(cap) * (1 + rank) + 0.01)
        mine = torch.tensor([tree.sum(), tree._min[0]], dtype=torch.float64)
        roots = torch.zeros(2 * world, dtype=torch.float64)
        dist.all_gather_into_tensor(roots, mine)
        out = []
        for step in range(5):
            u = np.random.default_rng(7 + step).random(B)        # the same uniforms on every rank
            first, count, local, pmin = sharded.global_quota(roots.numpy(), rank, B, u)
            idx = np.array([tree.find_index(t) for t in local], np.int64)
            w = po.is_weights(tree._val[idx], pmin, 0.4)
            out.append((first, count, idx, w, sharded.owner_round_robin(step, world)))
        q.put((rank, roots.numpy().copy(), out))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_mode_g_gloo_world2():
    import torch.multiprocessing as mp
    from oracle import per_oracle as po
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    world = 2
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    [p.start() for p in procs]
    res = {}
    for _ in range(world):
        rank, roots, out = q.get(timeout=120)
        res[rank] = (roots, out)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert np.array_equal(res[0][0], res[1][0])                  # every rank holds the same gathered roots
    # single-process oracle over the same two trees
    cap, B = 3000, 64
    trees = []
    for rank in range(world):
        rng = np.random.default_rng(100 + rank)
        t = po.OracleSumTree(cap)
        t.update(np.arange(cap), rng.random(cap) * (1 + rank) + 0.01)
        trees.append(t)
    assert res[0][0][0] == trees[0].sum() and res[0][0][2] == trees[1].sum()
    for step in range(5):
        u = np.random.default_rng(7 + step).random(B)
        covered = 0
        for rank in range(world):
            first, count, idx, w, owner = res[rank][1][step]
            ofirst, oidx, ow, _ = po.global_sample(trees, rank, B, u, 0.4)
            assert first == covered and count == len(oidx)
            if count:
                assert first == ofirst
            assert np.array_equal(idx, oidx)
            assert np.allclose(w, ow, rtol=1e-12)
            assert owner == step % world
            covered += count
        assert covered == B                                      # the slices partition the global batch


def test_quota_edges():
    sys.path.insert(0, ROOT)
    from reth_b200 import sharded
    B = 16
    u = np.full(B, 0.5)
    # an empty shard in the middle serves nothing; weights ignore its min
    roots = np.array([10.0, 0.5, 0.0, 0.0, 30.0, 0.25])
    tot, prefix, pmin = sharded.combine_roots(roots)
    assert tot == 40.0 and list(prefix) == [0.0, 10.0, 10.0] and pmin == 0.25
    f0, c0, l0, _ = sharded.global_quota(roots, 0, B, u)
    f1, c1, l1, _ = sharded.global_quota(roots, 1, B, u)
    f2, c2, l2, _ = sharded.global_quota(roots, 2, B, u)
    assert (f0, c0) == (0, 4) and c1 == 0 and (f2, c2) == (4, 12)
    assert np.all(l0 < 10.0) and np.all(l2 >= 0) and np.all(l2 < 30.0)
    # all shards empty: total 0 -> every target is 0 and lands on the last shard's range rule
    z = np.zeros(4)
    assert sharded.combine_roots(z)[2] == 1.0


def test_packed_quota_multi_matches_oracle():
    """Host statement of the Mode-G prefetch group (sharded.global_quota_multi): G global draws from the same
    roots, every shard's shares packed back to back -- against the n-tree oracle draw by draw."""
    sys.path.insert(0, ROOT)
    from oracle import per_oracle as po
    from reth_b200 import sharded
    cap, B, G, n = 2000, 48, 5, 3
    rng = np.random.default_rng(9)
    trees = []
    for s in range(n):
        t = po.OracleSumTree(cap)
        m = cap if s != 1 else 300
        t.update(np.arange(m), rng.random(m) * (1 + 4 * s) + 0.01)
        trees.append(t)
    roots = np.array([x for t in trees for x in (t.sum(), t._min[0])])
    u = rng.random(G * B)
    covered = np.zeros(G, np.int64)
    for s in range(n):
        quota, local, pmin = sharded.global_quota_multi(roots, s, B, u, G)
        idx = np.array([trees[s].find_index(t) for t in local], np.int64)
        off = 0
        for g in range(G):
            ofirst, oidx, ow, _ = po.global_sample(trees, s, B, u[g * B:(g + 1) * B], 0.4)
            first, count, o = quota[g]
            assert count == len(oidx) and o == off and first == covered[g]
            assert np.array_equal(idx[off:off + count], oidx)
            assert np.allclose(po.is_weights(trees[s]._val[oidx], pmin, 0.4), ow, rtol=1e-12)
            off += count
            covered[g] += count
        assert off == len(local)
    assert np.all(covered == B)
