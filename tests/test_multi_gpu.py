"""Two B200s, one process per GPU (torch.distributed / NCCL): Mode G end to end with cross-shard
batch assembly over NVLink (reth_replay_gather_to), and the shard-root exchange by peer stores into
every rank's roots board (reth_replay_roots_exchange) against the NCCL all-gather it replaces.  Every rank draws its slice of the global
stratified batch and writes rows + indices + weights straight into the learner rank's batch slot;
the assembled batch must equal the two-tree Mode-G oracle row for row.  Skipped with < 2 GPUs
(run with `gpurun --gpus 2`)."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STRUCT = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ())]
CAP, B, STEPS = 4000, 256, 4


def _data(rank):
    rng = np.random.default_rng(500 + rank)
    n = CAP if rank == 0 else 2500
    cols = [rng.integers(0, 256, (n, 4, 84, 84), dtype=np.uint8), np.arange(n) + 100000 * rank,
            rng.standard_normal(n).astype(np.float32)]
    p = rng.random(n) * (1 + 2 * rank) + 0.01
    return cols, p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    try:
        from reth_b200 import ReplayShard
        from reth_b200.sharded import ShardGroup
        sh = ReplayShard(CAP, STRUCT, max_batch=B, ring_depth=4, device=rank)
        cols, p = _data(rank)
        sh.append(cols, None)
        sh.update(np.arange(len(p)), p, transform=False)
        torch.cuda.synchronize()
        sg = ShardGroup(sh)
        sg.connect_learner(0)
        out = []
        for step in range(STEPS):
            u = np.random.default_rng(900 + step).random(B)
            sg.exchange_roots()
            first, count = sg.sample_global(B, beta=0.5, uniforms=u, slot=0)
            sg.assemble(slot=0, dst_slot=2 + step % 2)
            sg.exchange_roots()                                   # orders the learner's reads after every rank's stores
            torch.cuda.synchronize()
            if rank == 0:
                data, idx, w, _ = sh.slot_tensors(2 + step % 2)
                out.append(([d.cpu().numpy() for d in data], idx.cpu().numpy(), w.cpu().numpy()))
            dist.barrier()
        q.put((rank, out))
        dist.barrier()
        sh.close()
    finally:
        dist.destroy_process_group()


def test_mode_g_assembly_over_nvlink():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import __graft_entry__ as entry
    entry.build()
    import torch.multiprocessing as mp
    from oracle import per_oracle as po
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q), daemon=True) for r in range(2)]
    [p.start() for p in procs]
    try:
        res = dict(q.get(timeout=120) for _ in range(2))
        [p.join(30) for p in procs]
    finally:
        for p in procs:
            if p.is_alive():
                p.kill()
    assert all(p.exitcode == 0 for p in procs)
    trees, datas = [], []
    for rank in range(2):
        cols, p = _data(rank)
        t = po.OracleSumTree(CAP)
        t.update(np.arange(len(p)), p)
        trees.append(t)
        datas.append(cols)
    for step in range(STEPS):
        u = np.random.default_rng(900 + step).random(B)
        data, idx, w = res[0][step]
        covered = 0
        for rank in range(2):
            first, oidx, ow, _ = po.global_sample(trees, rank, B, u, 0.5)
            n = len(oidx)
            assert first == covered or n == 0
            sl = slice(covered, covered + n)
            assert np.array_equal(idx[sl], oidx)
            assert np.allclose(w[sl], ow, rtol=1e-5, atol=0)
            for c in range(3):
                assert np.array_equal(data[c][sl], datas[rank][c][oidx]), (step, rank, c)
            covered += n
        assert covered == B


# ---- roots board between two GPUs (CUDA IPC + NVLink peer stores) ----------------------------------
GB2, STEPS2, SEED2 = 512, 12, 4321


def _board_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    try:
        from reth_b200 import ReplayShard
        from reth_b200.sharded import ShardGroup
        cols, p = _data(rank)

        def make():
            sh = ReplayShard(CAP, STRUCT, max_batch=GB2, ring_depth=4, device=rank)
            sh.append(cols, None)
            sh.update(np.arange(len(p)), p, transform=False)
            return sh
        a, b = make(), make()                      # a: roots board + device loop; b: NCCL, call by call
        ga, gb = ShardGroup(a), ShardGroup(b)
        ga.connect_board()
        prio = torch.rand(STEPS2 * GB2, device="cuda", generator=torch.Generator(device="cuda").manual_seed(70 + rank)) + 0.01
        # (1) the exchange alone: same 4 doubles as the NCCL all-gather, for changing roots
        same = True
        for k in range(5):
            ga.exchange_roots()
            gb.exchange_roots()
            torch.cuda.synchronize()
            same = same and torch.equal(ga.all_roots, gb.all_roots)
            idx = torch.arange(64 * k, 64 * k + 64, device="cuda")
            newp = torch.full((64,), 0.5 + k + rank, device="cuda", dtype=torch.float64)
            a.update(idx, newp, transform=False)
            b.update(idx, newp, transform=False)
        # (2) Mode G device loop on the board vs the same steps call by call over NCCL
        a.step_loop(STEPS2, GB2, prio, seed=SEED2, alpha=0.6, beta=0.4, global_mode=True)
        metas = []
        for k in range(STEPS2):
            gb.exchange_roots()
            metas.append(gb.sample_global(GB2, beta=0.4, seed=SEED2, slot=k % 4))
            b.gather(k % 4)
            b.update_slot(k % 4, prio[k * GB2:(k + 1) * GB2], alpha=0.6)
        torch.cuda.synchronize()
        dist.barrier()
        last = (STEPS2 - 1) % 4
        da, ia, wa, _ = a.slot_tensors(last)
        db, ib, wb, _ = b.slot_tensors(last)
        n = metas[-1][1]
        rows_same = all(torch.equal(x[:n], y[:n]) for x, y in zip(da, db)) and torch.equal(ia[:n], ib[:n]) \
            and torch.equal(wa[:n], wb[:n])
        q.put((rank, {"exchange_same": bool(same), "tree_a": a.tree_arrays(), "tree_b": b.tree_arrays(),
                      "meta_a": tuple(a.slot_meta(last)), "metas_b": [tuple(m) for m in metas], "rows_same": bool(rows_same)}))
        dist.barrier()
        a.close()
        b.close()
    finally:
        dist.destroy_process_group()


def test_roots_board_and_mode_g_loop_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import __graft_entry__ as entry
    entry.build()
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_board_worker, args=(r, 2, port, q), daemon=True) for r in range(2)]
    [p.start() for p in procs]
    try:
        res = dict(q.get(timeout=180) for _ in range(2))
        [p.join(30) for p in procs]
    finally:
        for p in procs:
            if p.is_alive():
                p.kill()
    assert all(p.exitcode == 0 for p in procs)
    for rank in range(2):
        r = res[rank]
        assert r["exchange_same"]
        for x, y in zip(r["tree_a"], r["tree_b"]):
            assert np.array_equal(x, y), rank
        assert r["meta_a"] == r["metas_b"][-1] and r["rows_same"]
    # the two shares partition every global batch
    for k in range(STEPS2):
        (f0, c0), (f1, c1) = res[0]["metas_b"][k], res[1]["metas_b"][k]
        assert f0 == 0 and f1 == c0 and c0 + c1 == GB2
