"""Mode G leg of `bench.py --gpus N` (N > 1): the coupled multi-GPU path of BASELINE.json config 3.

Mode G = north_star's "shard priority totals combined ...": stratified targets over the GLOBAL CDF of all
shards, global p_min in the importance weights; every rank draws, gathers and writes back its share of each
global batch (data-parallel learners).  The reference has no counterpart: its learner polls whole batches shard
by shard with no cross-shard normalisation (test/apex-dqn/trainer.py:33-36,52-61).

Two parts, both run by every rank on the very shard the Mode R numbers were measured on:

  parity leg (untimed)   roots over the NVLink board == NCCL all-gather bit for bit; each rank's share
                         {first, count} == the host range rule sharded.global_quota(_multi); sampled indices
                         == a numpy restatement of the reference descent (utils/sumtree.py:34-58) over the
                         rank's own tree arrays with the rule's local targets; IS weights with the global
                         minimum (1e-5); gathered rows == the ring rows; after the write-back every _sum/_min
                         entry == a level-by-level recompute from the leaves, touched leaves == (w+1e-6)^alpha;
                         the ranks' shares partition the global batch.  Result all-reduced -> parity_ok.
  timed leg              reth_replay_step_loop in Mode G (bit6): per iteration ONE roots exchange (peer stores
                         into every rank's board + a wait on the own board, no host round trip), G global
                         draws with packed shares, one gather || one write-back of the share.  CUDA events,
                         max over ranks.

The numpy checker below is bench-side verification only (the product path never calls it).
"""
import numpy as np

from reth_b200 import sharded


# ---------------------------------------------------------------- numpy checkers --
def np_find_index(s_sum, s_val, targets):
    """Vectorised _numba_find_index (utils/sumtree.py:34-58) over float64 heap-order arrays."""
    cap = len(s_val)
    t = np.array(targets, dtype=np.float64)
    cur = np.zeros(len(t), np.int64)
    done = np.zeros(len(t), bool)
    for _ in range(66):
        if done.all():
            break
        left = 2 * cur + 1
        has_l = (left < cap) & ~done
        sl = np.where(has_l, s_sum[np.minimum(left, cap - 1)], 0.0)
        go_l = has_l & (t < sl)
        cur = np.where(go_l, left, cur)
        rest = ~done & ~go_l
        t = np.where(rest & has_l, t - sl, t)
        v = s_val[cur]
        hit = rest & (t < v + 1e-5)
        done |= hit
        rest &= ~hit
        t = np.where(rest, t - v, t)
        right = 2 * cur + 2
        no_r = rest & (right >= cap)
        done |= no_r
        rest &= ~no_r
        cur = np.where(rest, right, cur)
    return cur


def np_tree_recompute(val):
    """_sum / _min of a tree from its leaves (_numba_maintain_node, utils/sumtree.py:5-21):
    sum = (val + sum_l) + sum_r;  min = val (1 when 0), then the children's non-zero minima.  A node has been
    maintained -- and carries a non-zero min -- iff it or a descendant was ever written (written values are
    positive: (w + 1e-6) ** alpha); never-maintained nodes keep sum = min = 0 as in the reference."""
    cap = len(val)
    s = np.zeros(cap + 2)
    m = np.zeros(cap + 2)
    live = np.zeros(cap + 2, bool)
    d = int(np.floor(np.log2(cap)))
    while d >= 0:
        lo, hi = (1 << d) - 1, min((2 << d) - 1, cap)
        i = np.arange(lo, hi)
        l, r = np.minimum(2 * i + 1, cap), np.minimum(2 * i + 2, cap + 1)   # s[cap], s[cap+1] stay 0
        s[i] = (val[i] + s[l]) + s[r]
        mn = np.where(val[i] != 0.0, val[i], 1.0)
        ml, mr = m[l], m[r]
        mn = np.where((ml != 0.0) & (ml < mn), ml, mn)
        mn = np.where((mr != 0.0) & (mr < mn), mr, mn)
        live[i] = (val[i] != 0.0) | live[l] | live[r]
        m[i] = np.where(live[i], mn, 0.0)
        d -= 1
    return s[:cap], m[:cap]


def _fail(msgs, rank, what):
    msgs.append(f"rank {rank}: {what}")
    return False


def parity_leg(torch, dist, shard, sg, world, rank, GB, alpha, beta, log):
    """-> (ok: bool, detail: dict).  Collective: every rank runs the same steps."""
    msgs, ok = [], True
    dev = torch.device("cuda", shard.device)
    roots_nccl = torch.zeros(2 * world, dtype=torch.float64, device=dev)
    mine = torch.as_tensor(shard.root_dev(), device=dev)
    checked = 0
    for step in range(3):
        u = np.random.default_rng(7000 + step).random(GB)            # the same uniforms on every rank
        sg.exchange_roots()                                           # roots board (NVLink peer stores)
        dist.all_gather_into_tensor(roots_nccl, mine)                 # the collective it replaces
        torch.cuda.synchronize()
        roots = sg.all_roots.cpu().numpy().copy()
        if not np.array_equal(roots, roots_nccl.cpu().numpy()):
            ok = _fail(msgs, rank, f"step {step}: board roots != NCCL all-gather")
        s_sum, s_min, s_val = shard.tree_arrays()
        first, count = sg.sample_global(GB, beta=beta, uniforms=u, slot=0)
        qf, qc, local_t, pmin = sharded.global_quota(roots, rank, GB, u)
        if (first, count) != (qf, qc):
            ok = _fail(msgs, rank, f"step {step}: share {(first, count)} != range rule {(qf, qc)}")
            continue
        shard.gather(0)
        data, idx, w = shard.slot_read(0, with_rows=True, rows=count)
        want = np_find_index(s_sum, s_val, local_t)
        if not np.array_equal(idx, want):
            ok = _fail(msgs, rank, f"step {step}: {int((idx != want).sum())} of {count} sampled indices differ")
        ww = (s_val[want] / pmin) ** (-beta)
        if count and not np.allclose(w, ww, rtol=1e-5, atol=0):
            ok = _fail(msgs, rank, f"step {step}: IS weights differ (max rel {np.max(np.abs(w / ww - 1)):.2e})")
        for j in np.unique(np.linspace(0, max(count - 1, 0), 6).astype(int)) if count else []:
            for c in (0, 1, 3):
                if not np.array_equal(data[c][j], shard.ring_read(c, int(idx[j]), 1)[0]):
                    ok = _fail(msgs, rank, f"step {step}: gathered row {j} column {c} != ring row {int(idx[j])}")
        # the shares partition the global batch
        fc = torch.tensor([first, count], device=dev, dtype=torch.int64)
        allfc = torch.zeros(2 * world, device=dev, dtype=torch.int64)
        dist.all_gather_into_tensor(allfc, fc)
        allfc = allfc.cpu().numpy().reshape(world, 2)
        if allfc[0, 0] != 0 or allfc[:, 1].sum() != GB or not np.array_equal(allfc[1:, 0], np.cumsum(allfc[:-1, 1])):
            ok = _fail(msgs, rank, f"step {step}: shares do not partition the global batch: {allfc.tolist()}")
        # write-back of the share, then the tree against a recompute
        newp = np.abs(np.random.default_rng(8000 + 10 * step + rank).standard_normal(GB)).astype(np.float32) + 0.01
        shard.update_slot(0, newp, alpha=alpha)
        t_sum, t_min, t_val = shard.tree_arrays()
        r_sum, r_min = np_tree_recompute(t_val)
        if not (np.array_equal(t_sum, r_sum) and np.array_equal(t_min, r_min)):
            ok = _fail(msgs, rank, f"step {step}: tree sums / minima differ from the recompute")
        if count:
            last = np.r_[idx[1:] != idx[:-1], True]                    # duplicates are adjacent: the last one wins
            leaf = ((newp[:count][last] + np.float32(1e-6)).astype(np.float64) ** float(np.float32(alpha))).astype(np.float32)
            if not np.allclose(t_val[idx[last]], leaf.astype(np.float64), rtol=3e-7, atol=0):
                ok = _fail(msgs, rank, f"step {step}: written leaves differ from (w+1e-6)^alpha")
            untouched = np.ones(len(t_val), bool)
            untouched[idx] = False
            if not np.array_equal(t_val[untouched], s_val[untouched]):
                ok = _fail(msgs, rank, f"step {step}: a leaf outside the share changed")
        checked += count
    # packed prefetch group: G global draws in one launch, shares back to back
    G = 4
    if shard.ring_depth >= G:
        ud = torch.rand(G * GB, dtype=torch.float64, device=dev, generator=torch.Generator(device=dev).manual_seed(99))
        sg.exchange_roots()
        torch.cuda.synchronize()
        roots = sg.all_roots.cpu().numpy().copy()
        s_sum, s_min, s_val = shard.tree_arrays()
        quota = sg.sample_global(GB, beta=beta, uniforms=ud, slot=0, groups=G)
        want_q, local_t, pmin = sharded.global_quota_multi(roots, rank, GB, ud.cpu().numpy(), G)
        if not np.array_equal(quota, want_q):
            ok = _fail(msgs, rank, f"packed group: quota {quota.tolist()} != range rule {want_q.tolist()}")
        else:
            n = int(want_q[:, 1].sum())
            shard.gather_multi(0, G)
            idx, w = shard.group_read(0, G, with_weights=True)
            want = np_find_index(s_sum, s_val, local_t)
            if not np.array_equal(idx[:n], want):
                ok = _fail(msgs, rank, f"packed group: {int((idx[:n] != want).sum())} of {n} indices differ")
            if n and not np.allclose(w[:n], (s_val[want] / pmin) ** (-beta), rtol=1e-5, atol=0):
                ok = _fail(msgs, rank, "packed group: IS weights differ")
            cols, _, _, _ = shard.slot_tensors(0, rows=n) if n else ([], None, None, None)
            for j in np.unique(np.linspace(0, max(n - 1, 0), 6).astype(int)) if n else []:
                if not np.array_equal(cols[0][j].cpu().numpy(), shard.ring_read(0, int(idx[j]), 1)[0]):
                    ok = _fail(msgs, rank, f"packed group: gathered row {j} != ring row {int(idx[j])}")
            checked += n
    flag = torch.tensor([1 if ok else 0], device=dev, dtype=torch.int32)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    for m in msgs:
        log("[mode_g parity] " + m)
    return bool(flag.item()), {"steps": 3, "packed_group": G, "rows_checked_rank0": int(checked), "global_batch": GB}


def timed_case(torch, dist, shard, GB, G, iters, alpha, beta, seed):
    """Mode G device loop: `iters` iterations of G global batches each -> us per global batch (max over ranks)."""
    depth = max(4, 3 * G)                                         # the tree chain may run two groups ahead of the gather
    if shard.max_batch != GB or shard.ring_depth != depth:
        shard.resize_batch_ring(GB, depth)
    steps = iters * G
    dev = torch.device("cuda", shard.device)
    prio = torch.rand(min(steps, 64 * G) * GB, device=dev, generator=torch.Generator(device=dev).manual_seed(seed)) + 0.01
    chunk = prio.numel() // GB                                    # steps per step_loop call (a multiple of G)
    shard.step_loop(min(chunk, 8 * G), GB, prio, seed=7, alpha=alpha, beta=beta, global_mode=True, group=G)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = shard.stats()["kernel_launches"]
    e0.record()
    done = 0
    while done < steps:
        n = min(chunk, steps - done)
        shard.step_loop(n, GB, prio, seed=7 + done, alpha=alpha, beta=beta, global_mode=True, group=G)
        done += n
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    dist.barrier()
    us = float(ms.item()) * 1e3 / steps
    return {"global_batch": GB, "prefetch_group": G, "us_per_batch": us, "value": GB / us * 1e6,
            "iterations": iters, "gpu_launches": int(shard.stats()["kernel_launches"] - l0)}


def run(torch, dist, shard, world, rank, batch, alpha, beta, mode_r_value, log, iters=400):
    """Parity leg + timed cases on `shard` (its batch ring is re-shaped for the global batch).
    Cases: global batch = world x batch (each rank's share is about the Mode R batch) and 4x that,
    strict alternation (G = 1) and prefetch group 8."""
    from reth_b200.sharded import ShardGroup
    GB = world * batch
    shard.resize_batch_ring(GB, 8)
    sg = ShardGroup(shard)
    sg.connect_board()
    ok, detail = parity_leg(torch, dist, shard, sg, world, rank, GB, alpha, beta, log)
    cases = []
    for gb, G, it in ((GB, 8, iters), (GB, 1, iters), (4 * GB, 8, max(50, iters // 4)), (4 * GB, 1, max(50, iters // 4))):
        cases.append(timed_case(torch, dist, shard, gb, G, it, alpha, beta, seed=300 + rank))
        if rank == 0:
            log(f"[mode_g] global batch {gb} group {G}: {cases[-1]['us_per_batch']:.1f} us / batch")
    head = cases[0]
    return {"value": head["value"], "unit": "transitions/s", "us_per_batch": head["us_per_batch"],
            "global_batch": head["global_batch"], "prefetch_group": head["prefetch_group"], "exchange": "board",
            "parity_ok": ok, "parity": detail, "vs_mode_r": head["value"] / mode_r_value if mode_r_value else None,
            "cases": cases,
            "note": "global stratified draw over all shards (global CDF + global p_min), data-parallel learners: per "
                    "iteration one roots exchange by NVLink peer stores (roots_exchange_kernel), G global draws "
                    "(shares packed), gather || write-back of this rank's share; the reference polls shards "
                    "round-robin without normalisation (trainer.py:33-36)"}
