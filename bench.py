#!/usr/bin/env python
"""Benchmark of the prioritized-experience-replay hot path (BASELINE.json metric):
sampled transitions/s over  sample B -> gather B rows -> update B priorities.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

N=1 workload = BASELINE.json configs[1]: Atari 84x84x4 uint8 transitions (56 464 B rows),
PER capacity 2^20 (59.2 GB of frames resident in HBM), batch 512, alpha 0.6, beta 0.4.
One bench "step" = ITERS consecutive learner iterations (sample+gather+update of one batch
of 512 each); `value` = sampled transitions / s over the K timed steps.

Keys (see DESIGN.md "Measurement"): value = device-resident loop (reth_replay_step_loop, CUDA
events on the launching stream, max over ranks); e2e = the public Python API with HOST index /
priority buffers crossing PCIe every iteration; roofline = the gather kernel against the
measured HBM copy peak; cpu_baseline = the oracle (a C port of the reference's njit tree + numpy
ring) on the box's host cores.  --impl reference prints the CPU arm as its own line.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
ROW_BYTES = 28224 + 8 + 4 + 28224 + 4          # 56 464
ALPHA, BETA = 0.6, 0.4


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# stdout carries exactly ONE JSON line: everything else that libraries print there (NCCL prints its
# version banner on fd 1) is sent to stderr by swapping the descriptors for the whole run.
_REAL_STDOUT = None


def _claim_stdout():
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------ clocks --
class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 50 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "50", "-i", str(self.gpu)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        time.sleep(0.25)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(",") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        sm, reasons = [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            r = [x.strip() for x in r]
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                out["sm_max_mhz"] = float(r[2])
            except ValueError:
                continue
            for n, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if sm:
            out["sm_mhz"] = float(np.median(sm))
        out["reasons"] = sorted(reasons)
        out["samples"] = len(sm)
        return out


# ------------------------------------------------------------- CPU baseline --
def cpu_reference_arm(steps, warmup, iters, batch, cap, threads, ring_rows=1 << 14):
    """The reference's CPU path (oracle port: C restatement of the njit tree + memcpy row gather,
    SURVEY.md 8d Tier A) as `threads` independent replicas, one per host core.  Bounded sample:
    the tree has the full capacity, the frame ring holds `ring_rows` rows (slot & mask) so the
    host allocation stays below 1 GB; rows are still gathered from random slots."""
    from oracle import per_oracle as po
    rng = np.random.default_rng(0)
    ring = po.OracleRing(ring_rows, ATARI)
    for c in (0, 3):
        ring.column(c)[:] = rng.integers(0, 256, ring.column(c).shape, dtype=np.uint8)
    prio0 = po.normalize_weights((np.abs(rng.standard_normal(cap)) * (1 + 9 * (rng.random(cap) < 0.05))
                                  ).astype(np.float32), ALPHA)
    trees = []
    for _ in range(threads):
        t = po.OracleSumTree(cap)
        t.update(np.arange(cap), prio0)
        trees.append(t)
    total_steps = steps + warmup
    u = np.random.default_rng(1).random(total_steps * iters * batch)
    newp = np.abs(np.random.default_rng(2).standard_normal(total_steps * iters * batch)).astype(np.float32)

    def run(nsteps, offset):
        def work(t):
            n = nsteps * iters
            po.step_loop(t, ring, ring_rows - 1, n, batch, u[offset:], newp[offset:], ALPHA, BETA)
        th = [threading.Thread(target=work, args=(t,)) for t in trees]
        t0 = time.perf_counter()
        [x.start() for x in th]
        [x.join() for x in th]
        return time.perf_counter() - t0

    if warmup:
        run(warmup, 0)
    dt = run(steps, warmup * iters * batch)
    return threads * steps * iters * batch / dt, dt


# ---- the reference's OWN code (SURVEY.md 8d Tier A), from oracle/_ref (oracle/make_ref.py places it there at build time)
def ref_path():
    p = os.path.join(ROOT, "oracle", "_ref")
    return p if os.path.exists(os.path.join(p, "reth_ref", "utils", "sumtree.py")) else None


def _ref_numba_worker(a):
    """One replica = one process (the njit tree holds the GIL; the reference scales by processes too):
    FIFOPolicy.get_indices + NumpyBuffer.append_batch + PERSampler.update for the fill, then per iteration
    PERSampler.sample -> NumpyBuffer.select -> PERSampler.update, all the reference's own code."""
    path, cap, batch, warm, iters, ring_rows, seed, barrier = a
    sys.path.insert(0, path)
    from reth_ref.buffer import NumpyBuffer
    from reth_ref.cache_policy import FIFOPolicy
    from reth_ref.sampler import PERSampler
    rng = np.random.default_rng(seed)
    sampler, policy = PERSampler(cap, ALPHA, BETA), FIFOPolicy(cap)
    buf = NumpyBuffer(ring_rows, struct=[(np.dtype(dt).name, sh) for dt, sh in ATARI])
    blk = [rng.integers(0, 256, (64, 4, 84, 84), dtype=np.uint8), rng.integers(0, 18, 64), np.zeros(64, np.float32),
           rng.integers(0, 256, (64, 4, 84, 84), dtype=np.uint8), np.zeros(64, np.float32)]
    for _ in range(0, ring_rows, 64):                     # rows arrive 64 at a time (test/apex-dqn/config.yaml:3)
        buf.append_batch(blk)
    for st in range(0, cap, 1 << 14):
        n = min(1 << 14, cap - st)
        w = (np.abs(rng.standard_normal(n)) * (1 + 9 * (rng.random(n) < 0.05))).astype(np.float32)
        sampler.update(policy.get_indices(n), w)
    newp = np.abs(rng.standard_normal((64, batch))).astype(np.float32)
    mask = ring_rows - 1

    def loop(n):
        rows = None
        for i in range(n):
            idx, w = sampler.sample(batch)
            rows = buf.select(idx & mask)                 # bounded frame ring: slot & mask (the tree has full capacity)
            sampler.update(idx, newp[i & 63])
        return rows
    loop(warm)                                            # absorbs the numba JIT
    barrier.wait()
    t0 = time.perf_counter()
    loop(iters)
    return time.perf_counter() - t0


def reference_numba_arm(iters, batch, cap, procs, ring_rows=1 << 12, warm=3):
    """-> (transitions/s over all replicas, slowest replica's seconds) or None when numba / oracle/_ref is missing."""
    path = ref_path()
    if path is None:
        return None
    try:
        import numba  # noqa: F401
    except Exception:
        return None
    import multiprocessing as mp
    ctx = mp.get_context("spawn")                          # the parent may hold a CUDA context
    with ctx.Manager() as mgr:
        barrier = mgr.Barrier(procs)
        with ctx.Pool(procs) as pool:
            dts = pool.map(_ref_numba_worker, [(path, cap, batch, warm, iters, ring_rows, 100 + k, barrier) for k in range(procs)])
    return procs * iters * batch / max(dts), max(dts)


def service_tier_b():
    """BASELINE.md 3 Tier B: the reference's real multi-process service needs lmdb, portpicker, netifaces."""
    missing = []
    for m in ("lmdb", "portpicker", "netifaces", "lz4"):
        try:
            __import__(m)
        except Exception:
            missing.append(m)
    return ("service end-to-end: unavailable on this box (missing " + ", ".join(missing) + ")") if missing else "importable"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return                                        # rank 0 alone runs the CPU arm
    import __graft_entry__ as entry
    from oracle import per_oracle as po
    po.build()
    del entry
    cores = os.cpu_count() or 1
    iters = args.cpu_iters
    pvalue, pdt = cpu_reference_arm(args.steps, args.warmup, iters, args.batch, args.capacity, cores)
    psample = (f"{cores} replicas x {args.steps} steps x {iters} iterations of batch {args.batch}; tree capacity "
               f"{args.capacity}, frame ring bounded to 16384 rows (slot & mask)")
    port = {"value": pvalue, "unit": "transitions/s", "cores": cores, "kind": "port", "sample": psample}
    ref = None
    try:
        ref = reference_numba_arm(args.steps * iters, args.batch, args.capacity, cores, warm=max(3, args.warmup))
    except Exception as e:
        log(f"reference numba arm failed: {e!r}")
    if ref is not None:
        # the reference's own njit tree + NumPy ring (oracle/_ref) is the baseline of record; the C port beside it
        value, dt = ref
        dt = dt / max(1, args.steps)
        base = {"value": value, "unit": "transitions/s", "cores": cores, "kind": "reference",
                "sample": f"{cores} processes x {args.steps * iters} iterations of batch {args.batch}: the reference's own "
                          f"PERSampler (njit sum tree) + NumpyBuffer.select + PERSampler.update; tree capacity "
                          f"{args.capacity}, frame ring bounded to 4096 rows per process (slot & mask)",
                "port": port}
    else:
        value, dt = pvalue, pdt / max(1, args.steps)
        base = dict(port, note="numba or oracle/_ref unavailable on this box: C port of the reference's tree + memcpy ring")
    base["service_e2e"] = service_tier_b()
    line = {
        "impl": "reference", "metric": "sampled transitions/sec (sample+gather+priority update)",
        "value": value, "unit": "transitions/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8 rows / f64 tree", "data": "synthetic",
        "config": workload_config(args), "iterations_per_step": iters,
        "cpu_baseline": base,
        "e2e": {"value": value, "unit": "transitions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(args):
    return {"workload": f"Atari 84x84x4 uint8 transitions ({ROW_BYTES} B rows), PER capacity {args.capacity} per GPU, "
                        f"batch {args.batch}, alpha {ALPHA} beta {BETA} (BASELINE.json configs[1])",
            "capacity_per_gpu": args.capacity, "batch": args.batch, "row_bytes": ROW_BYTES,
            "prefetch_group": args.group,
            "l2": "inputs larger than L2: rows are gathered from random slots of a 59 GB ring"}


# ----------------------------------------------------------------- our arm --
def fill_shard(torch, shard, cap, seed):
    """Synthetic Atari-shaped rows generated on the device (a 59 GB host array is avoided)."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    chunk = 1 << 14
    for st in range(0, cap, chunk):
        n = min(chunk, cap - st)
        s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
        s1 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
        a = torch.randint(0, 18, (n,), device="cuda", generator=g)
        r = torch.randint(-1, 2, (n,), device="cuda", generator=g).float()
        d = (torch.rand(n, device="cuda", generator=g) < 0.01).float()
        w = torch.randn(n, device="cuda", generator=g).abs() * (1 + 9 * (torch.rand(n, device="cuda", generator=g) < 0.05))
        shard.append([s0, a, r, s1, d], w.float(), alpha=ALPHA)
    torch.cuda.synchronize()


def timed_loop(torch, dist, shard, steps, warmup, iters, batch, prio, group, world, time_gather, shard_group=None):
    def one_step(k):
        off = (k % prio_steps) * iters * batch
        shard.step_loop(iters, batch, prio[off:], seed=1234, alpha=ALPHA, beta=BETA, group=group,
                        time_gather=time_gather, one_gather_stream=time_gather)
        if shard_group is not None:
            # the path's only exchange: NCCL all-gather of the 16-byte shard roots {total, min}
            # (sharded.py Mode G), once per bench step, asynchronous on the torch stream
            shard_group.exchange_roots()
    prio_steps = prio.numel() // (iters * batch)
    for k in range(warmup):
        one_step(k)
    torch.cuda.synchronize()
    shard.gather_time()
    l0 = shard.stats()["kernel_launches"]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(steps):
        one_step(warmup + k)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        dist.barrier()
    launches = shard.stats()["kernel_launches"] - l0
    return ms, launches, shard.gather_time()


def e2e_loop(torch, dist, service, addr, iters_total, warm, batch, world, group, ahead, reference_call=None):
    """Public reth_buffer-style API with HOST buffers, per learner iteration:
        data, indices, weights = loader.sample()       # TorchCudaLoader: batch lent as CUDA tensors,
                                                       # indices = host int64 array (D2H inside)
        client.update_priorities(indices, loss)        # host int64 indices + host float32 |TD| -> H2D
    group/ahead = the loader's prefetch (the reference's loader ring + sampler queue);
    group 1 / ahead 0 = strict alternation."""
    rng = np.random.default_rng(3)
    newp = np.abs(rng.standard_normal((64, batch))).astype(np.float32)
    client = service.Client(addr)
    if reference_call:      # exactly as the reference's Ape-X trainer builds it (test/apex-dqn/trainer.py:22-24)
        loader = service.TorchCudaLoader(addr, **reference_call)
    else:
        loader = service.TorchCudaLoader(addr, buffer_size=(ahead + 1) * group, prefetch=group, ahead=ahead)
    eng = service._ENGINES[addr]

    def it(i):
        data, indices, weights = loader.sample()
        client.update_priorities(indices, newp[i % 64])
    for i in range(warm):
        it(i)
    eng.drain()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for i in range(iters_total):
        it(warm + i)
    eng.drain()
    e1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) * 1e3
    ms = max(e0.elapsed_time(e1), wall)                 # host-driven loop: the wall clock is the honest one
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    geom = (loader.prefetch, loader.ahead)
    loader.close()
    if reference_call:
        return world * iters_total * batch / (ms / 1e3), geom
    return world * iters_total * batch / (ms / 1e3), batch * (8 + 4), batch * (8 + 8)


def append_rate(torch, shard, n=16384, reps=10):
    cols = [torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda"),
            torch.zeros(n, dtype=torch.int64, device="cuda"), torch.zeros(n, device="cuda"),
            torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda"), torch.zeros(n, device="cuda")]
    w = torch.rand(n, device="cuda")
    for _ in range(3):
        shard.append(cols, w, alpha=ALPHA)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        shard.append(cols, w, alpha=ALPHA)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    return n / (ms / 1e3), 2 * n * ROW_BYTES / (ms / 1e3) / 1e9


def append_rate_host(torch, shard, n=4096, reps=5):
    """Host-producer variant (config 4): the same append from PINNED host columns, i.e. what an actor
    message costs once decoded -- PCIe-bound, reported beside the HBM-bound device-source figure."""
    cols = [torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8).pin_memory(),
            torch.zeros(n, dtype=torch.int64).pin_memory(), torch.zeros(n).pin_memory(),
            torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8).pin_memory(), torch.zeros(n).pin_memory()]
    w = torch.rand(n).pin_memory()
    shard.append(cols, w, alpha=ALPHA)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        shard.append(cols, w, alpha=ALPHA)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    return n / (ms / 1e3), n * ROW_BYTES / (ms / 1e3) / 1e9


def config4_loop(torch, shard, batch, group, rows_per_step, groups):
    """BASELINE.json config 4, device-resident producer: per group of `group` learner steps the pending actor rows
    (rows_per_step per step) are appended, then `group` batches are drawn, gathered and written back.
    alpha 0.5, beta 0.4 (test/apex-dqn/config.yaml:26-27).  -> (sampled transitions/s, appended rows/s, HBM GB/s)."""
    n = rows_per_step * group
    g = torch.Generator(device="cuda").manual_seed(5)
    s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
    z = torch.zeros(n, device="cuda")
    cols, w = [s0, z.long(), z, s0, z], torch.rand(n, device="cuda", generator=g)
    prio = torch.rand(group * batch, device="cuda", generator=g) + 0.01

    chunk = 8                                     # groups per call of the device loop (prio covers one group, reused)
    prio = prio.repeat(chunk)

    def loop(k):
        for _ in range(0, k, chunk):
            # reth_replay_step_loop_ingest: after every group the actor rows are appended without draining the loop
            shard.step_loop(group * chunk, batch, prio, alpha=0.5, beta=0.4, group=group, ingest=(cols, w))
    groups = max(chunk, groups // chunk * chunk)
    loop(chunk)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loop(groups)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (groups * group)
    return batch / us * 1e6, rows_per_step / us * 1e6, 2 * ROW_BYTES * (batch + rows_per_step) / us / 1e3


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the replay engine has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import __graft_entry__ as entry
    if rank == 0:
        entry.build()
    if world > 1:
        dist.barrier()
    from reth_b200 import service

    B, cap, iters, G = args.batch, args.capacity, args.iters, args.group
    if iters % G:
        iters += G - iters % G
    t0 = time.time()
    # the shard is owned by a reth_buffer-style server (start_per) so that the device loop and the
    # end-to-end loader/client loop run on the very same buffer
    proc, addr = service.start_per(cap, B, alpha=ALPHA, beta=BETA, sample_start=B, device=local,
                                   max_ring=max(16, (args.ahead + 1) * G))
    eng = service._ENGINES[addr]
    eng._ensure_shard(None, struct=ATARI)
    shard = eng.shard
    fill_shard(torch, shard, cap, seed=rank)
    eng.cnt = cap
    log(f"[rank {rank}] filled {cap} rows ({shard.stats()['hbm_bytes'] / 1e9:.1f} GB) in {time.time() - t0:.1f}s")
    pool_steps = min(args.steps + args.warmup, 64)
    g = torch.Generator(device="cuda").manual_seed(100 + rank)
    prio = torch.randn(pool_steps * iters * B, device="cuda", generator=g).abs().float()

    # multi-GPU: shards are independent (reference: one server per shard, trainer.py:52-61); the only
    # exchange is the 16-byte shard root {total, min}, combined with an NCCL all-gather per bench step
    sg = None
    if world > 1:
        from reth_b200.sharded import ShardGroup
        sg = ShardGroup(shard)

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms, launches, _ = timed_loop(torch, dist, shard, args.steps, args.warmup, iters, B, prio, G, world, False, sg)
    clk = clocks.stop() if rank == 0 else None
    value = world * args.steps * iters * B / (ms / 1e3)

    # roofline pass: the same loop with every gather launch bracketed by CUDA events on ONE gather stream
    # (in the headline loop consecutive gathers alternate between two streams and overlap, so a per-launch
    # bracket there would not be a kernel duration); the tree kernels still run concurrently
    _, _, (gt_ms, gt_n) = timed_loop(torch, dist, shard, max(2, args.steps // 2), 1, iters, B, prio, G, world, True)
    # strict alternation (no prefetch): sample k+1 sees update k
    ms1, _, _ = timed_loop(torch, dist, shard, max(1, args.steps // 2), 1, iters, B, prio, 1, world, False)
    strict = world * max(1, args.steps // 2) * iters * B / (ms1 / 1e3)
    # end to end through the public API with host buffers
    e2e_iters = min(20480, max(400, args.steps * iters))
    e2e_iters -= e2e_iters % G
    e2e, h2d, d2h = e2e_loop(torch, dist, service, addr, e2e_iters, 64, B, world, G, args.ahead)
    e2e_strict, _, _ = e2e_loop(torch, dist, service, addr, min(4000, e2e_iters), 16, B, world, 1, 0)
    ref_call = {"buffer_size": 4, "num_procs": 2}
    e2e_ref, ref_geom = e2e_loop(torch, dist, service, addr, min(8192, e2e_iters), 64, B, world, 0, 0, reference_call=ref_call)
    app_rows, app_gbs = append_rate(torch, shard) if rank == 0 or world == 1 else (None, None)
    apph_rows, apph_gbs = append_rate_host(torch, shard) if rank == 0 and world == 1 else (None, None)
    config4 = None
    if world == 1:
        c4a = config4_loop(torch, shard, B, G, 64, 150)
        c4s = config4_loop(torch, shard, B, G, 256 * 64, 30)
        config4 = {"apex_like": {"value": c4a[0], "appended_rows_per_s": c4a[1], "hbm_gbs": c4a[2], "rows_appended_per_learner_step": 64},
                   "stress": {"value": c4s[0], "appended_rows_per_s": c4s[1], "hbm_gbs": c4s[2], "rows_appended_per_learner_step": 256 * 64},
                   "unit": "sampled transitions/s", "prefetch_group": G,
                   "note": "learner loop with actor ingest on one shard, device-resident producer (reth_replay_step_loop_ingest): "
                           "per group G x (sample 512 + gather + update), then the rows that arrived are appended -- their leaves in "
                           "the write-back's repair launch, the row copy behind the gathers in flight, later gathers behind the copy; alpha 0.5 "
                           "beta 0.4 (config.yaml:26-27); 256 actor connections over ZeroMQ: see ingest_256"}

    peak, peak_src = measured_peak()
    rows_per_launch = G * B
    gather_us = gt_ms * 1e3 / max(1, gt_n)
    achieved = 2 * rows_per_launch * ROW_BYTES / (gather_us * 1e-6) / 1e9 if gt_n else None
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "gather_traffic.json")
    if os.path.exists(tpath):
        try:
            tj = json.load(open(tpath))
            # only a capture of THIS configuration counts (capacity, rows per launch); anything else stays null
            if int(tj.get("rows_per_launch", -1)) == rows_per_launch and int(tj.get("capacity", -1)) == cap:
                traffic = tj["dram_bytes_per_launch"]
        except Exception:
            pass
    line = {
        "metric": "sampled transitions/sec (sample+gather+priority update)", "value": value, "unit": "transitions/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8 rows / f64 tree",
        "data": "synthetic", "config": workload_config(args), "iterations_per_step": iters,
        "clocks": clk, "gpu_launches": int(launches),
        "e2e": {"value": e2e, "unit": "transitions/s", "h2d_bytes_per_step": h2d * iters, "d2h_bytes_per_step": d2h * iters,
                "strict_alternation": e2e_strict,
                "e2e_ref_geometry": {"value": e2e_ref, "call": "TorchCudaLoader(addr, buffer_size=4, num_procs=2)",
                                     "prefetch_group": ref_geom[0], "groups_ahead": ref_geom[1],
                                     "batches_in_flight": ref_geom[0] * (ref_geom[1] + 1),
                                     "note": "the loader exactly as test/apex-dqn/trainer.py:22-24 builds it: the reference keeps "
                                             "buffer_size + num_procs + 2 = 8 batches between draw and consumption"},
                "loader": {"prefetch_group": G, "groups_ahead": args.ahead, "buffer_size": (args.ahead + 1) * G},
                "note": "reth_buffer-style public API per iteration: TorchCudaLoader.sample() (batch lent as CUDA "
                        "tensors; indices + weights D2H into pinned host rings, indices returned as a host int64 "
                        "array) + Client.update_priorities(host int64 indices, host float32 |TD|) -> pinned "
                        "staging -> H2D -> tree update; strict_alternation = prefetch 1 / 0 ahead"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": (achieved / peak) if achieved else None, "traffic": traffic, "peak_source": peak_src,
                     "kernel": "row_move_kernel<false> (frame gather)", "rows_per_launch": rows_per_launch,
                     "algorithmic_bytes_per_launch": 2 * rows_per_launch * ROW_BYTES, "avg_launch_us": gather_us,
                     "launches_timed": int(gt_n),
                     "how": "second pass of the same loop with all gathers on ONE stream (tree kernels concurrent): per "
                            "bench step one CUDA-event bracket on the gather stream from the first to the last gather "
                            "launch, divided by the launches it spans (includes the gaps between launches); the "
                            "headline loop alternates two gather streams, see hbm_gbs_whole_loop for its aggregate"},
        "hbm_gbs_whole_loop": 2 * ROW_BYTES * value / 1e9 / world,
        "strict_alternation": {"value": strict, "unit": "transitions/s", "prefetch_group": 1},
        "append": {"rows_per_s": app_rows, "gbs": app_gbs, "rows_per_call": 16384, "source": "device"},
        "append_host": {"rows_per_s": apph_rows, "pcie_gbs": apph_gbs, "rows_per_call": 4096, "source": "pinned host"},
        "config4": config4,
        "shard_mode": "one shard per GPU, no data-path collective; NCCL all-gather of the 16-byte shard roots {total,min} once per bench step inside the timed region" if world > 1 else "single shard",
    }
    if world > 1:
        line["shard_totals"] = [float(x) for x in sg.all_roots.cpu()[0::2]]
        # Mode G (BASELINE.json config 3, north_star's coupled multi-GPU path): parity leg + timed device loop on
        # the same shards, after every Mode R figure above has been taken (the batch ring is re-shaped)
        import bench_modeg
        try:
            line["mode_g"] = bench_modeg.run(torch, dist, shard, world, rank, B, ALPHA, BETA, value, log,
                                             iters=args.modeg_iters)
        except Exception as e:                                  # keep the Mode R line if the coupled leg fails
            log(f"[rank {rank}] mode_g failed: {e!r}")
            line["mode_g"] = {"error": repr(e), "parity_ok": False}
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        v, dt = cpu_reference_arm(args.cpu_steps, 1, args.cpu_iters, B, cap, cores)
        v1, _ = cpu_reference_arm(max(1, args.cpu_steps // 6), 1, args.cpu_iters, B, cap, 1)
        line["cpu_baseline"] = {
            "value": v, "unit": "transitions/s", "cores": cores, "kind": "port", "one_core": v1,
            "sample": f"{cores} replicas x {args.cpu_steps} steps x {args.cpu_iters} iterations of batch {B} "
                      f"({dt:.1f} s); tree capacity {cap}, frame ring bounded to 16384 rows (slot & mask)"}
        try:                                            # the reference's own code on the same cores (oracle/_ref)
            rn = reference_numba_arm(max(64, args.cpu_steps * args.cpu_iters // 4), B, cap, cores)
            rn1 = reference_numba_arm(max(64, args.cpu_steps * args.cpu_iters // 4), B, cap, 1, ring_rows=1 << 16)
        except Exception as e:
            log(f"reference numba arm failed: {e!r}")
            rn = rn1 = None
        line["cpu_baseline"]["reference_numba"] = None if rn is None else {
            "value": rn[0], "unit": "transitions/s", "cores": cores, "one_core_ring_65536": rn1[0] if rn1 else None,
            "sample": f"{cores} processes: reference PERSampler (njit tree, capacity {cap}) + NumpyBuffer.select "
                      f"(ring bounded to 4096 rows per process; one-core figure with a 65536-row = 3.7 GB ring) + "
                      f"PERSampler.update, {rn[1]:.1f} s"}
        line["cpu_baseline"]["service_e2e"] = service_tier_b()
    proc.terminate()
    if rank == 0 and world == 1 and not args.no_process:
        import bench_process
        line["e2e_process"] = bench_process.run_leg("e2e_process", 180, batch=B)
        line["ingest_256"] = bench_process.run_leg("ingest_256", 240)
        line["apex_process"] = bench_process.run_leg("apex_process", 150)      # both at once: config 4 over the wire
        for k in ("e2e_process", "ingest_256", "apex_process"):
            if "error" in line[k]:
                log(f"{k} failed: {line[k]['error']}")
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--capacity", type=int, default=1 << 20)
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--iters", type=int, default=2048,
                    help="learner iterations per bench step (2048 x batch 512 = 1 M sampled transitions, about 20 ms on one "
                         "B200: the default 10 timed steps then span several clock samples)")
    ap.add_argument("--group", type=int, default=8,
                    help="prefetch group: batches drawn per launch (one draw kernel, one gather launch, one write-back)")
    ap.add_argument("--ahead", type=int, default=2,
                    help="e2e loader: groups drawn ahead of the one being consumed; the lending ring holds "
                         "(ahead + 1) * group batches.  The reference keeps buffer_size + num_procs + 2 batches in "
                         "flight (TorchCudaLoader ring + one per worker + sampler side: 16 at its defaults, "
                         "torch_cuda_loader.py:70, sampler_loop.py:14); 8 x (2 + 1) = 24 here, --ahead 1 = 16")
    ap.add_argument("--cpu-steps", type=int, default=24)
    ap.add_argument("--cpu-iters", type=int, default=64, help="iterations per step of the CPU arm (bounded sample)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-process", action="store_true", help="skip the server-process legs (e2e_process, ingest_256)")
    ap.add_argument("--modeg-iters", type=int, default=400, help="Mode G leg (N > 1): timed iterations per case")
    args = ap.parse_args()
    _claim_stdout()
    if args.warmup < 3 and args.impl == "ours":
        log(f"--warmup {args.warmup} raised to 3 (timing rule: at least three untimed warm-up steps)")
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
