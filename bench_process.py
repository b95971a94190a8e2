"""Process-layout legs of `bench.py` (N = 1, rank 0): the path the `reth_buffer` import shim defaults to.

The reference runs its replay as a server process reached by address (test/apex-dqn/trainer.py:49-66); the
shim's default (`RETH_B200_SERVER=process`, reth_b200/service_mp.py) does the same: ZeroMQ for append / update /
loader requests in the reference wire format, CUDA IPC for the batch ring.  Measured here, beside the in-process
engine the headline `e2e` uses:

  e2e_process   TorchCudaLoader(addr).sample() + Client.update_priorities() against a spawned server, loader
                built with the DEFAULT arguments a caller of the reference API passes; capacity bounded to
                2^16 Atari rows (3.7 GB: the rows travel to the server to get in)
  ingest        rows/s taken in by the server: same-host shared-memory ring (this build's Client default), and
                BASELINE.json config 4 as worded -- 256 concurrent PUSH connections x 64-row messages in the
                reference's ROW wire format (client/client.py:26-35), plain and lz4-framed (worker.py:60)
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))


def e2e_process(batch=512, capacity=1 << 16, iters=4000, warm=128):
    from reth_b200 import service_mp
    proc, addr = service_mp.start_per(capacity, batch, alpha=0.6, beta=0.4, sample_start=batch)
    try:
        client = service_mp.Client(addr)                      # wire "auto": shared-memory ring on the server's host
        rng = np.random.default_rng(0)
        n = 2048
        s0 = rng.integers(0, 256, (n, 4, 84, 84), dtype=np.uint8)
        cols = [s0, rng.integers(0, 18, n), np.zeros(n, np.float32), s0, np.zeros(n, np.float32)]
        wts = np.abs(rng.standard_normal(n)).astype(np.float32)
        loader = service_mp.TorchCudaLoader(addr)             # the reference's defaults: buffer_size=8, num_procs=6
        client.append(cols, wts)
        loader.drain()
        t0 = time.perf_counter()
        for _ in range(n, capacity, n):
            client.append(cols, wts)
        loader.drain()
        t_fill = time.perf_counter() - t0
        newp = np.abs(rng.standard_normal((64, batch))).astype(np.float32)
        for i in range(warm):
            data, idx, w = loader.sample()
            client.update_priorities(idx, newp[i % 64])
        loader.drain()
        t0 = time.perf_counter()
        for i in range(iters):
            data, idx, w = loader.sample()
            client.update_priorities(idx, newp[i % 64])
        loader.drain()
        dt = time.perf_counter() - t0
        out = {"value": iters * batch / dt, "unit": "transitions/s", "us_per_iteration": dt / iters * 1e6,
               "call": "reth_buffer.start_per(...) in its own process; TorchCudaLoader(addr) [defaults]; Client(addr)",
               "prefetch_group": loader.group, "groups_ahead": loader.ahead, "capacity": capacity,
               "ingest_shm_rows_per_s": (capacity - n) / t_fill, "ingest_shm_GBps": (capacity - n) * 56464 / t_fill / 1e9,
               "h2d_bytes_per_iteration": batch * 12, "d2h_bytes_per_iteration": batch * 16}
        loader.close()
        client.close()
        return out
    finally:
        proc.terminate()
        proc.join()


def ingest_256(log=None):
    """256 connections x 64-row messages, reference row format (scripts/ingest_rate.py)."""
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import ingest_rate
    out = {}
    for name, comp, msgs in (("rows", False, 6), ("rows_lz4", True, 6)):
        try:
            r = ingest_rate.run(256, msgs, compress=comp, wire="rows")
            out[name] = {k: r[k] for k in ("rows_per_s", "wire_GBps", "messages", "message_bytes", "dropped")}
        except Exception as e:                                  # keep the bench line
            out[name] = {"error": repr(e)}
            if log:
                log(f"ingest_256 {name} failed: {e!r}")
    out["connections"], out["rows_per_message"] = 256, 64
    out["ingest_threads"] = int(os.environ.get("RETH_B200_INGEST_THREADS", 0)) or max(2, min(8, (os.cpu_count() or 4) // 2))
    out["note"] = ("256 concurrent ZeroMQ PUSH connections (one per actor) x 64-row messages in the reference's row wire "
                   "format into the server process: ZeroMQ I/O threads -> decode threads (header, lz4, row->column "
                   "transpose in one library call) -> pinned staging -> asynchronous H2D -> ring append + tree update; "
                   "messages are encoded once and re-sent (the per-row Python encode is the actor's cost)")
    return out


def _apex_sender(meta_addr, append_addr, n_conn, rows, pace_rows_per_s, stop, sent, window=512):
    """Actor side of apex_process, in a process of its own: n_conn PUSH connections, one pre-encoded `rows`-row message
    in the reference's row format re-sent round robin, paced to pace_rows_per_s (None: as fast as the server takes
    them).  At most `window` messages are outstanding (the server's row counter is read every 32 sends): 256
    connections x HWM 10 on both ends would otherwise park up to 19 GB of messages in host memory."""
    import zmq
    from reth_b200 import pack, service_mp
    ctl = service_mp.NumpyLoader(meta_addr)
    cnt0 = ctl.stat()["appended_rows"]
    taken = 0
    rng = np.random.default_rng(1)
    cols = [rng.integers(0, 256, (rows, 4, 84, 84), dtype=np.uint8), rng.integers(0, 18, rows),
            rng.standard_normal(rows).astype(np.float32), rng.integers(0, 256, (rows, 4, 84, 84), dtype=np.uint8),
            (rng.random(rows) < 0.01).astype(np.float32)]
    msg = bytes(pack.encode_append_rows(cols, np.abs(rng.standard_normal(rows)).astype(np.float32)))
    ctx = zmq.Context(4)
    socks = []
    for _ in range(n_conn):
        sk = ctx.socket(zmq.PUSH)
        sk.set_hwm(10)
        sk.setsockopt(zmq.LINGER, 0)
        sk.connect(append_addr)
        socks.append(sk)
    gap = rows / pace_rows_per_s if pace_rows_per_s else 0.0
    nxt, k = time.perf_counter(), 0
    while not stop.is_set():
        if sent.value - taken >= window or (k % 32 == 0 and sent.value - taken >= window // 2):
            taken = (ctl.stat()["appended_rows"] - cnt0) // rows
            if sent.value - taken >= window:
                time.sleep(0.0005)
                continue
        if gap:
            now = time.perf_counter()
            if now < nxt:
                time.sleep(min(nxt - now, 0.001))
                continue
            nxt = max(nxt + gap, now - 10 * gap)
        try:
            socks[k % n_conn].send(msg, zmq.NOBLOCK, copy=False)
            with sent.get_lock():
                sent.value += 1
        except zmq.Again:
            time.sleep(0.0002)
        k += 1
    for sk in socks:
        sk.close(0)
    ctx.term()
    ctl.close()


def apex_process(batch=512, capacity=1 << 16, seconds=1.5, n_conn=256, rows=64):
    """BASELINE.json config 4 over the wire: the learner loop of e2e_process (TorchCudaLoader defaults +
    Client.update_priorities) WHILE n_conn actor connections push 64-row messages in the reference's row format into
    the same server process (test/apex-dqn: trainer.py:22-36 and worker.py:57-61 at once).  Two actor rates: paced to
    50 k rows/s in total (256 actors x ~200 env steps/s) and unpaced (the server's ingest limit)."""
    import multiprocessing as mp
    from reth_b200 import service_mp
    proc, addr = service_mp.start_per(capacity, batch, alpha=0.6, beta=0.4, sample_start=batch)
    ctx = mp.get_context("spawn")
    try:
        meta = service_mp.get_meta(addr)
        client = service_mp.Client(addr)
        rng = np.random.default_rng(0)
        n = 2048
        s0 = rng.integers(0, 256, (n, 4, 84, 84), dtype=np.uint8)
        cols = [s0, rng.integers(0, 18, n), np.zeros(n, np.float32), s0, np.zeros(n, np.float32)]
        wts = np.abs(rng.standard_normal(n)).astype(np.float32)
        loader = service_mp.TorchCudaLoader(addr)
        for _ in range(0, capacity, n):
            client.append(cols, wts)
        loader.drain()
        newp = np.abs(rng.standard_normal((64, batch))).astype(np.float32)
        out = {"connections": n_conn, "rows_per_message": rows, "capacity": capacity,
               "call": "start_per(...) in its own process; learner: TorchCudaLoader(addr) [defaults] + Client(addr).update_priorities; "
                       "actors: PUSH connections in a second process, reference row format"}
        for name, pace in (("paced_50k_rows_per_s", 50_000), ("unpaced", None)):
            stop, sent = ctx.Event(), ctx.Value("q", 0)
            snd = ctx.Process(target=_apex_sender, args=(addr, meta["append_addr"], n_conn, rows, pace, stop, sent), daemon=True)
            snd.start()
            t_wait = time.time() + 60
            while sent.value < n_conn and time.time() < t_wait:      # every connection is up, slabs pinned
                time.sleep(0.01)
            for i in range(256):
                data, idx, w = loader.sample()
                client.update_priorities(idx, newp[i % 64])
            c0 = loader.stat()
            t0 = time.perf_counter()
            it = 0
            while time.perf_counter() - t0 < seconds:
                for _ in range(64):
                    data, idx, w = loader.sample()
                    client.update_priorities(idx, newp[it % 64])
                    it += 1
            c1 = loader.stat()
            dt = time.perf_counter() - t0
            stop.set()
            snd.join(20)
            if snd.is_alive():
                snd.kill()
            loader.drain()
            out[name] = {"learner_transitions_per_s": it * batch / dt,
                         "ingest_rows_per_s": (c1["appended_rows"] - c0["appended_rows"]) / dt,
                         "us_per_iteration": dt / max(1, it) * 1e6, "decode_backlog": c1.get("backlog"), "dropped": c1.get("dropped")}
        loader.close()
        client.close()
        return out
    finally:
        proc.terminate()
        proc.join()


# ---- isolation: each leg runs in a child process (own session) under a deadline, so that a hung server or a lost
# message can cost the bench line one key, never the line
def _leg_main(name, kwargs, q):
    os.setsid()
    sys.path.insert(0, ROOT)
    try:
        q.put(("ok", globals()[name](**kwargs)))
    except BaseException as e:                                  # noqa: BLE001 -- reported to the parent
        q.put(("error", repr(e)))


def run_leg(name, timeout_s=180, **kwargs):
    """-> the leg's dict, or {"error": ...}.  name: "e2e_process" | "ingest_256" | "apex_process"."""
    import multiprocessing as mp
    import signal
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    p = ctx.Process(target=_leg_main, args=(name, kwargs, q))
    p.start()
    try:
        kind, val = q.get(timeout=timeout_s)
    except Exception:
        kind, val = "error", f"no result within {timeout_s} s"
    p.join(10)
    if p.is_alive() or kind == "error":
        try:
            os.killpg(p.pid, signal.SIGKILL)                    # the leg and whatever it spawned
        except Exception:
            pass
        p.join(5)
    return val if kind == "ok" else {"error": val}
