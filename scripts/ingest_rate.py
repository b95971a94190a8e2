"""Ingest at Ape-X scale (BASELINE.json config 4 as worded: "256 synthetic actor append streams"): N_CONN
concurrent ZeroMQ PUSH connections (one per actor, test/apex-dqn/worker.py:22,57-61) send 64-row messages
(rollout_batch_size, test/apex-dqn/config.yaml:3) in the reference's ROW wire format (client/client.py:26-35) to
the server process; the rate is rows taken in per second until the server has committed every message.
Messages are encoded once per sender thread and re-sent (the per-row Python encode is the ACTOR's cost,
29 k rows/s per producer in BASELINE.md 2, and is not what is measured).  One JSON line."""
import json, os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import zmq


def run(n_conn=256, msgs_per_conn=8, rows=64, compress=False, sender_threads=8, wire="rows", capacity=1 << 18, batch=512,
        io_threads=4):
    from reth_b200 import pack, service_mp
    proc, addr = service_mp.start_per(capacity, batch, sample_start=batch)
    try:
        meta = service_mp.get_meta(addr)
        rng = np.random.default_rng(0)
        cols = [rng.integers(0, 256, (rows, 4, 84, 84), dtype=np.uint8), rng.integers(0, 18, rows),
                rng.standard_normal(rows).astype(np.float32), rng.integers(0, 256, (rows, 4, 84, 84), dtype=np.uint8),
                (rng.random(rows) < 0.01).astype(np.float32)]
        if compress:                                   # Atari frames compress; uniform noise does not
            cols[0] = (cols[0] // 64) * 64
            cols[3] = cols[0]
        w = np.abs(rng.standard_normal(rows)).astype(np.float32)
        enc = pack.encode_append_rows if wire == "rows" else pack.encode_append_columns
        msg = bytes(enc(cols, w, compress=compress))
        ctx = zmq.Context(io_threads)
        socks = []
        for _ in range(n_conn):
            s = ctx.socket(zmq.PUSH)
            s.set_hwm(10)                              # utils/__init__.py:12
            s.connect(meta["append_addr"])
            socks.append(s)
        loader = service_mp.NumpyLoader(addr)
        # warm-up: one message per connection (connections established, row struct known, slabs pinned)
        for s in socks:
            s.send(msg, copy=False)
        total = n_conn * (msgs_per_conn + 1) * rows
        deadline = time.time() + 60
        while loader.drain()["cnt"] < n_conn * rows:
            if time.time() > deadline:
                raise TimeoutError("ingest warm-up: the server did not take in one message per connection within 60 s")
            time.sleep(0.01)

        def sender(k):
            mine = socks[k::sender_threads]
            for _ in range(msgs_per_conn):
                for s in mine:
                    s.send(msg, copy=False)
        th = [threading.Thread(target=sender, args=(k,)) for k in range(sender_threads)]
        t0 = time.perf_counter()
        [t.start() for t in th]
        [t.join() for t in th]
        t_sent = time.perf_counter() - t0
        deadline = time.time() + 120
        while True:
            st = loader.drain()
            if st["cnt"] >= total:
                break
            if time.time() > deadline or st.get("dropped", {}).get("append"):
                raise TimeoutError(f"ingest: {st['cnt']} of {total} rows taken in (dropped: {st.get('dropped')})")
            time.sleep(0.002)
        dt = time.perf_counter() - t0
        n = n_conn * msgs_per_conn * rows
        out = {"connections": n_conn, "rows_per_message": rows, "messages": n_conn * msgs_per_conn, "wire": wire,
               "compress": bool(compress), "message_bytes": len(msg), "rows_per_s": n / dt, "wire_GBps": n_conn * msgs_per_conn * len(msg) / dt / 1e9,
               "seconds": dt, "send_seconds": t_sent, "dropped": st.get("dropped"),
               "ingest_threads": int(os.environ.get("RETH_B200_INGEST_THREADS", 0)) or max(2, min(8, (os.cpu_count() or 4) // 2))}
        loader.close()
        for s in socks:
            s.close(0)
        ctx.term()
        return out
    finally:
        proc.terminate()
        proc.join()


if __name__ == "__main__":
    import __graft_entry__ as entry
    entry.build()
    for wire, comp in (("rows", False), ("rows", True), ("columns", False)):
        print(json.dumps(run(int(os.environ.get("N_CONN", 256)), int(os.environ.get("MSGS", 8)), compress=comp, wire=wire)), flush=True)
