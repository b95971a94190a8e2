"""The poll loop of the server process (reth_b200/service_mp.server_main) on the CPU: the real loop, sockets, decode
threads, Client and NumpyLoader, around a stand-in engine that records the order in which the loop hands it work.
What is pinned here is ORDER (the reference gets it from its sockets, server/sampler_loop.py:29-36,
server/main_loop.py:44-58):
  * a draw sees every write its own process sent before the request, however long the decode of an append takes;
  * a write-back waits for the appends ITS sender sent before it, and for nobody else's;
  * drain() waits for everything received."""
import signal
import threading
import time

import numpy as np
import pytest


class _Shard:
    out_struct = [(np.dtype(np.int64), ()), (np.dtype(np.float32), ())]
    ring_depth = 4
    _h = None

    def _ck(self, rc):
        pass

    rows = 0

    def stats(self):
        return {"appended_rows": self.rows}

    def slot_read(self, slot, with_rows=False, stream=None):
        return [np.zeros(4, np.int64), np.zeros(4, np.float32)], None, None


class _Engine:
    """The surface of service._Engine the loop uses; every call lands in `log`."""
    log = []

    def __init__(self, *a, **k):
        self.topic, self.topics = "default", {"default": {"kind": "per", "sample_start": 1}}
        self.shard = self.stream = self.pipe = None
        self.loaders, self.max_ring, self.group, self.ahead, self.lend, self.cnt, self._owed = 0, 8, 1, 0, 1, 0, 0

    def append(self, cols, w, stream_key=None):
        self.shard = self.shard or _Shard()
        self.shard.rows += len(w)
        self.cnt += len(w)
        _Engine.log.append(("append", int(cols[0][0]), len(w)))

    def update(self, idx, w, step):
        _Engine.log.append(("update", [int(i) for i in idx], self.shard is not None))

    def update_block(self, idx, w, sizes, steps):
        _Engine.log.append(("update", [int(i) for i in idx], self.shard is not None))

    def ready(self):
        return self.shard is not None

    def request_group(self):
        _Engine.log.append(("draw",))
        self._owed += 1

    def try_deliver(self):
        if not self._owed:
            return None
        self._owed -= 1
        return 0, np.zeros((1, 4), np.int64), np.zeros((1, 4))

    def open_pipe(self, group, ahead, lend):
        self.loaders += 1

    def release_pipe(self):
        self.loaders -= 1

    def drain(self):
        pass

    def close(self):
        pass


@pytest.fixture
def server(monkeypatch):
    """-> (address, set_decode_delay): server_main in a thread of this process, no CUDA anywhere."""
    import faulthandler
    import torch
    from reth_b200 import _cabi, service, service_mp as m
    real = _cabi.lib()

    class Lib:
        def __getattr__(self, k):
            return getattr(real, k)

        def reth_replay_ipc_export(self, h, ex):
            return 0
    delay = {"first_value": None, "seconds": 0.0}
    handlers = {}

    def slab(self, nbytes, seq):                              # stands in for the pinned slab (pin_memory needs CUDA)
        return torch.empty(max(1 << 16, nbytes), dtype=torch.uint8)
    real_open = m.pack.open_append

    def slow_open(buf, alloc=None):                           # a decode that takes long for the marked message
        res = real_open(buf, alloc)
        if delay["first_value"] is not None and res[0] == "columns" and int(res[1][0][0]) == delay["first_value"]:
            time.sleep(delay["seconds"])
        return res
    monkeypatch.setattr(service, "_Engine", _Engine)
    monkeypatch.setattr(m._Ingest, "_slab", slab)
    monkeypatch.setattr(m.pack, "open_append", slow_open)
    monkeypatch.setattr(_cabi, "lib", lambda: Lib())
    monkeypatch.setattr(signal, "signal", lambda sig, h: handlers.__setitem__(sig, h))    # not the main thread's business
    monkeypatch.setattr(faulthandler, "dump_traceback_later", lambda *a, **k: None)
    _Engine.log = []
    ports = {k: m._pick_port("127.0.0.1") for k in ("meta", "append", "update", "loader")}
    ready = threading.Event()
    th = threading.Thread(target=m.server_main, args=(4096, 4, "127.0.0.1", ports, None, None, 0, 8, ready), daemon=True)
    th.start()
    assert ready.wait(20)

    def set_delay(first_value, seconds):
        delay["first_value"], delay["seconds"] = first_value, seconds
    yield f"tcp://127.0.0.1:{ports['meta']}", set_delay
    handlers[signal.SIGTERM]()                                # what proc.terminate() delivers
    th.join(10)
    assert not th.is_alive()


def _cols(first, n=16):
    ids = first + np.arange(n, dtype=np.int64)
    return [ids, ids.astype(np.float32)], np.full(n, 1e-3, np.float32)


@pytest.mark.parametrize("wire", ["columns", "rows"])
def test_write_back_waits_for_a_slow_decode_of_its_own_append(server, wire):
    """The first decode of a server pins its staging slab, which starts CUDA: seconds.  The write-back sent right
    behind the append must still land after it (it was applied to an empty buffer -- and lost -- when the wait for
    the named append gave up after 2 s)."""
    from reth_b200 import service_mp as m
    addr, set_delay = server
    set_delay(1000, 2.6)
    client = m.Client(addr, wire=wire, coalesce=8 if wire == "columns" else 1)
    loader = m.NumpyLoader(addr)
    client.append(*_cols(1000))
    client.update_priorities(np.array([5]), np.array([1e9], np.float32))
    loader.sample()
    assert _Engine.log[:3] == [("append", 1000, 16), ("update", [5], True), ("draw",)]
    # steady state: write-backs held for coalescing are flushed by the request and taken in before the draw
    for it in range(3):
        client.update_priorities(np.array([100 + it]), np.array([1e-3], np.float32))
        client.update_priorities(np.array([6 + it]), np.array([1e9], np.float32))
        loader.sample()
        assert _Engine.log[-1] == ("draw",)
        seen = [i for e in _Engine.log[-3:-1] if e[0] == "update" for i in e[1]]
        assert seen[-2:] == [100 + it, 6 + it], _Engine.log[-4:]
    loader.close()
    client.close()


def test_a_learner_does_not_wait_for_other_producers(server):
    """An actor's message sits in the decode threads for a while; the learner's write-back and its draw go ahead
    (appends and write-backs of different producers are unrelated over the reference's sockets), drain() waits."""
    import zmq
    from reth_b200 import pack, service_mp as m
    addr, set_delay = server
    learner = m.Client(addr, wire="columns")
    loader = m.NumpyLoader(addr)
    # the actor lives in another process as far as the server can tell: its own producer id, its own sockets, and
    # not among the clients whose writes this process's loader requests name
    ctx = zmq.Context.instance()
    a_app, a_upd = ctx.socket(zmq.PUSH), ctx.socket(zmq.PUSH)
    a_app.connect(learner.meta["append_addr"])
    a_upd.connect(learner.meta["update_addr"])
    sent = {"a": 0, "u": 0}

    def actor_append(first):
        sent["a"] += 1
        cols, w = _cols(first)
        a_app.send(pack.encode_append_columns(cols, w, extra={pack.KEY_SRC: ["a" * 16, sent["a"]]}))

    def actor_update(idx):
        sent["u"] += 1                                        # the reference's message + the header field it ignores
        a_upd.send(pack.encode([np.array([idx]), np.array([1.0], np.float32), True],
                               extra={pack.KEY_SRC: ["a" * 16, sent["u"], sent["a"]]}))
    learner.append(*_cols(0))
    loader.sample()                                           # buffer exists, loader open
    set_delay(2000, 1.5)
    actor_append(2000)
    time.sleep(0.2)                                           # the server has received it: it is in the decode backlog
    n0 = len(_Engine.log)
    t0 = time.time()
    learner.update_priorities(np.array([7]), np.array([1.0], np.float32))
    loader.sample()
    assert time.time() - t0 < 1.0
    assert _Engine.log[n0:] == [("update", [7], True), ("draw",)]
    assert loader.stat()["backlog"] == 1
    loader.drain()                                            # everything received so far
    assert _Engine.log[-1] == ("append", 2000, 16)
    # the actor's own write-back stays behind the actor's append
    set_delay(3000, 1.0)
    actor_append(3000)
    actor_update(9)
    deadline = time.time() + 5
    while _Engine.log[-1] != ("update", [9], True) and time.time() < deadline:
        time.sleep(0.01)
    assert _Engine.log[-2:] == [("append", 3000, 16), ("update", [9], True)]
    learner.close()
    loader.close()
    a_app.close(0)
    a_upd.close(0)


def test_a_streaming_backlog_does_not_starve_the_learner(server, monkeypatch):
    """Hundreds of actor messages decoded and waiting to be committed (each commit costs the loop a few hundred
    microseconds of host work; 3 ms here): the learner's write-backs and draws are served between small batches of
    commits, not after the backlog."""
    import zmq
    from reth_b200 import pack, service_mp as m
    addr, _ = server
    real_append = _Engine.append

    def slow_append(self, cols, w, stream_key=None):
        time.sleep(0.003)
        real_append(self, cols, w, stream_key)
    monkeypatch.setattr(_Engine, "append", slow_append)
    learner = m.Client(addr, wire="columns")
    loader = m.NumpyLoader(addr)
    learner.append(*_cols(0))
    loader.sample()
    socks = []
    msg = pack.encode_append_columns(*_cols(5000))
    for _ in range(16):                                       # 16 connections x HWM 10 on either end: 300 messages fit
        sk = zmq.Context.instance().socket(zmq.PUSH)
        sk.connect(learner.meta["append_addr"])
        socks.append(sk)
    n_msgs = 0
    for k in range(300):
        try:
            socks[k % 16].send(msg, zmq.NOBLOCK)
            n_msgs += 1
        except zmq.Again:
            pass
    assert n_msgs >= 200                                      # about 0.6-0.9 s of commits
    time.sleep(0.05)
    worst = 0.0
    for it in range(10):
        t0 = time.time()
        learner.update_priorities(np.array([it]), np.array([1.0], np.float32))
        loader.sample()
        worst = max(worst, time.time() - t0)
    assert loader.stat()["backlog"] > 0                       # the learner ran while the backlog was still there
    assert worst < 0.3, worst                                 # the whole backlog is 0.6-0.9 s of commits
    loader.drain()
    assert sum(1 for e in _Engine.log if e[0] == "append" and e[1] == 5000) == n_msgs
    learner.close()
    loader.close()
    for sk in socks:
        sk.close(0)


def test_an_append_waits_for_the_write_backs_its_sender_sent_before_it(server):
    """update_priorities() then append() of one producer travel on two sockets; if the append overtakes the write-back
    and recycles the slot, the stale priority would land on the new row.  The append names the update frames sent
    before it and the loop applies those first; frames of other producers that arrive meanwhile keep their order."""
    import zmq
    from reth_b200 import pack, service_mp as m
    addr, _ = server
    boot = m.Client(addr, wire="columns")
    loader = m.NumpyLoader(addr)
    boot.append(*_cols(0))
    loader.sample()
    ctx = zmq.Context.instance()
    app, upd, other = ctx.socket(zmq.PUSH), ctx.socket(zmq.PUSH), ctx.socket(zmq.PUSH)
    app.connect(boot.meta["append_addr"])
    upd.connect(boot.meta["update_addr"])
    other.connect(boot.meta["update_addr"])
    n0 = len(_Engine.log)
    cols, w = _cols(7000)
    app.send(pack.encode_append_columns(cols, w, extra={pack.KEY_SRC: ["b" * 16, 1, 1]}))   # names update frame 1 ...
    time.sleep(0.3)
    assert len(_Engine.log) == n0                             # ... and is held back for it
    other.send(pack.encode([np.array([42]), np.array([1.0], np.float32), True], extra={pack.KEY_SRC: ["c" * 16, 1, 0]}))
    time.sleep(0.1)
    upd.send(pack.encode([np.array([3]), np.array([1.0], np.float32), True], extra={pack.KEY_SRC: ["b" * 16, 1, 0]}))
    deadline = time.time() + 5
    while len(_Engine.log) < n0 + 3 and time.time() < deadline:
        time.sleep(0.01)
    assert _Engine.log[n0:] == [("update", [3], True), ("append", 7000, 16), ("update", [42], True)]
    # the real client does the naming itself
    boot.update_priorities(np.array([8]), np.array([1.0], np.float32))
    boot.append(*_cols(8000))
    loader.drain()
    assert _Engine.log[-2:] == [("update", [8], True), ("append", 8000, 16)]
    # a frame that never arrives costs the append 2 s, not more
    t0 = time.time()
    app.send(pack.encode_append_columns(*_cols(9000), extra={pack.KEY_SRC: ["b" * 16, 2, 5]}))
    while _Engine.log[-1] != ("append", 9000, 16) and time.time() - t0 < 6:
        time.sleep(0.02)
    assert _Engine.log[-1] == ("append", 9000, 16) and 1.5 < time.time() - t0 < 4
    assert loader.stat()["dropped"]["late"] == 1
    t0 = time.time()                                          # ... and it is waited for once: the next append goes straight in
    app.send(pack.encode_append_columns(*_cols(9100), extra={pack.KEY_SRC: ["b" * 16, 3, 5]}))
    while _Engine.log[-1] != ("append", 9100, 16) and time.time() - t0 < 3:
        time.sleep(0.02)
    assert _Engine.log[-1] == ("append", 9100, 16) and time.time() - t0 < 1.0
    boot.close()
    loader.close()
    for sk in (app, upd, other):
        sk.close(0)


def test_shared_memory_ring_appends_through_the_loop(server):
    """Same-host fast path: the columns travel in the client's shared-memory ring, only a descriptor over ZeroMQ; the
    server maps the segment, hands the engine views into it and reports consumption, so the ring (1 MB here) is
    reused many times over; a message larger than the ring falls back to one ZeroMQ frame."""
    from reth_b200 import service_mp as m
    addr, _ = server
    client = m.Client(addr, wire="shm", shm_mb=1)
    loader = m.NumpyLoader(addr)
    n0 = len(_Engine.log)
    for k in range(40):                                       # 40 x 192 KB through a 1 MB ring
        ids = 10_000 * (k + 1) + np.arange(16384, dtype=np.int64)
        client.append([ids, ids.astype(np.float32)], np.ones(16384, np.float32))
    big = 5_000_000 + np.arange(200_000, dtype=np.int64)      # 2.4 MB: does not fit the ring
    client.append([big, big.astype(np.float32)], np.ones(len(big), np.float32))
    loader.drain()
    got = [e for e in _Engine.log[n0:] if e[0] == "append"]
    assert got == [("append", 10_000 * (k + 1), 16384) for k in range(40)] + [("append", 5_000_000, 200_000)]
    assert client._ring is not None and client._ring.seq == 40
    client.close()
    loader.close()


def test_malformed_messages_are_dropped_and_counted(server):
    """One bad message from one actor must not stop the server (the reference has no error channel either: messages
    are fire-and-forget, client/client.py:21-39): garbage on the append and update sockets, a truncated append, an
    unknown loader request -- counted, answered where there is someone to answer, and the loop keeps serving."""
    import zmq
    from reth_b200 import pack, service_mp as m
    addr, _ = server
    client = m.Client(addr, wire="columns")
    loader = m.NumpyLoader(addr)
    client.append(*_cols(0))
    loader.sample()
    ctx = zmq.Context.instance()
    app, upd = ctx.socket(zmq.PUSH), ctx.socket(zmq.PUSH)
    app.connect(client.meta["append_addr"])
    upd.connect(client.meta["update_addr"])
    good = bytes(pack.encode_append_rows(*_cols(100)))
    app.send(b"\x00\x00\x00\x05garbage")
    app.send(good[:-7])                                       # truncated body
    upd.send(b"RBU2" + b"\x00" * 3)                           # a frame cut short
    upd.send(b"\x00\x00\x00\x02{}")
    app.send(good)
    deadline = time.time() + 5
    while ("append", 100, 16) not in _Engine.log and time.time() < deadline:
        time.sleep(0.01)
    assert ("append", 100, 16) in _Engine.log
    st = loader.drain()
    assert st["dropped"]["append"] == 2 and st["dropped"]["update"] == 2
    with pytest.raises(RuntimeError, match="unknown op"):
        loader._call({"op": "frobnicate"})
    loader.sample()                                           # still serving
    client.close()
    loader.close()
    app.close(0)
    upd.close(0)


def test_frames_set_aside_for_one_append_are_found_by_the_next(server):
    """Two producers, each with a write-back and then an append.  While the first append waits for its sender's frame
    the other producer's frame arrives and is set aside; the second append must find it there (not wait for the
    socket to deliver it again)."""
    import zmq
    from reth_b200 import pack, service_mp as m
    addr, _ = server
    boot = m.Client(addr, wire="columns")
    loader = m.NumpyLoader(addr)
    boot.append(*_cols(0))
    loader.sample()
    ctx = zmq.Context.instance()
    app, upd = ctx.socket(zmq.PUSH), ctx.socket(zmq.PUSH)
    app.connect(boot.meta["append_addr"])
    upd.connect(boot.meta["update_addr"])

    def frame(who, idx, seq):
        return pack.encode([np.array([idx]), np.array([1.0], np.float32), True], extra={pack.KEY_SRC: [who * 16, seq, 0]})
    n0, t0 = len(_Engine.log), time.time()
    app.send(pack.encode_append_columns(*_cols(1100), extra={pack.KEY_SRC: ["x" * 16, 1, 1]}))
    time.sleep(0.1)                                           # the loop now waits for x's frame 1 and reads the update socket
    upd.send(frame("y", 21, 1))
    upd.send(frame("y", 22, 2))
    app.send(pack.encode_append_columns(*_cols(2200), extra={pack.KEY_SRC: ["y" * 16, 1, 2]}))   # behind y's two frames
    time.sleep(0.2)
    upd.send(frame("x", 11, 1))
    while len(_Engine.log) < n0 + 5 and time.time() - t0 < 5:
        time.sleep(0.01)
    assert time.time() - t0 < 1.5                             # nobody sat out a 2 s wait
    assert _Engine.log[n0:] == [("update", [11], True), ("append", 1100, 16), ("update", [21], True),
                                ("update", [22], True), ("append", 2200, 16)]
    assert loader.stat()["dropped"]["late"] == 0
    boot.close()
    loader.close()
    app.close(0)
    upd.close(0)


def test_requests_are_parked_until_the_buffer_exists(server):
    """sampler_loop.py:25-28: nothing is emitted before the buffer holds sample_start rows; a loader that asks earlier
    is parked (its struct is unknown until the first append, torch_cuda_loader.py:106-112) and served as soon as the
    first rows are in -- from another producer, here."""
    from reth_b200 import service_mp as m
    addr, _ = server
    loader = m.NumpyLoader(addr)
    got = {}

    def ask():
        got["batch"] = loader.sample()
    th = threading.Thread(target=ask, daemon=True)
    th.start()
    time.sleep(0.3)
    assert th.is_alive() and not _Engine.log                  # parked: no draw on an empty buffer
    client = m.Client(addr, wire="rows")
    client.append(*_cols(50))
    th.join(5)
    assert not th.is_alive() and _Engine.log == [("append", 50, 16), ("draw",)]
    data, idx, w = got["batch"]
    assert idx.dtype == np.int64 and w.dtype == np.float64 and len(data) == 2
    client.close()
    loader.close()


def test_bench_actor_sender_is_paced_and_flow_controlled(server, monkeypatch):
    """The actor side of bench.py's apex_process leg (bench_process._apex_sender): paced sends in the reference's row
    format over many connections, never more than `window` messages outstanding (the server's row counter is read
    through the stat request), and it stops when told to."""
    import multiprocessing as mp
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench_process
    from reth_b200 import service_mp as m
    addr, _ = server
    real_append = _Engine.append
    seen_max = {"outstanding": 0}
    stop, sent = mp.Event(), mp.Value("q", 0)

    def slow_append(self, cols, w, stream_key=None):
        time.sleep(0.004)                                     # the server takes in at most 250 messages/s
        real_append(self, cols, w, stream_key)
        seen_max["outstanding"] = max(seen_max["outstanding"], sent.value - self.shard.rows // 8)
    monkeypatch.setattr(_Engine, "append", slow_append)
    meta = m.get_meta(addr)
    th = threading.Thread(target=bench_process._apex_sender, args=(addr, meta["append_addr"], 8, 8, None, stop, sent, 32),
                          daemon=True)
    th.start()                                                # unpaced: only the window holds it back
    time.sleep(1.5)
    stop.set()
    th.join(10)
    assert not th.is_alive()
    assert 100 < sent.value < 600 and seen_max["outstanding"] <= 32 + 1
    # paced: 8-row messages at 800 rows/s = 100 messages/s
    stop2, sent2 = mp.Event(), mp.Value("q", 0)
    th = threading.Thread(target=bench_process._apex_sender, args=(addr, meta["append_addr"], 8, 8, 800, stop2, sent2, 32),
                          daemon=True)
    th.start()
    time.sleep(1.5)
    stop2.set()
    th.join(10)
    assert not th.is_alive() and 50 <= sent2.value <= 170
