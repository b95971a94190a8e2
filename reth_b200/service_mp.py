"""The replay server in its own process, reachable by address (SURVEY.md 7 "hard part 8").

The reference's callers start the server in one process and build `Client` / loaders in others
from the address string alone (test/apex-dqn/trainer.py:49-66, worker.py:22,
reth_buffer/test/test_rb.py:63-85).  This module keeps that layout:

  server process   owns the GPU shard (service._Engine) and runs ONE poll loop over
                   meta   REP    tcp   -> JSON config                (server/main_loop.py:64-76)
                   append PULL   tcp   <- reference wire format      (main_loop.py:21-61, utils/pack.py)
                   update PULL   tcp   <- [indices, weights, step]   (main_loop.py:79-88, sampler_loop.py:29-36)
                   loader ROUTER tcp   <-> loaders in other processes (replaces the per-sampler PUSH
                                           sockets + the loader worker processes)
                   A single loop = a single arrival order for appends, updates and draws, which is
                   what the reference gets from funnelling everything through one SUB socket.
  Client           PUSH sockets, fire-and-forget, HWM 10 like the reference (utils/__init__.py:12)
  loaders          ask the server for the next pre-sampled GROUP; indices/weights come back in the
                   reply, the batch itself stays in HBM and is mapped into the loader's process
                   through CUDA IPC handles of the batch ring (reth_ipc_open) -- the mechanism of
                   client/torch_cuda_loader.py:58,63,157-160.  NumpyLoader copies the rows D2H.

Append messages may be in the reference's row format (what its Client sends) or this build's
columnar format (what `Client(..., wire="columns")`, the default here, sends): see pack.py.
"""
import ctypes as C
import json
import multiprocessing as mp
import os
import signal
import socket
import struct
import sys
import threading
import time
import weakref

import numpy as np

from . import pack

ZMQ_DEFAULT_HWM = 10            # utils/__init__.py:12

# compact priority write-back frame of this build's own Client:
#   [magic, client id, frame number, count] [step, float32?, reserved, n] * count
#   [int64 indices of all messages] [float32|float64 priorities of all messages]  (one dtype per frame)
_UPD_MAGIC = b"RBU1"            # a reference-format message starts with a big-endian header length: first byte 0
_UPD_FRAME = struct.Struct("<4s8sQI")
_UPD_MAGIC2 = b"RBU2"           # ... and the number of append messages the sender had sent before this frame
_UPD_FRAME2 = struct.Struct("<4s8sQQI")
_UPD_HEAD = struct.Struct("<BBHI")


def _pick_port(host):
    s = socket.socket()
    s.bind((host, 0))
    port = s.getsockname()[1]
    s.close()
    return port


# ------------------------------------------------------------------ server --
def _default_ingest_threads():
    """Decode threads of the ingest: half the host's cores, between 2 and 8 (the decode -- lz4 frame, row scatter --
    runs in the library with the GIL released)."""
    return max(2, min(8, (os.cpu_count() or 4) // 2))


class _Ingest:
    """Decode threads of the server's ingest (replaces the single append_loop process,
    server/main_loop.py:21-61, as the ceiling of the service's append rate).

    The poll loop only RECEIVES append frames (zero-copy ZeroMQ frames) and numbers them; `n_threads` workers
    undo the wire format -- header, lz4 body, and for the reference's row format the row -> column transpose
    (reth_pack_scatter_rows, one library call per message, GIL released) -- writing the columns straight into
    a slab of PINNED host memory; the poll loop then commits finished messages strictly in arrival order
    (slot = arrival order, cache_policy/fifo_policy.py:11-18) with an asynchronous host->device copy out of the
    slab, which is recycled once a CUDA event behind the copy has fired."""

    def __init__(self, n_threads, device, max_slabs=24):
        import queue
        self.device = device
        self.q = queue.Queue()
        self.lock = threading.Lock()
        self.cv = threading.Condition(self.lock)
        self.done = {}                 # message number -> result tuple or exception
        self.submitted = self.committed = 0
        self.free = []                 # pinned slabs (torch uint8 tensors), largest first
        self.slabs = 0
        self.max_slabs = max_slabs
        self.window = max(1, max_slabs // 2)
        self.busy = []                 # (cuda event, slab): copies still reading a slab
        self.stop = False
        self.threads = [threading.Thread(target=self._work, daemon=True) for _ in range(max(1, n_threads))]
        for t in self.threads:
            t.start()

    def _slab(self, nbytes, seq):
        import torch
        with self.cv:
            while True:
                # slabs go to the messages that commit next: a message more than `window` ahead of the commit point
                # waits, so the decoded-but-uncommitted messages can never hold every slab while the message the commit
                # point waits for has none (slabs come back only through commits)
                if seq - self.committed >= self.window:
                    self.cv.wait(0.01)
                    continue
                for k, t in enumerate(self.free):
                    if t.numel() >= nbytes:
                        return self.free.pop(k)
                if self.slabs < self.max_slabs:
                    self.slabs += 1
                    break
                if self.free:                     # every free slab is too small: replace the smallest
                    self.free.pop()
                    break
                self.cv.wait(0.01)                # back-pressure: every slab is being filled or copied from
        size = max(1 << 20, 1 << int(np.ceil(np.log2(max(1, nbytes)))))
        return torch.empty(size, dtype=torch.uint8).pin_memory()

    def _work(self):
        while True:
            item = self.q.get()
            if item is None:
                return
            seq, frame = item
            slab = []

            def alloc(nbytes):
                t = self._slab(nbytes, seq)
                slab.append(t)
                return t.numpy()[:nbytes]
            try:
                res = pack.open_append(frame.buffer, alloc) + (slab[0] if slab else None,)
            except Exception as e:                 # reported and dropped at commit time, in order
                res = e
                for t in slab:
                    self.release(t)
            with self.cv:
                self.done[seq] = res
                self.cv.notify_all()

    def submit(self, frame):
        self.q.put((self.submitted, frame))
        self.submitted += 1

    def pop_ready(self):
        """The next message in arrival order if its decode has finished, else None."""
        with self.lock:
            if self.committed in self.done:
                res = self.done.pop(self.committed)
                self.committed += 1
                self.cv.notify_all()
                return res
        return None

    @property
    def backlog(self):
        return self.submitted - self.committed

    def wait_next(self, timeout):
        with self.cv:
            if self.committed not in self.done and self.backlog:
                self.cv.wait(timeout)

    def hold(self, slab, stream):
        """The host->device copy out of `slab` was just enqueued on `stream`: recycle it once that has run."""
        import torch
        ev = torch.cuda.Event()
        ev.record(torch.cuda.ExternalStream(stream, device=self.device))
        self.busy.append((ev, slab))

    def release(self, slab):
        with self.cv:
            self.free.append(slab)
            self.free.sort(key=lambda t: -t.numel())
            self.cv.notify_all()

    def retire(self):
        while self.busy and self.busy[0][0].query():
            self.release(self.busy.pop(0)[1])

    def close(self):
        for _ in self.threads:
            self.q.put(None)


def server_main(capacity, batch_size, host, ports, samplers, cache_policy, device, max_ring, ready_evt, frames=None, float_frames=None,
                nstep=None):
    """Body of the server process (main_loop.py:110-216 + append_loop + sampler_loop in one loop)."""
    import zmq
    from . import _cabi, service

    signal.signal(signal.SIGINT, signal.SIG_IGN)
    stop = {"flag": False}

    def on_term(*_):
        # leave the loop; if that has not happened 10 s later, say where every thread stands and exit anyway:
        # proc.terminate() + proc.join() (trainer.py:72-74) must never hang its caller
        import faulthandler
        stop["flag"] = True
        faulthandler.dump_traceback_later(10, exit=True)
        if os.environ.get("RETH_B200_DEBUG") and "ingest" in stop:
            g = stop["ingest"]
            print(f"reth_b200 server: ingest submitted {g.submitted} committed {g.committed} decoded-waiting {sorted(g.done)[:8]} "
                  f"slabs {g.slabs} free {len(g.free)} busy {len(g.busy)} "
                  f"busy[0] done {g.busy[0][0].query() if g.busy else None}", file=sys.stderr, flush=True)
    signal.signal(signal.SIGTERM, on_term)

    eng = service._Engine(capacity, batch_size, samplers, cache_policy, device, max_ring, float_frames=float_frames, frames=frames,
                          nstep=nstep)
    ctx = zmq.Context.instance()
    ctx.set(zmq.IO_THREADS, int(os.environ.get("RETH_B200_IO_THREADS", "4")))   # many actor connections (worker.py:22,57-61)
    ctx.setsockopt(zmq.LINGER, 0)
    addrs = {k: f"tcp://{host}:{p}" for k, p in ports.items()}
    meta = ctx.socket(zmq.REP)
    meta.bind(addrs["meta"])
    app = ctx.socket(zmq.PULL)
    app.set_hwm(ZMQ_DEFAULT_HWM)
    app.bind(addrs["append"])
    upd = ctx.socket(zmq.PULL)
    upd.set_hwm(ZMQ_DEFAULT_HWM)
    upd.bind(addrs["update"])
    ldr = ctx.socket(zmq.ROUTER)
    ldr.bind(addrs["loader"])
    config = {"engine": "reth_b200", "capacity": capacity, "batch_size": batch_size, "device": device,
              "meta_addr": addrs["meta"], "append_addr": addrs["append"], "update_addr": addrs["update"],
              "loader_addr": addrs["loader"], "pid": os.getpid(), "host_id": _host_id(),
              "sampler_info": {t: {"topic": t, "sampler_cls": st["kind"], "sample_start": st["sample_start"],
                                   "batch_size": batch_size, "primary": t == eng.topic}
                               for t, st in eng.topics.items()}}
    poller = zmq.Poller()
    for s in (meta, app, upd, ldr):
        poller.register(s, zmq.POLLIN)
    ingest = _Ingest(int(os.environ.get("RETH_B200_INGEST_THREADS", str(_default_ingest_threads()))), device)
    stop["ingest"] = ingest
    waiting = []                 # loader requests parked until the buffer is ready (sampler_loop.py:25-28)
    owed = []                    # loaders that asked for a group which is still being drawn / gathered
    seen = {}                    # client id -> [last append message, last update frame] taken in
    segments = {}                # shared-memory ingest rings of same-host clients: name -> _ShmSegment
    pending = []                 # (cuda event, segment, seq): H2D copies out of a ring still in flight
    dropped = {"append": 0, "update": 0, "late": 0}   # malformed messages dropped / requests that went ahead of late writes
    trace_on = bool(os.environ.get("RETH_B200_TRACE"))

    def trace(*what):                                  # diagnostic: the order in which the loop takes things in
        if trace_on:
            print("reth_b200 server trace:", *what, file=sys.stderr, flush=True)
    ready_evt.set()

    def append_from_ring(head, weights, key=None):
        import torch
        seg = segments.get(head[pack.KEY_SHM])
        if seg is None:
            seg = segments[head[pack.KEY_SHM]] = _ShmSegment(head[pack.KEY_SHM], head["size"], device)
        cols = [seg.buf[at:at + nb].view(np.dtype(dt)).reshape(sh) for dt, sh, at, nb in head["cols"]]
        eng.append(cols, weights, stream_key=key)
        if seg.pinned:                                  # the copies run asynchronously out of the ring
            ev = torch.cuda.Event()
            ev.record(torch.cuda.ExternalStream(eng.stream, device=device))
            pending.append((ev, seg, int(head["seq"])))
        else:                                           # pageable source: staged before the call returned
            seg.consumed[0] = int(head["seq"])

    def retire_ring_copies():
        while pending and pending[0][0].query():
            _, seg, seq = pending.pop(0)
            seg.consumed[0] = seq

    def reply(ident, obj, *blobs):
        ldr.send_multipart([ident, b"", json.dumps(obj).encode()] + list(blobs))

    def deliver():
        """Send every group that is owed and complete; nothing here blocks on the device."""
        while owed:
            got = eng.try_deliver()
            if got is None:
                return
            g, idx, w = got
            reply(owed.pop(0), {"ok": True, "op": "next", "group": int(g), "first_slot": int(g * eng.group),
                                "n": int(eng.group)}, idx.tobytes(), w.tobytes())

    def open_update(buf):
        """One update frame: the reference's [indices, weights, step] message (client.py:37-39), or this build's
        compact frame holding one or more such messages back to back.
        -> (client id | None, frame number, appends the sender had sent before it, apply())."""
        mv = memoryview(buf)
        if mv[:4] not in (_UPD_MAGIC, _UPD_MAGIC2):
            (indices, weights, step), src = pack.decode_src(mv)
            def apply():
                eng.update(np.asarray(indices), np.asarray(weights), bool(step))
            if not src:
                return None, 0, 0, apply
            return src[0], int(src[1]), int(src[2]) if len(src) > 2 else 0, apply
        if mv[:4] == _UPD_MAGIC2:
            _, cid, seq, aseq, count = _UPD_FRAME2.unpack_from(mv, 0)
            at = _UPD_FRAME2.size
        else:
            _, cid, seq, count = _UPD_FRAME.unpack_from(mv, 0)
            aseq, at = 0, _UPD_FRAME.size
        sizes, steps, f32 = [], [], 1
        for _ in range(count):
            step, f32, _, n = _UPD_HEAD.unpack_from(mv, at)
            at += _UPD_HEAD.size
            sizes.append(n)
            steps.append(step)
        total = sum(sizes)
        idx = np.frombuffer(mv, np.int64, total, at)
        w = np.frombuffer(mv, np.float32 if f32 else np.float64, total, at + 8 * total)
        return cid.hex(), seq, aseq, lambda: eng.update_block(idx, w, sizes, steps)

    def drain_appends(wait=False, budget=4):
        """Receive every pending append frame (decoded by the ingest threads), then commit what has been decoded,
        strictly in arrival order.  wait: block until everything received so far is committed (drain()).  Otherwise
        at most `budget` messages per call: with the actors streaming, the decode threads finish messages as fast as
        this loop commits them, and a call that committed "everything that is ready" did not return for milliseconds
        while a learner's request sat in its socket."""
        while True:
            try:
                ingest.submit(app.recv(zmq.NOBLOCK, copy=False))
            except zmq.Again:
                break
        taken = 0
        while True:
            if not wait and taken >= budget:
                break
            res = ingest.pop_ready()
            if res is None:
                if wait and ingest.backlog:
                    retire_ring_copies()
                    ingest.retire()                    # the decode threads may be waiting for a staging slab
                    ingest.wait_next(0.002)
                    continue
                break
            src = None
            taken += 1
            try:
                if isinstance(res, Exception):
                    raise res
                named = res[3]
                if named and len(named) > 2 and seen.get(named[0], (0, 0))[1] < int(named[2]):
                    await_updates(named[0], int(named[2]))
                src = take_append(*res)
                trace("append", src, "rows", len(res[2]))
            except Exception as e:                     # one bad message from one actor must not stop the server
                dropped["append"] += 1
                print(f"reth_b200 server: append message dropped ({type(e).__name__}: {e})", file=sys.stderr, flush=True)
            if src:
                seen.setdefault(src[0], [0, 0])[0] = int(src[1])
            retire_ring_copies()
            ingest.retire()
        retire_ring_copies()
        ingest.retire()

    def take_append(kind, payload, weights, src, slab):
        if kind == "steps":
            head, frames, others = payload
            reset = head.get("reset")
            eng.append_steps((src[0] if src else "", head[pack.KEY_STEPS]), frames, others, weights,
                             None if reset is None else np.asarray(reset))
        elif kind == "shm":
            try:
                append_from_ring(payload, weights, src[0] if src else None)
            except FileNotFoundError:              # the producer vanished with its ring: the rows are lost
                print(f"reth_b200 server: ingest ring {payload[pack.KEY_SHM]} is gone, message dropped",
                      file=sys.stderr, flush=True)
        else:
            eng.append(payload, weights, stream_key=src[0] if src else None)
        if slab is not None:                       # the columns sit in pinned staging: the copy is asynchronous
            if eng.stream is not None:
                ingest.hold(slab, eng.stream)
            else:
                ingest.release(slab)
        return src

    deferred = []                # update frames of OTHER producers received while an append waited for its sender's

    def await_updates(cid, useq, timeout=2.0):
        """An append names the update frames its sender had sent before it (they travel on another socket): apply
        those first, or a write-back could land on the row that replaces the one it was meant for.  Runs inside
        drain_appends, between two commits: frames of other producers are set aside, in order, for drain_updates
        (applying them here could require commits of their own)."""
        t0 = time.time()
        aside, deferred[:] = list(deferred), []        # frames set aside during an earlier wait may be this sender's
        while seen.get(cid, (0, 0))[1] < useq and time.time() - t0 < timeout:
            if aside:
                frame = aside.pop(0)
            else:
                try:
                    frame = upd.recv(zmq.NOBLOCK, copy=False)
                except zmq.Again:
                    upd.poll(1)
                    continue
            try:
                ocid, seq, _, apply = open_update(frame.buffer)
                if ocid != cid:
                    deferred.append(frame)
                    continue
                apply()                                # the appends this frame names came before this one: they are in
                trace("update frame", cid, seq, "ahead of append")
                seen.setdefault(cid, [0, 0])[1] = seq
            except Exception as e:
                dropped["update"] += 1
                print(f"reth_b200 server: update message dropped ({type(e).__name__}: {e})", file=sys.stderr, flush=True)
        deferred.extend(aside)                         # not looked at: they stay in order behind the ones that were
        if seen.get(cid, (0, 0))[1] < useq:
            dropped["late"] += 1
            seen.setdefault(cid, [0, 0])[1] = useq     # lost: later messages do not wait for it again
            print(f"reth_b200 server: an append went ahead of update frame {useq} of its sender, which never arrived",
                  file=sys.stderr, flush=True)

    def drain_updates():
        while True:
            if deferred:
                frame = deferred.pop(0)
            else:
                try:
                    frame = upd.recv(zmq.NOBLOCK, copy=False)
                except zmq.Again:
                    break
            try:
                cid, seq, aseq, apply = open_update(frame.buffer)
                # Order is per producer, as over the reference's sockets: a write-back stays behind the appends ITS sender
                # sent before it (they travel on another socket and pass the decode threads), not behind every actor's
                # -- with 256 actors streaming there is always a decode backlog, and a learner's write-backs (and its
                # draws, handle_loader) waiting for all of it ran at a few thousand transitions/s (bench apex_process)
                if cid is not None and aseq:
                    await_appends(cid, aseq)
                apply()
                trace("update frame", cid, seq, "names appends", aseq, "shard", eng.shard is not None, "seen", seen.get(cid))
                if cid is not None:
                    seen.setdefault(cid, [0, 0])[1] = seq
            except Exception as e:
                dropped["update"] += 1
                print(f"reth_b200 server: update message dropped ({type(e).__name__}: {e})", file=sys.stderr, flush=True)

    def await_appends(cid, aseq, timeout=2.0):
        """Commit appends (arrival order) until message `aseq` of producer `cid` is in."""
        t0 = hard = time.time()                        # the clock runs only while nothing received is waiting to go in:
        while seen.get(cid, (0, 0))[0] < aseq and time.time() - t0 < timeout and time.time() - hard < 30:   # a decode may
            drain_appends()                            # take seconds (the first one pins its staging slab: CUDA start-up);
                                                       # 30 s in all, or streaming actors would keep a lost message awaited
            if seen.get(cid, (0, 0))[0] < aseq:
                if ingest.backlog:
                    ingest.wait_next(0.002)
                    t0 = time.time()
                else:
                    poller.poll(1)
        if seen.get(cid, (0, 0))[0] < aseq:
            dropped["late"] += 1
            seen.setdefault(cid, [0, 0])[0] = aseq     # lost (or dropped as malformed): do not wait for it again
            print(f"reth_b200 server: a write-back went ahead of append {aseq} of its sender, which never arrived",
                  file=sys.stderr, flush=True)

    def await_writes(after, timeout=2.0):
        """A loader's request names the appends / write-backs its own process sent before it (they
        travel on other sockets): take them in first, so the request sees them -- the arrival order
        the reference gets from one SUB socket (sampler_loop.py:29-36)."""
        def behind():
            return any(seen.get(c, (0, 0))[0] < a or seen.get(c, (0, 0))[1] < u for c, (a, u) in after.items())
        if not after or not behind():
            return
        t0 = hard = time.time()
        while behind() and time.time() - t0 < timeout and time.time() - hard < 30:
            drain_appends()
            drain_updates()
            if behind():
                if ingest.backlog:
                    ingest.wait_next(0.002)
                    t0 = time.time()                   # received and being decoded: not lost
                else:
                    poller.poll(1)
        if behind():                                   # a write named by the request never arrived (dropped / lost)
            dropped["late"] += 1
            for c, (a, u) in after.items():
                have = seen.setdefault(c, [0, 0])
                have[0], have[1] = max(have[0], a), max(have[1], u)
            print(f"reth_b200 server: a loader request went ahead without writes sent before it ({timeout:.1f} s "
                  "wait): ordering against those writes is not guaranteed", file=sys.stderr, flush=True)

    def handle_loader(ident, req):
        op = req.get("op")
        if op == "stat":                               # counters as they stand: no draining, no waiting (producers' flow control)
            rows = eng.shard.stats()["appended_rows"] if eng.shard is not None else 0
            reply(ident, {"ok": True, "op": "stat", "cnt": eng.cnt, "appended_rows": rows, "backlog": ingest.backlog,
                          "dropped": dropped})
            return
        if op == "drain" and ingest.backlog:
            drain_appends(wait=True)                   # drain(): everything received so far
        else:
            drain_appends()                            # what is decoded goes in before a draw (sampler_loop.py:29-36)
        await_writes(req.get("after"))                 # ... and everything the requesting process sent before its request
        if op == "open":
            if eng.shard is None:                      # row struct unknown until the first append
                waiting.append((ident, req))
                return
            ex = _cabi.IpcExport()
            eng.shard._ck(_cabi.lib().reth_replay_ipc_export(eng.shard._h, C.byref(ex)))
            struct = [[np.dtype(dt).str, list(sh)] for dt, sh in eng.shard.out_struct]
            if req.get("topic", eng.topic) != eng.topic:   # another sampler topic: served draw by draw
                reply(ident, {"ok": True, "op": "open", "struct": struct, "batch_size": batch_size, "group": 1,
                              "ahead": 0, "lend": 1, "ring_depth": eng.shard.ring_depth}, bytes(ex))
                return
            # fit the loader's geometry into the batch ring: three groups when it asks ahead
            group = max(1, int(req.get("prefetch", 1)))
            ahead = max(0, int(req.get("ahead", 0)))
            if ahead > 0 and 3 * group > eng.max_ring:
                group = max(1, eng.max_ring // 3)
            ahead = min(ahead, max(0, eng.max_ring // group - 1))
            eng.open_pipe(group, ahead, req.get("lend", 1))
            reply(ident, {"ok": True, "op": "open", "struct": struct, "batch_size": batch_size, "group": eng.group,
                          "ahead": eng.ahead, "lend": eng.lend, "ring_depth": eng.shard.ring_depth}, bytes(ex))
        elif op == "next" and req.get("topic", eng.topic) != eng.topic:
            topic = req["topic"]
            if not eng.topic_ready(topic):
                waiting.append((ident, req))
                return
            slot, idx, w = eng.sample_topic(topic)
            reply(ident, {"ok": True, "op": "next", "group": -1, "first_slot": int(slot), "n": 1},
                  np.ascontiguousarray(idx[:batch_size]).tobytes(), np.ascontiguousarray(w[:batch_size]).tobytes())
        elif op == "next":
            if not eng.ready():
                waiting.append((ident, req))
                return
            trace("draw", req.get("after"), "seen", seen)
            eng.request_group()
            owed.append(ident)
            deliver()
        elif op == "rows":                             # NumpyLoader: rows of a lent slot, device -> host
            data, _, _ = eng.shard.slot_read(int(req["slot"]), with_rows=True, stream=eng.stream)
            reply(ident, {"ok": True, "op": "rows"}, *[np.ascontiguousarray(c).tobytes() for c in data])
        elif op == "ring":                             # parity / debug: rows [first, first + n) of a ring column
            col = eng.shard.ring_read(int(req["column"]), int(req["first"]), int(req["n"]), stream=eng.stream)
            reply(ident, {"ok": True, "op": "ring", "dtype": col.dtype.str, "shape": list(col.shape)}, col.tobytes())
        elif op == "close":
            owed[:] = [i for i in owed if i != ident]
            if req.get("topic", eng.topic) == eng.topic:
                eng.release_pipe()
            reply(ident, {"ok": True, "op": "close"})
        elif op == "drain":
            eng.drain()
            reply(ident, {"ok": True, "op": "drain", "cnt": eng.cnt, "dropped": dropped})
        else:
            reply(ident, {"ok": False, "op": op, "error": f"unknown op {op!r}"})

    last_traffic = 0.0
    try:
        while not stop["flag"]:
            # a delivery is owed, or traffic within the last 2 ms: spin (a blocking poll costs a scheduler
            # wake-up per message; the reference's sampler loop spins in zmq.select too, sampler_loop.py:22-24)
            busy = owed or ingest.backlog or (eng.loaders > 0 and time.perf_counter() - last_traffic < 0.002)
            events = dict(poller.poll(0 if busy else (1 if pending else 50)))
            if events:
                last_traffic = time.perf_counter()
            if meta in events:
                meta.recv()
                meta.send_json(config)
            # drain every pending write before serving a draw (sampler_loop.py:29-36)
            drain_appends()
            drain_updates()
            if waiting and eng.shard is not None:
                parked, waiting[:] = list(waiting), []
                for ident, req in parked:
                    handle_loader(ident, req)
            while True:
                try:
                    ident, _, payload = ldr.recv_multipart(zmq.NOBLOCK)
                except zmq.Again:
                    break
                req = {}
                try:
                    req = json.loads(payload)
                    handle_loader(ident, req)
                except Exception as e:                 # keep serving; the loader raises on its side
                    reply(ident, {"ok": False, "op": req.get("op"), "error": f"{type(e).__name__}: {e}"})
            deliver()
    finally:
        ingest.close()
        eng.close()
        for seg in segments.values():
            seg.close()
        for s in (meta, app, upd, ldr):
            s.close(0)


class _ShmSegment:
    """Server side of a client's shared-memory ingest ring: attached once, registered with CUDA so the
    host->device copies of the rows are true asynchronous DMA out of the ring."""

    def __init__(self, name, size, device):
        # mapped directly: SharedMemory(name=...) would register the segment with this process's resource tracker,
        # which a spawned server SHARES with the client that owns (and unlinks) it -- the tracker then complains
        # about the second unregister at exit
        import _posixshmem
        import mmap
        fd = _posixshmem.shm_open("/" + name.lstrip("/"), os.O_RDWR, mode=0o600)
        try:
            self.shm = mmap.mmap(fd, os.fstat(fd).st_size)
        finally:
            os.close(fd)
        if len(self.shm) < int(size):
            raise ValueError(f"ingest ring {name}: {len(self.shm)} bytes mapped, the message says {size}")
        self.buf = np.frombuffer(self.shm, np.uint8)
        self.consumed = np.frombuffer(self.shm, np.uint64, count=1)
        self.pinned = False
        try:
            import torch
            torch.cuda.set_device(device)
            rc = torch.cuda.cudart().cudaHostRegister(self.buf.ctypes.data, self.buf.nbytes, 0)
            self.pinned = int(rc) == 0
        except Exception:
            self.pinned = False

    def close(self):
        try:
            if self.pinned:
                import torch
                torch.cuda.cudart().cudaHostUnregister(self.buf.ctypes.data)
            self.buf = self.consumed = None
            self.shm.close()
        except Exception:
            pass


# ------------------------------------------------------------ public surface --
def start_server(capacity, batch_size, host=None, port=None, samplers=None, cache_policy=None, device=0,
                 max_ring=32, frames=None, float_frames=None, nstep=None):
    """reth_buffer.start_server (__init__.py:11-27): spawn the server process, return
    (proc, meta_addr).  `proc` supports terminate()/join() as callers expect (trainer.py:72-74)."""
    host = host or "127.0.0.1"
    ports = {"meta": port or _pick_port(host)}
    for k in ("append", "update", "loader"):
        ports[k] = _pick_port(host)
    if samplers is not None:
        samplers = [dict(s, sampler_cls=getattr(s["sampler_cls"], "__name__", str(s["sampler_cls"]))) for s in samplers]
    ctx = mp.get_context("spawn")
    ready = ctx.Event()
    proc = ctx.Process(target=_server_entry, args=(capacity, batch_size, host, ports, samplers, cache_policy, device,
                                                   max_ring, ready, frames, float_frames, nstep), daemon=True)
    proc.start()
    if not ready.wait(120):
        proc.terminate()
        raise RuntimeError("replay server did not come up")
    return proc, f"tcp://{host}:{ports['meta']}"


def _server_entry(capacity, batch_size, host, ports, samplers, cache_policy, device, max_ring, ready, frames=None,
                  float_frames=None, nstep=None):
    from .sampler import PERSampler
    if samplers is not None:
        for s in samplers:
            pass                                       # class names are resolved by service._Engine
    prof_path = os.environ.get("RETH_B200_SERVER_PROFILE")      # diagnostic: cProfile of the server loop
    if not prof_path:
        server_main(capacity, batch_size, host, ports, samplers, cache_policy, device, max_ring, ready, frames, float_frames, nstep)
        return
    import cProfile
    import io
    import pstats
    pr = cProfile.Profile()
    try:
        pr.runcall(server_main, capacity, batch_size, host, ports, samplers, cache_policy, device, max_ring, ready, frames,
                   float_frames, nstep)
    finally:
        out = io.StringIO()
        pstats.Stats(pr, stream=out).sort_stats("tottime").print_stats(25)
        pstats.Stats(pr, stream=out).sort_stats("cumtime").print_stats("reth_b200", 40)
        with open(prof_path, "w") as f:
            f.write(out.getvalue())


def start_per(capacity, batch_size, alpha=0.6, beta=0.4, sample_start=1000, num_sampler_procs=1,
              host=None, port=None, cache_policy=None, device=0, max_ring=32, frames=None, float_frames=None, nstep=None):
    """reth_buffer.start_per (__init__.py:30-49)."""
    from .sampler import PERSampler
    samplers = [{"sampler_cls": PERSampler, "num_procs": num_sampler_procs, "sample_start": sample_start,
                 "kwargs": {"alpha": alpha, "beta": beta}}]
    return start_server(capacity, batch_size, host, port, samplers, cache_policy, device, max_ring, frames, float_frames, nstep)


_META_CACHE = {}


def get_meta(meta_addr, timeout_ms=30000):
    """utils/__init__.py:50-59 (REQ to the meta endpoint, cached per address)."""
    if meta_addr in _META_CACHE:
        return _META_CACHE[meta_addr]
    import zmq
    ctx = zmq.Context.instance()
    sock = ctx.socket(zmq.REQ)
    sock.setsockopt(zmq.LINGER, 0)
    sock.setsockopt(zmq.IMMEDIATE, 1)
    sock.connect(meta_addr)
    sock.send(b"")
    if not sock.poll(timeout_ms):
        sock.close(0)
        raise ConnectionError(f"no replay server answers at {meta_addr!r}")
    res = sock.recv_json()
    sock.close(0)
    _META_CACHE[meta_addr] = res
    return res


def _host_id():
    """Identifies the machine (same value in every process of one host / container)."""
    try:
        with open("/proc/sys/kernel/random/boot_id") as f:
            return socket.gethostname() + ":" + f.read().strip()
    except OSError:
        return socket.gethostname()


class _ShmRing:
    """Client side of the shared-memory ingest ring: a POSIX shm segment owned by one client.
    Bytes 0..7 hold the number of messages the server has finished copying out (written by the
    server); message bodies are bump-allocated after a 64-byte header and reused once consumed.
    The reference is an "intra-node shared-mem based" service too (rows live in an LMDB file in
    /dev/shm, main_loop.py:128-132); here shared memory carries the rows straight to the H2D copy."""

    def __init__(self, size):
        from multiprocessing import shared_memory
        self.shm = shared_memory.SharedMemory(create=True, size=int(size))
        self.buf = np.frombuffer(self.shm.buf, np.uint8)
        self.consumed = np.frombuffer(self.shm.buf, np.uint64, count=1)
        self.consumed[0] = 0
        self.size = int(size)
        self.cur = 64
        self.seq = 0                      # messages produced
        self.inflight = []                # (seq, start, end) not yet known to be consumed
        self._closed = False
        import atexit
        atexit.register(self.close)       # before interpreter teardown: views must go before the mapping

    def _retire(self):
        done = int(self.consumed[0])
        self.inflight = [m for m in self.inflight if m[0] > done]

    def reserve(self, nbytes, timeout=120.0):
        """-> offset of a free contiguous region (blocks while the server still reads an overlapping one)."""
        nbytes = (int(nbytes) + 63) & ~63
        if nbytes + 64 > self.size:
            return None
        if self.cur + nbytes > self.size:
            self.cur = 64
        a, b = self.cur, self.cur + nbytes
        t0 = time.time()
        while True:
            self._retire()
            if not any(s < b and a < e for _, s, e in self.inflight):
                break
            if time.time() - t0 > timeout:
                raise TimeoutError("shared-memory ingest ring: the server does not drain it")
            time.sleep(0.0002)
        self.cur = b
        self.seq += 1
        self.inflight.append((self.seq, a, b))
        return a

    def close(self, timeout=30.0):
        """Fire-and-forget appends may still sit in the ring: wait until the server has copied them out
        (or `timeout`), then unlink the segment."""
        if self._closed:
            return
        self._closed = True
        try:
            t0 = time.time()
            while self.consumed is not None and int(self.consumed[0]) < self.seq and time.time() - t0 < timeout:
                time.sleep(0.001)
            self.buf = self.consumed = None
            self.shm.close()
            self.shm.unlink()
        except Exception:
            pass


class Client:
    """client/client.py:8-39 against a server in another process.  wire:
      "rows"    exactly what the reference's Client sends (one inner message per transition);
      "columns" the columns as they are, in one ZeroMQ frame;
      "shm"     same-host fast path: the columns are written once into a shared-memory ring the
                server copies host->device from, only a small descriptor travels over ZeroMQ;
      "auto"    (default) "shm" when the server runs on this host, else "columns"."""

    def __init__(self, meta_addr, wire="auto", shm_mb=None, coalesce=None):
        import zmq
        self.ctx = zmq.Context.instance()
        self.meta = get_meta(meta_addr)
        # priority write-backs are fire-and-forget (client.py:37-39); up to `coalesce` consecutive messages
        # share one ZeroMQ frame (each keeps its own step flag and is applied separately, in order).  A
        # partial frame is sent before this process asks for a new group, at append(), close(), and by a
        # timer once it is 2 ms old.  coalesce=1 (and wire="rows") sends every message on its own.
        self.coalesce = max(1, int(coalesce if coalesce is not None else os.environ.get("RETH_B200_UPDATE_COALESCE", 8)))
        self._lock = threading.Lock()
        self._id = os.urandom(8)
        self._aseq = self._useq = 0                            # append messages / update frames sent so far
        self._heads, self._idxs, self._ws, self._held, self._held_t0, self._held_f32 = [], [], [], 0, 0.0, True
        self._timer = None
        self._closed = False
        _CLIENTS.setdefault(self.meta["meta_addr"], weakref.WeakSet()).add(self)
        if wire == "auto":
            wire = "shm" if self.meta.get("host_id") == _host_id() else "columns"
        self.wire = wire
        self._ring = None
        self._shm_bytes = int(shm_mb if shm_mb is not None else os.environ.get("RETH_B200_SHM_MB", 512)) << 20
        self.append_sock = self.ctx.socket(zmq.PUSH)
        self.append_sock.set_hwm(ZMQ_DEFAULT_HWM)
        self.append_sock.connect(self.meta["append_addr"])
        self.update_sock = self.ctx.socket(zmq.PUSH)
        self.update_sock.set_hwm(ZMQ_DEFAULT_HWM)
        self.update_sock.connect(self.meta["update_addr"])

    def _append_shm(self, data, weights):
        if self._ring is None:
            self._ring = _ShmRing(self._shm_bytes)
        cols = [np.ascontiguousarray(c) for c in data]
        total = sum((c.nbytes + 63) & ~63 for c in cols)
        at = self._ring.reserve(total)
        if at is None:
            return False                                      # larger than the ring: fall back to one frame
        desc = []
        for c in cols:
            self._ring.buf[at:at + c.nbytes] = c.reshape(-1).view(np.uint8)
            desc.append([c.dtype.str, list(c.shape), at, c.nbytes])
            at += (c.nbytes + 63) & ~63
        head = {pack.KEY_SHM: self._ring.shm.name, "size": self._ring.size, "seq": self._ring.seq, "cols": desc}
        self.append_sock.send(pack.encode([head, weights], extra=self._next_append()), copy=False)
        return True

    def _next_append(self):
        self._aseq += 1                                       # + the update frames sent before it (append() flushes first)
        return {pack.KEY_SRC: [self._id.hex(), self._aseq, self._useq]}

    def append(self, data, weights, compress=False):
        assert isinstance(data, (list, tuple))                # client.py:22
        for col in data:
            assert isinstance(col, np.ndarray)                # :24
            assert len(col) == len(weights)                   # :25
        weights = np.asarray(weights)
        self.flush()                                          # keep write-backs ahead of later appends
        if self.wire == "rows":
            msg = pack.encode_append_rows(data, weights, compress=compress, extra=self._next_append())
        elif self.wire == "shm" and self._append_shm(data, weights):
            return                                            # same host: nothing to gain from compress
        else:
            msg = pack.encode_append_columns(data, weights, compress=compress, extra=self._next_append())
        self.append_sock.send(msg, copy=False)

    def append_steps(self, env, frames, others, weights, reset=None):
        """Frame-de-duplicated ingest for a server started with frames=...: the observation frame after
        each of n steps of environment stream `env`, the other columns of the n transitions, |TD|
        weights; `reset` = first frame of an episode the stream starts.  Each frame travels once."""
        weights = np.asarray(weights)
        assert len(frames) == len(weights) and all(len(o) == len(weights) for o in others)
        self.flush()
        head = {pack.KEY_STEPS: str(env), "reset": None if reset is None else np.ascontiguousarray(reset, dtype=np.uint8)}
        msg = pack.encode([head, np.ascontiguousarray(frames, dtype=np.uint8), [np.ascontiguousarray(o) for o in others],
                           weights], extra=self._next_append())
        self.append_sock.send(msg, copy=False)

    def update_priorities(self, indices, weights, step=True):
        assert len(indices) == len(weights)                   # :38
        if self.wire == "rows":                               # the reference's message (+ a header field it ignores)
            with self._lock:
                self._useq += 1
                self.update_sock.send(pack.encode([np.asarray(indices), np.asarray(weights), step],
                                                  extra={pack.KEY_SRC: [self._id.hex(), self._useq, self._aseq]}), copy=False)
            return
        idx = indices if (type(indices) is np.ndarray and indices.dtype == np.int64) else np.asarray(indices, dtype=np.int64)
        w = weights if type(weights) is np.ndarray else np.asarray(weights)
        f32 = w.dtype == np.float32
        if not f32 and w.dtype != np.float64:
            w = w.astype(np.float64)
        with self._lock:
            if self._held and f32 != self._held_f32:           # one priority dtype per frame
                self._flush_locked()
            self._held_f32 = f32
            self._heads.append(_UPD_HEAD.pack(1 if step else 0, 1 if f32 else 0, 0, idx.size))
            self._idxs.append(idx.tobytes())                  # copies: the caller may reuse its arrays
            self._ws.append(w.tobytes())
            self._held += 1
            if self._held >= self.coalesce:
                self._flush_locked()
            elif self._held == 1:
                self._held_t0 = time.perf_counter()
                if self._timer is None:
                    self._timer = threading.Thread(target=_flush_timer, args=(weakref.ref(self),), daemon=True)
                    self._timer.start()

    def _flush_locked(self):
        if self._held:
            self._useq += 1
            frame = b"".join([_UPD_FRAME2.pack(_UPD_MAGIC2, self._id, self._useq, self._aseq, self._held)] + self._heads
                             + self._idxs + self._ws)
            self._heads, self._idxs, self._ws, self._held = [], [], [], 0
            self.update_sock.send(frame, copy=False)

    def flush(self):
        """Send the write-backs held back for coalescing.  -> (append messages, update frames) sent so far."""
        with self._lock:
            if not self._closed:
                self._flush_locked()
            return self._aseq, self._useq

    def close(self):
        with self._lock:
            if self._closed:
                return
            self._flush_locked()
            self._closed = True
            self.append_sock.close(0)
            self.update_sock.close(2000)                      # let the last write-backs leave
        if self._ring is not None:
            self._ring.close()
            self._ring = None

    def __del__(self):
        try:
            self.flush()
            if self._ring is not None:
                self._ring.close()
        except Exception:
            pass


_CLIENTS = {}                    # meta address -> clients of this process (flushed before a loader asks for a group)


def _flush_clients(meta_addr):
    """-> {client id: [append messages, update frames]} sent by this process: a loader request carries it
    so that the server takes those writes in before it acts on the request."""
    after = {}
    for c in list(_CLIENTS.get(meta_addr, ())):
        after[c._id.hex()] = list(c.flush())
    return after


def _flush_timer(ref):
    """Bounds how long a partial frame of write-backs may wait: 2 ms."""
    while True:
        time.sleep(0.002)
        c = ref()
        if c is None or c._closed:
            return
        with c._lock:
            if c._held and time.perf_counter() - c._held_t0 > 0.002 and not c._closed:
                try:
                    c._flush_locked()
                except Exception:
                    return
        del c


class _RemoteLoader:
    """Loader-side connection: one DEALER socket.  With ahead >= 1 the request for group k+1 leaves
    when group k ARRIVES (so the server draws and gathers while this process consumes group k); it
    hands back group k-1, and groups k and k+1 may both be lent meanwhile.  With ahead == 0 (strict
    alternation) the request leaves when the next group is needed and hands back the previous one --
    the borrow rule of client/torch_cuda_loader.py:152-163 at group granularity."""

    def __init__(self, meta_addr, topic="default", prefetch=1, ahead=0, map_ring=True, timeout_s=120.0):
        import zmq
        from . import _cabi
        self.meta = get_meta(meta_addr)
        assert topic in self.meta["sampler_info"]
        self.topic = topic
        self._zmq = zmq
        self.sock = zmq.Context.instance().socket(zmq.DEALER)
        self.sock.setsockopt(zmq.LINGER, 0)
        self.sock.connect(self.meta["loader_addr"])
        self.timeout_ms = int(timeout_s * 1000)
        self._view = None
        self._lib = _cabi.lib()
        self._open = False
        self._map_ring = map_ring
        self.prefetch, self.ahead = int(prefetch), int(ahead)
        self.group = int(prefetch)
        self.batch_size = self.meta["batch_size"]
        self.device = self.meta["device"]
        self.struct = None
        self._queue = []
        self._asked = False          # a "next" request is outstanding
        self._ask_ahead = False
        self._stash = []             # "next" replies that arrived while another call was waiting

    def _ensure_open(self):
        """The row struct is only known after the first append (torch_cuda_loader.py:106-112 discovers
        it from the first row), so the ring is opened at the first sample(), not in the constructor."""
        if self._open:
            return
        from . import _cabi
        head, blobs = self._call({"op": "open", "topic": self.topic, "prefetch": self.prefetch, "ahead": self.ahead,
                                  "lend": 2 if self.ahead > 0 else 1})
        self._open = True
        self.struct = [(np.dtype(dt), tuple(sh)) for dt, sh in head["struct"]]
        self.group = head["group"]
        self._ask_ahead = head.get("lend", 1) >= 2     # else the ring is too small to lend two groups
        if self._map_ring:
            import torch
            torch.cuda.set_device(self.device)         # the ring is mapped for the current device
            ex = _cabi.IpcExport.from_buffer_copy(blobs[0])
            v = C.c_void_p()
            rc = self._lib.reth_ipc_open(C.byref(ex), C.byref(v))
            if rc != 0:
                _cabi.check(None, rc)
            self._view = v
            self._ncols = ex.n_columns

    def _send(self, obj):
        self.sock.send_multipart([b"", json.dumps(obj).encode()])

    def _recv(self, op):
        """The reply to request `op`; replies to an outstanding "next" that arrive first are kept."""
        if op == "next" and self._stash:
            return self._stash.pop(0)
        zmq = self._zmq
        while True:
            spin_until = time.perf_counter() + 0.0005          # a blocking poll costs a scheduler wake-up
            while True:
                try:
                    parts = self.sock.recv_multipart(zmq.NOBLOCK)
                    break
                except zmq.Again:
                    if time.perf_counter() < spin_until:
                        continue
                    if not self.sock.poll(self.timeout_ms):
                        raise TimeoutError("the replay server did not answer (no batch available: fewer than "
                                           "sample_start transitions were appended?)") from None
            head = json.loads(parts[1])
            if head.get("op") == "next" and op != "next" and head.get("ok"):
                self._stash.append((head, parts[2:]))
                continue
            if not head.get("ok"):
                if head.get("op") == "next":
                    self._asked = False
                raise RuntimeError(f"replay server: {head.get('error')}")
            return head, parts[2:]

    def _call(self, obj):
        self._send(obj)
        return self._recv(obj["op"])

    def _ask(self):
        # this process's appends and write-backs are taken in before the draw
        self._send({"op": "next", "topic": self.topic, "after": _flush_clients(self.meta["meta_addr"])})
        self._asked = True

    def _next(self):
        """-> (batch slot, indices int64[B], weights float64[B]) of the next pre-sampled batch."""
        if not self._queue:
            self._ensure_open()
            if not self._asked:
                self._ask()
            head, blobs = self._recv("next")
            self._asked = False
            if self._ask_ahead:
                self._ask()                            # the server works on the next group while this one is consumed
            G, B = head["n"], self.batch_size
            idx = np.frombuffer(blobs[0], np.int64).reshape(G, B).copy()      # writable, like the reference's
            w = np.frombuffer(blobs[1], np.float64).reshape(G, B).copy()
            first = head["first_slot"]
            self._queue = [(first + k, idx[k], w[k]) for k in range(G - 1, -1, -1)]
        return self._queue.pop()

    def slot_pointers(self, slot):
        cols = (C.c_void_p * max(1, self._ncols))()
        ip, wp = C.c_void_p(), C.c_void_p()
        rc = self._lib.reth_ipc_slot_ptrs(self._view, int(slot), cols, C.byref(ip), C.byref(wp))
        assert rc == 0
        return [cols[c] for c in range(self._ncols)], ip.value, wp.value

    def ring_read(self, column, first, n):
        """Parity / debug (reth_replay_ring_read): rows [first, first + n) of ring column `column` as stored."""
        head, blobs = self._call({"op": "ring", "column": int(column), "first": int(first), "n": int(n)})
        return np.frombuffer(blobs[0], np.dtype(head["dtype"])).reshape(head["shape"])

    def stat(self):
        """The server's counters as they stand (rows taken in, decode backlog, dropped messages); nothing is drained."""
        return self._call({"op": "stat"})[0]

    def drain(self):
        """Block until the server has executed everything it has received so far; -> {"cnt": ...}."""
        return self._call({"op": "drain", "after": _flush_clients(self.meta["meta_addr"])})[0]

    def close(self):
        if self._open:
            self._open = False
            try:
                self._call({"op": "close", "topic": self.topic})
            except Exception:
                pass
        if self._view is not None:
            self._lib.reth_ipc_close(self._view)
            self._view = None
        self.sock.close(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __iter__(self):
        return self

    def __next__(self):
        return self.sample()


class NumpyLoader(_RemoteLoader):
    """client/numpy_loader.py:8-56 against a server in another process: fresh host arrays."""

    def __init__(self, meta_addr, topic="default", prefetch=1, ahead=0):
        super().__init__(meta_addr, topic, prefetch, ahead, map_ring=False)

    def sample(self):
        slot, idx, w = self._next()
        _, blobs = self._call({"op": "rows", "slot": slot})
        B = self.batch_size
        data = [np.frombuffer(b, dt).reshape((B,) + sh).copy() for b, (dt, sh) in zip(blobs, self.struct)]
        return data, idx, w

    def get_indices(self):
        _, idx, w = self._next()
        return idx, w


class TorchCudaLoader(_RemoteLoader):
    """client/torch_cuda_loader.py:69-163 against a server in another process on the same GPU:
    the batch ring is mapped through CUDA IPC, data/weights are zero-copy CUDA tensors lent until
    the group is released, indices come back on the host."""

    def __init__(self, meta_addr, topic="default", buffer_size=8, num_procs=6, prefetch=None, ahead=None):
        # across processes every group costs a request/reply exchange (~100 us of host work on each
        # side), so the default group is the whole buffer_size and two more groups are kept in flight;
        # the server shrinks the group if its batch ring (start_server(max_ring=32)) cannot hold three
        buffer_size = max(1, int(buffer_size))
        if prefetch is None:
            prefetch = buffer_size
        if ahead is None:
            ahead = 2 if buffer_size > 1 else 0
        super().__init__(meta_addr, topic, prefetch, ahead, map_ring=True)
        self.buffer_size = buffer_size
        self.num_procs = num_procs
        self._views = {}

    def sample(self):
        import torch
        from .shard import _DevArray
        slot, idx, _ = self._next()
        v = self._views.get(slot)
        if v is None:
            cols, _, wp = self.slot_pointers(slot)
            B = self.batch_size
            dev = torch.device("cuda", self.device)
            data = [torch.as_tensor(_DevArray(p, (B,) + sh, dt, self), device=dev)
                    for p, (dt, sh) in zip(cols, self.struct)]
            weights = torch.as_tensor(_DevArray(wp, (B,), np.float64, self), device=dev)
            v = self._views[slot] = (data, weights)
        return v[0], idx, v[1]
