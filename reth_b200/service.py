"""The reth_buffer client/server surface on top of a ReplayShard.

Mirrors (same names, kwargs, return conventions, assertion behaviour):
  start_server / start_per   reth_buffer/reth_buffer/__init__.py:11-49
  Client                     reth_buffer/reth_buffer/client/client.py:8-39
  NumpyLoader                reth_buffer/reth_buffer/client/numpy_loader.py:8-56
  TorchCudaLoader            reth_buffer/reth_buffer/client/torch_cuda_loader.py:69-163
  sampler gating / ordering  reth_buffer/reth_buffer/server/sampler_loop.py:25-39

What replaces what: the reference's server is five OS processes glued by ZeroMQ and an LMDB
file in /dev/shm; here the "server" is one object owning an HBM-resident shard, reached by
address.  Addresses of the form  inproc://reth-b200/<n>  name an engine living in this
process (the fast path: calls go straight to the C ABI under a lock, which reproduces the
total order the reference gets from its single SUB socket, sampler_loop.py:30-36).  Row
storage, slot allocation, tree, sampling, IS weights and the batch gather all run on the GPU.

Staleness semantics are kept: batches are drawn ahead of their consumption, `prefetch` per
launch group and `ahead` groups beyond the one being consumed -- the analogue of the sampler's
HWM-1 PUSH socket + the loader worker processes + the loader ring (sampler_loop.py:14,
torch_cuda_loader.py:70, trainer.py:23) -- so a priority update lands after up to
(ahead + 1) * prefetch further batches were drawn.  prefetch=1, ahead=0 gives the strict
alternation of the in-process twin.  Priority write-backs are fire-and-forget (client.py:37-39).
"""
import itertools
import threading
import time

import numpy as np

from .sampler import FIFOPolicy, FIFOSampler, PERSampler, Schedule, UniformSampler  # noqa: F401  (re-exported like reth_buffer does)
from .shard import MIN_SERVICE, ReplayShard

_ENGINES = {}
_COUNTER = itertools.count()
_SCHEME = "inproc://reth-b200/"


class _Engine:
    """Server-side state of one replay buffer: shard + PER schedules + gating counters.

    Every mutation (append, update, draw) is enqueued on the shard's own ordering stream, in
    call order, under `lock`: the total order the reference gets from its single SUB socket
    (sampler_loop.py:30-36).  Nothing here waits for the device except a loader taking delivery
    of a pre-sampled group (ShardPipe.wait)."""

    def __init__(self, capacity, batch_size, samplers=None, cache_policy=None, device=0, max_ring=16,
                 float_frames=None, frames=None, nstep=None):
        # nstep = {"step": n, "gamma": g, "streams": 256, "r_col": 2, "s1_col": 3, "done_col": 4}: producers send their
        # RAW 1-step rows and the shard assembles the n-step transitions while it appends (reth_replay_append_nstep;
        # replaces NStepAdder on the actor, reth/reth/utils/nstep_adder.py:11-28, test/apex-dqn/worker.py:35,53).
        # Every Client is one environment stream, as every reference worker owns one adder.
        self.nstep = dict(nstep) if nstep else None
        self._nstep_streams = {}                              # producer id -> stream id, least recently used first
        self._nstep_open = False
        self.float_frames = float_frames                      # None | "f32" | "f32/255": see ReplayShard(out_cast=)
        # frames = {"k": stack depth, "shape": frame shape, "capacity": frames in the frame ring,
        #           "columns": (0, 3)}: frame-de-duplicated rows, fed through append_steps()
        self.frames = dict(frames) if frames else None
        self._writer = None
        self.capacity = int(capacity)
        self.batch_size = int(batch_size)
        self.device = device
        if samplers is None:                                  # server/main_loop.py:116-123
            samplers = [{"sampler_cls": PERSampler, "num_procs": 1, "sample_start": 1000}]
        assert len(samplers) >= 1
        topics = {}
        for item in samplers:
            topic = item.get("topic", "default")
            assert topic not in topics                        # main_loop.py:137
            kwargs = dict(item.get("kwargs", {}))
            cls = item["sampler_cls"]
            name = cls if isinstance(cls, str) else getattr(cls, "__name__", "")
            kinds = {"PERSampler": "per", "UniformSampler": "uniform", "FIFOSampler": "fifo"}
            if name not in kinds:
                raise NotImplementedError(f"sampler {name!r} has no device implementation in this build")
            topics[topic] = {
                "kind": kinds[name],
                "alpha": Schedule.from_str(kwargs.get("alpha", 0.6)),
                "beta": Schedule.from_str(kwargs.get("beta", 0.4)),
                "sample_start": max(item.get("sample_start", 1000), self.batch_size),   # main_loop.py:144-145
                "num_procs": item.get("num_procs", 1),
            }
        # Several topics share the row store; every sampler sees every append and update
        # (main_loop.py:58,79-88).  The first topic is served by the row shard's own tree and the
        # prefetch pipeline; each further topic gets a tree-only index shard over the same slots and
        # one reserved batch slot, and is served draw by draw (sample_topic).
        self.topics = topics
        self.topic, self.sampler = next(iter(topics.items()))
        self.kind = self.sampler["kind"]                      # "per" | "uniform" | "fifo"
        self.secondary = [t for t in topics if t != self.topic]
        # a custom cache policy keeps deciding slots on the host (BasePolicy.get_indices);
        # the default FIFO cursor lives in the shard
        self.cache_policy = cache_policy
        self.shard = None
        self.struct = None
        self.stream = None                                    # the shard's ordering stream
        self.cnt = 0                                          # sampler_loop.py:20,36
        self.lock = threading.RLock()
        self.closed = False
        self.max_ring = int(max_ring)
        # loader pipeline (opened by the first loader): groups of `group` batches, `ahead` groups
        # issued beyond the one being consumed
        self.pipe = None
        self.group = 1
        self.ahead = 0
        self.lend = 1                                         # delivered groups that stay lent (2: loader in another process)
        self._want = 0                                        # deliveries a remote loader has asked for
        self.loaders = 0
        self._queue = []                                      # delivered, not yet consumed: (slot, indices)
        # priority write-backs being coalesced into one staging entry
        self._upd_e = -1
        self._upd_k = 0
        self._upd_f32 = True

    # -- server side of Client.append (server/main_loop.py:44-58) -----------
    def _ensure_shard(self, data, struct=None):
        if self.shard is not None:
            return
        self.struct = struct or [(np.asarray(col).dtype, tuple(np.asarray(col).shape[1:])) for col in data]
        # float_frames: image-like uint8 columns (rank >= 2 per row) are handed to the learner as float32,
        # the conversion DQNSolver does first (reth/reth/algorithm/dqn/dqn_solver.py:71-75), fused into the gather
        cast = [self.float_frames if (np.dtype(dt) == np.uint8 and len(sh) >= 2) else None for dt, sh in self.struct]
        fr = None
        if self.frames:
            fr = {"shape": tuple(self.frames["shape"]), "capacity": int(self.frames["capacity"]),
                  "columns": list(self.frames.get("columns", (0, 3)))}
        self.shard = ReplayShard(self.capacity, self.struct, max_batch=self.batch_size,
                                 ring_depth=self.max_ring + len(self.secondary), device=self.device,
                                 min_mode=MIN_SERVICE, out_cast=cast if self.float_frames else None, frames=fr)
        self.stream = self.shard.pipe_stream
        self.shard.pipe_sampler(self.kind)
        if self.secondary:
            self.shard.pipe_reserve(len(self.secondary))
        for j, t in enumerate(self.secondary):
            st = self.topics[t]
            st["shard"] = ReplayShard(self.capacity, [], max_batch=self.batch_size, ring_depth=2, device=self.device,
                                      min_mode=MIN_SERVICE, store_rows=False)
            st["slot"] = self.max_ring + j                     # its batch slot in the row shard's ring

    def _notify_append(self, first, slots, weights):
        """The append notice of the other topics' samplers (main_loop.py:58 -> sampler.update)."""
        for t in self.secondary:
            st = self.topics[t]
            sh = st["shard"]
            if slots is None:
                sh.append([], weights, alpha=st["alpha"].value(), transform=True, stream=self.stream)
            else:
                sh.append_at(slots, [], weights, alpha=st["alpha"].value(), transform=True, stream=self.stream)
            if st["kind"] == "uniform":
                sh.uniform_note(slots, first_slot=first, n=len(weights), stream=self.stream)
            elif st["kind"] == "fifo":
                sh.fifo_push(slots, weights, first_slot=first, stream=self.stream)

    def _notify_update(self, indices, weights, step):
        for t in self.secondary:
            st = self.topics[t]
            if step:
                st["alpha"].step()
                st["beta"].step()
            sh = st["shard"]
            if st["kind"] == "per":
                sh.update(np.asarray(indices), np.asarray(weights), alpha=st["alpha"].value(), transform=True,
                          stream=self.stream)
            elif st["kind"] == "uniform":
                sh.uniform_note(np.asarray(indices), stream=self.stream)
            else:
                sh.fifo_push(np.asarray(indices), np.asarray(weights), stream=self.stream)

    def topic_ready(self, topic):
        st = self.topics[topic]
        if self.shard is None or self.cnt < st["sample_start"]:
            return False
        if st["kind"] == "fifo":
            return st["shard"].sampler_state()["fifo_len"] > self.batch_size
        return True

    def sample_topic(self, topic, timeout=60.0):
        """One batch of a topic other than the first: drawn by its index shard, rows gathered from the
        shared row store into the topic's batch slot (lent until the topic's next draw).
        -> (batch slot, host indices int64[B], host weights float64[B])."""
        t0 = time.time()
        while not self.topic_ready(topic):
            if self.closed:
                raise RuntimeError("replay buffer terminated")
            if time.time() - t0 > timeout:
                raise TimeoutError("no batch available: fewer than sample_start transitions were appended")
            time.sleep(0.001)
        with self.lock:
            self.flush()
            st = self.topics[topic]
            sh, B = st["shard"], self.batch_size
            if st["kind"] == "per":
                sh.sample(B, beta=st["beta"].value(), seed=0x70C1, slot=0, stream=self.stream)
            elif st["kind"] == "uniform":
                sh.sample_uniform(B, first_slot=0, n_slots=1, seed=0x70C1, stream=self.stream)
            else:
                sh.sample_fifo(B, first_slot=0, n_slots=1, stream=self.stream)
            self.shard.gather_from(st["slot"], sh, 0, stream=self.stream)
            _, idx, w = self.shard.slot_read(st["slot"], stream=self.stream)
            return st["slot"], idx, w

    def append_steps(self, stream_key, frames, others, weights, reset=None):
        """Frame-de-duplicated ingest (this build's extension; the reference stores the full k-stacks):
        `frames[i]` is the observation frame after step i of environment stream `stream_key`, `others`
        the non-frame columns of the n transitions, `reset` the first frame of an episode when the
        stream starts one.  Stacks are formed like FrameStack does (reth/reth/env/util.py:179-209)."""
        from .frames import FrameRefWriter
        with self.lock:
            if self.frames is None:
                raise ValueError("this buffer was not started with frames=...")
            k, shape = int(self.frames["k"]), tuple(self.frames["shape"])
            cols = tuple(self.frames.get("columns", (0, 3)))
            if self.shard is None:
                it = iter(others)
                struct = []
                for c in range(len(others) + len(cols)):
                    if c in cols:
                        struct.append((np.uint8, (k,) + shape))
                    else:
                        o = np.asarray(next(it))
                        struct.append((o.dtype, tuple(o.shape[1:])))
                self._ensure_shard(None, struct=struct)
            self.flush()                                      # staged write-backs first (call order, as in append)
            if self._writer is None:
                self._writer = FrameRefWriter(self.shard, k=k, frame_cols=cols)
            w = self._writer
            if reset is not None:
                w.reset(stream_key, reset)
            n = len(weights)
            frames = np.asarray(frames).reshape((n,) + shape) if n else frames
            others = [np.asarray(o) for o in others]
            for i in range(n):
                w.step(stream_key, frames[i], [o[i] for o in others], prio=float(weights[i]))
            w.alpha = self.sampler["alpha"].value()
            first = w.flush(stream=self.stream)
            if n:
                wts = np.asarray(weights, dtype=np.float32)
                if self.kind == "uniform":
                    self.shard.uniform_note(None, first_slot=first, n=n, stream=self.stream)
                elif self.kind == "fifo":
                    self.shard.fifo_push(None, wts, first_slot=first, stream=self.stream)
                self._notify_append(first, None, wts)
                self.cnt += n

    def append(self, data, weights, stream_key=None):
        with self.lock:
            self._ensure_shard(data)
            self.flush()                                      # staged write-backs land before the rows that may recycle their slots
            alpha = self.sampler["alpha"].value()
            if self.nstep is not None:
                if self.cache_policy is not None or self.secondary or self.kind != "per":
                    raise ValueError("n-step assembly serves one PER topic with the FIFO policy")
                ns = self.nstep
                if stream_key is None:
                    raise ValueError("n-step assembly needs messages that name their producer (this build's Client does; "
                                     "an unmodified reference actor keeps folding with its own NStepAdder)")
                if not self._nstep_open:
                    self._nstep_open = True
                    self.shard.nstep_open(int(ns.get("streams", 256)), int(ns["step"]), float(ns["gamma"]),
                                          r_col=int(ns.get("r_col", 2)), s1_col=int(ns.get("s1_col", 3)),
                                          done_col=int(ns.get("done_col", 4)), max_rows=int(ns.get("max_rows", 4096)))
                sid = self._nstep_streams.pop(stream_key, None)
                if sid is None:
                    if len(self._nstep_streams) < int(ns.get("streams", 256)):
                        sid = len(self._nstep_streams)
                    else:                                     # producers come and go (a restarted actor has a new id): the
                        oldest = next(iter(self._nstep_streams))   # least recently used stream starts over for the new one,
                        sid = self._nstep_streams.pop(oldest)      # its unfinished rows are dropped like a dead actor's
                        self.shard.nstep_reset(sid)
                self._nstep_streams[stream_key] = sid         # most recently used last
                _, emitted = self.shard.append_nstep(sid, list(data), weights, alpha=alpha, transform=True, stream=self.stream)
                self.cnt += emitted                           # rows that left the adder, as the reference counts stored rows
                return
            if self.cache_policy is None:
                first = self.shard.append(list(data), weights, alpha=alpha, transform=True, stream=self.stream)
                if self.kind == "uniform":                    # the append notice every sampler receives (main_loop.py:58)
                    self.shard.uniform_note(None, first_slot=first, n=len(weights), stream=self.stream)
                elif self.kind == "fifo":
                    self.shard.fifo_push(None, weights, first_slot=first, stream=self.stream)
                self._notify_append(first, None, weights)
            else:
                import torch
                slots = np.asarray(self.cache_policy.get_indices(len(weights)), dtype=np.int64)
                dev = [torch.from_numpy(np.ascontiguousarray(c, dtype=dt)).cuda(self.device)
                       for c, (dt, _) in zip(data, self.struct)]
                torch.cuda.current_stream(self.device).synchronize()
                self.shard.append_at(slots, dev, weights, alpha=alpha, transform=True, stream=self.stream)
                # the scatter reads `dev` on the engine's stream (not a torch stream): tell the caching allocator,
                # or the blocks could be handed to the next .cuda() copy while the kernel still reads them
                ext = torch.cuda.ExternalStream(self.stream, device=self.device)
                for t in dev:
                    t.record_stream(ext)
                if self.kind == "uniform":
                    self.shard.uniform_note(slots, stream=self.stream)
                elif self.kind == "fifo":
                    self.shard.fifo_push(slots, weights, stream=self.stream)
                self._notify_append(0, slots, weights)
            self.cnt += len(weights)

    # -- loader pipeline --------------------------------------------------
    def open_pipe(self, group, ahead, lend=1):
        """Called by a loader: `group` batches per launch group, `ahead` groups in flight beyond
        the one being consumed (ahead 0 = strict alternation: a draw sees every earlier update).
        lend = delivered groups that stay valid: 1 for a loader in this process; 2 for a loader in
        another process that asks for the next group when it STARTS on the current one (the request
        then hands back the group before)."""
        with self.lock:
            group = max(1, int(group))
            ahead = max(0, int(ahead))
            lend = 2 if (int(lend) >= 2 and ahead > 0) else 1
            if lend == 2 and max(ahead + 1, 3) * group > self.max_ring:
                lend = 1                                      # the batch ring holds fewer than 3 groups: lazy requests
            if self.pipe is not None:
                if (group, ahead, lend) != (self.group, self.ahead, self.lend):
                    if self.loaders > 0:
                        raise ValueError(f"this buffer already serves loaders with prefetch group {self.group} / "
                                         f"{self.ahead} ahead")
                    self._close_pipe()
            self.group, self.ahead, self.lend = group, ahead, lend
            self.loaders += 1

    def _close_pipe(self):
        if self.pipe is not None:
            self.flush()
            self.pipe.close()
            self.pipe = None
            self._queue = []
            self._want = 0

    def release_pipe(self):
        with self.lock:
            self.loaders = max(0, self.loaders - 1)
            if self.loaders == 0:
                self._close_pipe()

    def _ensure_pipe(self):
        if self.pipe is None:
            need = max(self.ahead + 1, self.lend + 1) * self.group
            if need > self.shard.ring_depth:
                raise ValueError(f"prefetch of {need} batches exceeds the batch ring ({self.shard.ring_depth})")
            self.pipe = self.shard.open_pipe(self.batch_size, self.group, seed=0x5EED)
            if self.lend > 1:
                self.pipe.hold(self.lend)
        return self.pipe

    # -- loader in another process (service_mp): request and delivery are separate, nothing blocks --
    def request_group(self):
        """A remote loader asked for one more group: hand back the oldest lent group once `lend`
        groups are out, keep `ahead` groups in flight, and remember that a delivery is owed."""
        with self.lock:
            pipe = self._ensure_pipe()
            self.flush()                                      # staged write-backs land before the draw
            beta = self.sampler["beta"].value()
            handed_back = False
            while pipe.in_flight < max(1, self.ahead):
                if pipe.can_issue():
                    pipe.issue(beta)
                elif pipe.waited >= pipe.lend and not handed_back:
                    pipe.advance(beta)                        # the request hands ONE group back: draw into its slots
                    handed_back = True
                else:
                    break
            self._want += 1

    def try_deliver(self):
        """-> (group slot g, host indices [group, batch], host weights [group, batch]) when a requested
        group is complete, else None.  Never blocks."""
        with self.lock:
            pipe = self.pipe
            if pipe is None or self._want <= 0:
                return None
            if pipe.in_flight == 0:                           # e.g. the ring was full when the request came
                if not pipe.can_issue():
                    return None
                pipe.issue(self.sampler["beta"].value())
            if not pipe.ready():
                return None
            g = pipe.wait()
            self._want -= 1
            return g, pipe.idx_host[g], pipe.w_host[g]

    def next_group(self):
        """Take delivery of the next pre-sampled group and keep `ahead` more in flight.
        -> (group slot g, host indices [group, batch], host weights [group, batch]); batch slots
        g*group .. g*group+group-1 stay lent until the next call."""
        self.wait_ready()
        with self.lock:
            pipe = self._ensure_pipe()
            self.flush()                                      # partially staged write-backs land before the draw
            beta = self.sampler["beta"].value()
            if pipe.in_flight == self.ahead and self.ahead > 0:
                g = pipe.next(beta)                           # steady state: take delivery + draw one more group
            else:
                if pipe.in_flight == 0:                       # first delivery, or strict alternation
                    pipe.issue(beta)
                g = pipe.wait()
                while pipe.in_flight < self.ahead and pipe.can_issue():
                    pipe.issue(beta)
            return g, pipe.idx_host[g], pipe.w_host[g]

    def next_batch(self):
        """-> (batch slot, host indices view) of the next pre-sampled batch."""
        if self._queue:
            return self._queue.pop()
        with self.lock:
            if self._queue:
                return self._queue.pop()
            g, idx, _ = self.next_group()
            G = self.group
            self._queue = [(g * G + k, idx[k]) for k in range(G - 1, -1, -1)]
            return self._queue.pop()

    # -- server side of Client.update_priorities (sampler_loop.py:29-36) --
    def update(self, indices, weights, step):
        """Fire-and-forget like the reference (client.py:37-39): the write-back is staged in pinned
        host memory and enqueued behind everything issued so far; `group` consecutive messages of
        batch_size share one staging entry = one host->device copy + one tree-update launch (the
        frames the reference keeps in its PUSH/SUB pipes).  Order and last-writer-wins are those of
        the message sequence; alpha is read when the entry is committed (with a scheduled alpha the up to
        `group` coalesced messages share the value at commit time: a deviation of at most group - 1 schedule
        steps from the reference's per-message read, per_sampler.py:34-35)."""
        with self.lock:
            if self.shard is None:
                return                                        # nothing stored yet: the reference updates an empty tree
            if step:                                          # sampler_loop.py:33-34, per_sampler.py:30-32
                self.sampler["alpha"].step()
                self.sampler["beta"].step()
            n = len(indices)
            self.cnt += n
            if self.secondary:
                self._notify_update(indices, weights, step)
            if self.kind != "per":                            # sampler.update() of the plugin (sampler_loop.py:35)
                if self.kind == "uniform":
                    self.shard.uniform_note(np.asarray(indices), stream=self.stream)
                else:
                    self.shard.fifo_push(np.asarray(indices), np.asarray(weights), stream=self.stream)
                return
            pipe = self.pipe
            B = self.batch_size
            w = weights if isinstance(weights, np.ndarray) else np.asarray(weights)
            if pipe is None:
                self.shard.update(np.asarray(indices), w, alpha=self.sampler["alpha"].value(), transform=True,
                                  stream=self.stream)
                return
            f32 = w.dtype == np.float32
            if n != B:                                        # not a learner batch: one launch per 16 384 indices
                self.flush()
                self.shard.update(np.asarray(indices), w, alpha=self.sampler["alpha"].value(), transform=True,
                                  stream=self.stream)
                return
            k = self._upd_k
            if k and f32 != self._upd_f32:
                self.flush()
                k = 0
            if k == 0:
                self._upd_e = pipe.update_begin()
                self._upd_f32 = f32
            e = self._upd_e
            pipe.upd_idx[e, k * B:(k + 1) * B] = indices
            (pipe.upd_p32 if f32 else pipe.upd_p64)[e, k * B:(k + 1) * B] = w
            self._upd_k = k + 1
            if k + 1 >= self.group:
                self.flush()

    def update_block(self, indices, weights, sizes, steps):
        """Several update_priorities messages that arrived in one frame (service_mp): indices / weights
        are the messages' arrays back to back.  Same effect as update() per message, in order; whole
        runs of learner-batch messages are staged with one copy each."""
        with self.lock:
            B, pipe = self.batch_size, self.pipe
            if (self.shard is None or self.kind != "per" or pipe is None or self.secondary or not all(steps)
                    or any(n != B for n in sizes)):
                at = 0
                for n, st in zip(sizes, steps):
                    self.update(indices[at:at + n], weights[at:at + n], bool(st))
                    at += n
                return
            f32 = weights.dtype == np.float32
            alpha, beta = self.sampler["alpha"], self.sampler["beta"]
            K, pos = len(sizes), 0
            while pos < K:
                k = self._upd_k
                if k and f32 != self._upd_f32:
                    self.flush()
                    k = 0
                if k == 0:
                    self._upd_e = pipe.update_begin()
                    self._upd_f32 = f32
                take = min(self.group - k, K - pos)
                for _ in range(take):                         # sampler_loop.py:33-34, once per message
                    alpha.step()
                    beta.step()
                e = self._upd_e
                pipe.upd_idx[e, k * B:(k + take) * B] = indices[pos * B:(pos + take) * B]
                (pipe.upd_p32 if f32 else pipe.upd_p64)[e, k * B:(k + take) * B] = weights[pos * B:(pos + take) * B]
                self._upd_k = k + take
                self.cnt += take * B
                pos += take
                if self._upd_k >= self.group:
                    self.flush()

    def flush(self):
        with self.lock:
            if self._upd_k and self.pipe is not None:
                k, self._upd_k = self._upd_k, 0
                self.pipe.update_commit(self._upd_e, k * self.batch_size, self._upd_f32,
                                        self.sampler["alpha"].value())

    def drain(self):
        """Block until every enqueued append / update / draw has executed (tests, benchmarks)."""
        with self.lock:
            self.flush()
            if self.pipe is not None:
                self.pipe.drain()
            elif self.shard is not None:
                self.shard._ck(self.shard._lib.reth_replay_pipe_drain(self.shard._h))

    def ready(self):
        if self.shard is None or self.cnt < self.sampler["sample_start"]:
            return False
        if self.kind == "fifo":                               # FIFOSampler.ready_sample (fifo_sampler.py:17-18)
            return self.shard.sampler_state()["fifo_len"] > self.batch_size * self.group
        return True

    def wait_ready(self, timeout=60.0):
        if self.ready():
            return
        t0 = time.time()
        while not self.ready():
            if self.closed:
                raise RuntimeError("replay buffer terminated")
            if time.time() - t0 > timeout:
                raise TimeoutError("no batch available: fewer than sample_start transitions were appended")
            time.sleep(0.001)

    def close(self):
        with self.lock:
            self.closed = True
            if self.pipe is not None:
                self.pipe.close()
                self.pipe = None
            for t in self.secondary:
                sh = self.topics[t].pop("shard", None)
                if sh is not None:
                    sh.close()
            if self.shard is not None:
                self.shard.close()
                self.shard = None


class _ServerHandle:
    """What start_server returns as `proc`: callers only use terminate()/join()
    (test/apex-dqn/trainer.py:72-74, reth_buffer/test/test_rb.py:34-35)."""

    def __init__(self, addr):
        self.addr = addr

    def terminate(self):
        eng = _ENGINES.pop(self.addr, None)
        if eng is not None:
            eng.close()

    def join(self, timeout=None):
        return None

    def is_alive(self):
        return self.addr in _ENGINES

    kill = terminate


def start_server(capacity, batch_size, host=None, port=None, samplers=None, cache_policy=None, device=0,
                 max_ring=16, float_frames=None, frames=None, nstep=None):
    """reth_buffer.start_server (__init__.py:11-27).  host/port are accepted for signature
    compatibility; the engine is reached through the returned in-process address."""
    addr = f"{_SCHEME}{next(_COUNTER)}"
    _ENGINES[addr] = _Engine(capacity, batch_size, samplers, cache_policy, device, max_ring, float_frames, frames, nstep)
    return _ServerHandle(addr), addr


def start_per(capacity, batch_size, alpha=0.6, beta=0.4, sample_start=1000, num_sampler_procs=1,
              host=None, port=None, cache_policy=None, device=0, max_ring=16, float_frames=None, frames=None, nstep=None):
    """reth_buffer.start_per (__init__.py:30-49)."""
    samplers = [{"sampler_cls": PERSampler, "num_procs": num_sampler_procs, "sample_start": sample_start,
                 "kwargs": {"alpha": alpha, "beta": beta}}]
    return start_server(capacity, batch_size, host, port, samplers, cache_policy, device, max_ring, float_frames, frames, nstep)


def _engine(meta_addr):
    try:
        return _ENGINES[meta_addr]
    except KeyError:
        raise ConnectionError(f"no replay engine at {meta_addr!r} in this process") from None


def _remote(meta_addr):
    return isinstance(meta_addr, str) and not meta_addr.startswith(_SCHEME)


class Client:
    """client/client.py:8-39.  An address that does not name an in-process engine is a server in
    another process (service_mp): the same class name then builds the ZeroMQ client."""

    def __new__(cls, meta_addr, *args, **kwargs):
        if cls is Client and _remote(meta_addr):
            from . import service_mp
            return service_mp.Client(meta_addr, *args, **kwargs)
        return super().__new__(cls)

    def __init__(self, meta_addr):
        self._eng = _engine(meta_addr)
        self.meta = {"capacity": self._eng.capacity, "batch_size": self._eng.batch_size, "meta_addr": meta_addr}
        self._id = next(_COUNTER)                             # producer id (n-step stream key); id(self) could be reused

    def append(self, data, weights, compress=False):
        assert isinstance(data, (list, tuple))                # client.py:22
        for col in data:
            assert isinstance(col, np.ndarray)                # :24
            assert len(col) == len(weights)                   # :25
        self._eng.append(data, np.asarray(weights), stream_key=self._id)

    def append_steps(self, env, frames, others, weights, reset=None):
        """Frame-de-duplicated ingest for a buffer started with frames=...: see _Engine.append_steps."""
        assert len(frames) == len(weights) and all(len(o) == len(weights) for o in others)
        self._eng.append_steps((self._id, env), frames, others, np.asarray(weights), reset)

    def update_priorities(self, indices, weights, step=True):
        assert len(indices) == len(weights)                   # :38
        self._eng.update(indices, weights, step)


class _LoaderBase:
    """prefetch = batches per launch group, ahead = groups in flight beyond the one being consumed;
    (ahead + 1) * prefetch batches are resident in the lending ring (the reference's buffer_size)."""

    def __init__(self, meta_addr, topic="default", prefetch=1, ahead=0):
        self._eng = _engine(meta_addr)
        assert topic in self._eng.topics
        self.topic = topic
        self._primary = topic == self._eng.topic             # other topics are served draw by draw
        self.meta = {"capacity": self._eng.capacity, "batch_size": self._eng.batch_size}
        self.prefetch = max(1, int(prefetch))
        self.ahead = max(0, int(ahead))
        if self._primary:
            self._eng.open_pipe(self.prefetch, self.ahead)
        self._open = True

    def __iter__(self):
        return self

    def __next__(self):
        return self.sample()

    def _next(self):
        """-> (batch slot, indices, weights) on the host."""
        if not self._primary:
            return self._eng.sample_topic(self.topic)
        slot, idx = self._eng.next_batch()
        pipe = self._eng.pipe
        return slot, np.array(idx), np.array(pipe.w_host[slot // pipe.group, slot % pipe.group])

    def get_indices(self):
        """numpy_loader.py:53-56: (indices, weights) of the next batch."""
        _, idx, w = self._next()
        return idx, w

    def close(self):
        if self._open:
            self._open = False
            if self._primary:
                self._eng.release_pipe()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class NumpyLoader(_LoaderBase):
    """client/numpy_loader.py:8-56: fresh host arrays (data, indices int64, weights float64)."""

    def __new__(cls, meta_addr, *args, **kwargs):
        if cls is NumpyLoader and _remote(meta_addr):
            from . import service_mp
            return service_mp.NumpyLoader(meta_addr, *args, **kwargs)
        return super().__new__(cls)

    def __init__(self, meta_addr, topic="default", prefetch=1, ahead=0):
        super().__init__(meta_addr, topic, prefetch, ahead)

    def sample(self):
        eng = self._eng
        slot, idx, w = self._next()
        with eng.lock:
            data, _, _ = eng.shard.slot_read(slot, with_rows=True, stream=eng.stream)
            return data, idx, w


def loader_geometry(buffer_size, num_procs, max_ring):
    """(prefetch group, groups ahead) for a reference-style TorchCudaLoader(buffer_size, num_procs).
    The reference keeps up to buffer_size + num_procs + 2 pre-sampled batches between a draw and its
    consumption: `buffer_size` ring slots (torch_cuda_loader.py:70,114-135), one batch in each of the
    `num_procs` worker processes (:20-66), one in the sampler's HWM-1 PUSH socket and one being drawn
    (server/sampler_loop.py:14,37-39) -- 16 at its defaults (8, 6), 8 as the Ape-X trainer calls it
    (trainer.py:22-24: 4, 2).  The same budget here: half of it is the group being consumed, the other
    half is drawn + gathered ahead."""
    budget = max(2, min(int(buffer_size) + max(0, int(num_procs)) + 2, int(max_ring)))
    prefetch = max(1, budget // 2)
    return prefetch, (1 if budget >= 2 else 0)


class TorchCudaLoader(_LoaderBase):
    """client/torch_cuda_loader.py:69-163: data = list of CUDA tensors lent until the next
    sample() (:152-163), indices on the host (np.int64), weights a CUDA float64 tensor.
    buffer_size / num_procs set the staleness budget as in the reference (loader_geometry): by default
    half of the batches it allows in flight form the group being consumed and the other half is drawn +
    gathered ahead.  prefetch=1, ahead=0 is strict alternation."""

    def __new__(cls, meta_addr, *args, **kwargs):
        if cls is TorchCudaLoader and _remote(meta_addr):
            from . import service_mp
            return service_mp.TorchCudaLoader(meta_addr, *args, **kwargs)
        return super().__new__(cls)

    def __init__(self, meta_addr, topic="default", buffer_size=8, num_procs=6, prefetch=None, ahead=None):
        eng = _engine(meta_addr)
        buffer_size = max(1, min(int(buffer_size), eng.max_ring))
        if prefetch is None and ahead is None:
            prefetch, ahead = loader_geometry(buffer_size, num_procs, eng.max_ring)
        if prefetch is None:
            prefetch = max(1, buffer_size // 2)
        if ahead is None:
            ahead = max(0, buffer_size // int(prefetch) - 1)
        ahead = min(int(ahead), max(0, eng.max_ring // int(prefetch) - 1))
        super().__init__(meta_addr, topic, prefetch, ahead)
        self.buffer_size = buffer_size
        self.num_procs = num_procs                            # no gather processes are needed; counts towards the staleness budget
        self._views = {}                                      # slot -> cached zero-copy torch views

    def sample(self):
        if self._primary:
            slot, idx = self._eng.next_batch()
        else:
            slot, idx, _ = self._eng.sample_topic(self.topic)
        v = self._views.get(slot)
        if v is None:
            data, _, weights, _ = self._eng.shard.slot_tensors(slot)
            v = self._views[slot] = (data, weights)
        return v[0], np.array(idx), v[1]
