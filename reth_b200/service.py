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

Staleness semantics are kept: a loader pre-samples `prefetch` batches per launch group, the
analogue of sampler HWM 1 + the loader ring (torch_cuda_loader.py:70, trainer.py:23), so a
priority update lands after up to `prefetch` further batches were drawn.  prefetch=1 gives the
strict alternation of the in-process twin.
"""
import itertools
import threading
import time

import numpy as np

from .sampler import FIFOPolicy, PERSampler, Schedule  # noqa: F401  (re-exported like reth_buffer does)
from .shard import MIN_SERVICE, ReplayShard

_ENGINES = {}
_COUNTER = itertools.count()
_SCHEME = "inproc://reth-b200/"


class _Engine:
    """Server-side state of one replay buffer: shard + PER schedules + gating counters."""

    def __init__(self, capacity, batch_size, samplers=None, cache_policy=None, device=0):
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
            if cls is not PERSampler and getattr(cls, "__name__", "") != "PERSampler":
                raise NotImplementedError("only PERSampler runs on the device in this build (SURVEY.md 8f row 3)")
            topics[topic] = {
                "alpha": Schedule.from_str(kwargs.get("alpha", 0.6)),
                "beta": Schedule.from_str(kwargs.get("beta", 0.4)),
                "sample_start": max(item.get("sample_start", 1000), self.batch_size),   # main_loop.py:144-145
                "num_procs": item.get("num_procs", 1),
            }
        if len(topics) != 1:
            raise NotImplementedError("one sampler topic per buffer in this build")
        self.topic, self.sampler = next(iter(topics.items()))
        # a custom cache policy keeps deciding slots on the host (BasePolicy.get_indices);
        # the default FIFO cursor lives in the shard
        self.cache_policy = cache_policy
        self.shard = None
        self.struct = None
        self.cnt = 0                                          # sampler_loop.py:20,36
        self.lock = threading.RLock()
        self._pending = []
        self.closed = False
        self.max_ring = 16

    # -- server side of Client.append (server/main_loop.py:44-58) -----------
    def _ensure_shard(self, data, struct=None):
        if self.shard is not None:
            return
        self.struct = struct or [(np.asarray(col).dtype, tuple(np.asarray(col).shape[1:])) for col in data]
        self.shard = ReplayShard(self.capacity, self.struct, max_batch=self.batch_size,
                                 ring_depth=self.max_ring, device=self.device, min_mode=MIN_SERVICE)

    def append(self, data, weights):
        with self.lock:
            self._ensure_shard(data)
            alpha = self.sampler["alpha"].value()
            if self.cache_policy is None:
                self.shard.append(list(data), weights, alpha=alpha, transform=True)
            else:
                import torch
                slots = np.asarray(self.cache_policy.get_indices(len(weights)), dtype=np.int64)
                dev = [torch.from_numpy(np.ascontiguousarray(c, dtype=dt)).cuda(self.device)
                       for c, (dt, _) in zip(data, self.struct)]
                self.shard.append_at(slots, dev, weights, alpha=alpha, transform=True)
            self.cnt += len(weights)

    def update(self, indices, weights, step, coalesce=1):
        """sampler_loop.py:29-36.  coalesce > 1: updates are queued like the frames the reference
        keeps in its PUSH/SUB pipes (fire-and-forget, client.py:37-39) and written back in one
        launch once `coalesce` are pending or the sampler is about to draw (flush()); order and
        last-writer-wins are those of the message sequence.  alpha is constant over a run here
        when coalescing (the schedule still advances once per message)."""
        with self.lock:
            if self.shard is None:
                return                                        # nothing stored yet: the reference drops into an empty tree
            if step:                                          # sampler_loop.py:33-34, per_sampler.py:30-32
                self.sampler["alpha"].step()
                self.sampler["beta"].step()
            self.cnt += len(indices)
            if coalesce <= 1 and not self._pending:
                self.shard.update(indices, weights, alpha=self.sampler["alpha"].value(), transform=True)
                return
            self._pending.append((np.asarray(indices, dtype=np.int64), np.asarray(weights)))
            if len(self._pending) >= coalesce:
                self.flush()

    def flush(self):
        with self.lock:
            if not self._pending or self.shard is None:
                return
            idx = np.concatenate([p[0] for p in self._pending])
            dts = {p[1].dtype for p in self._pending}
            w = np.concatenate([p[1] if len(dts) == 1 else p[1].astype(np.float64) for p in self._pending])
            self._pending = []
            self.shard.update(idx, w, alpha=self.sampler["alpha"].value(), transform=True)

    def ready(self):
        return self.shard is not None and self.cnt >= self.sampler["sample_start"]

    def wait_ready(self, timeout=60.0):
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


def start_server(capacity, batch_size, host=None, port=None, samplers=None, cache_policy=None, device=0):
    """reth_buffer.start_server (__init__.py:11-27).  host/port are accepted for signature
    compatibility; the engine is reached through the returned in-process address."""
    addr = f"{_SCHEME}{next(_COUNTER)}"
    _ENGINES[addr] = _Engine(capacity, batch_size, samplers, cache_policy, device)
    return _ServerHandle(addr), addr


def start_per(capacity, batch_size, alpha=0.6, beta=0.4, sample_start=1000, num_sampler_procs=1,
              host=None, port=None, cache_policy=None, device=0):
    """reth_buffer.start_per (__init__.py:30-49)."""
    samplers = [{"sampler_cls": PERSampler, "num_procs": num_sampler_procs, "sample_start": sample_start,
                 "kwargs": {"alpha": alpha, "beta": beta}}]
    return start_server(capacity, batch_size, host, port, samplers, cache_policy, device)


def _engine(meta_addr):
    try:
        return _ENGINES[meta_addr]
    except KeyError:
        raise ConnectionError(f"no replay engine at {meta_addr!r} in this process") from None


class Client:
    """client/client.py:8-39."""

    def __init__(self, meta_addr, coalesce_updates=1):
        self._eng = _engine(meta_addr)
        self.coalesce_updates = int(coalesce_updates)
        self.meta = {"capacity": self._eng.capacity, "batch_size": self._eng.batch_size, "meta_addr": meta_addr}

    def append(self, data, weights, compress=False):
        assert isinstance(data, (list, tuple))                # client.py:22
        for col in data:
            assert isinstance(col, np.ndarray)                # :24
            assert len(col) == len(weights)                   # :25
        self._eng.append(data, np.asarray(weights))

    def update_priorities(self, indices, weights, step=True):
        assert len(indices) == len(weights)                   # :38
        self._eng.update(np.asarray(indices), np.asarray(weights), step, self.coalesce_updates)


class _LoaderBase:
    def __init__(self, meta_addr, topic="default", prefetch=1):
        self._eng = _engine(meta_addr)
        assert topic == self._eng.topic
        self.meta = {"capacity": self._eng.capacity, "batch_size": self._eng.batch_size}
        self.prefetch = max(1, int(prefetch))
        self._queue = []          # (slot, indices) of pre-sampled, not yet delivered batches
        self._next_group = 0

    def __iter__(self):
        return self

    def __next__(self):
        return self.sample()

    def _refill(self):
        eng = self._eng
        eng.wait_ready()
        with eng.lock:
            eng.flush()                                       # pending priority messages land before the draw
            sh, B, G = eng.shard, eng.batch_size, self.prefetch
            depth = sh.ring_depth
            first = self._next_group
            if first + G > depth:
                first = 0
            self._next_group = first + G if first + 2 * G <= depth else 0
            beta = eng.sampler["beta"].value()
            if G == 1:
                sh.sample(B, beta=beta, seed=0x5EED, slot=first)
                sh.gather(first)
            else:
                sh.sample_multi(B, first, G, beta=beta, seed=0x5EED)
                sh.gather_multi(first, G)
            if G == 1:
                idx, _ = sh.group_read(first, 1)
                self._queue.append((first, idx))
            else:                                             # one D2H of all indices, one sync per group
                idx, _ = sh.group_read(first, G)
                for g in range(G):
                    self._queue.append((first + g, idx[g * B:(g + 1) * B]))

    def get_indices(self):
        """numpy_loader.py:53-56: (indices, weights) of the next batch."""
        if not self._queue:
            self._refill()
        slot, idx = self._queue.pop(0)
        _, _, w = self._eng.shard.slot_read(slot)
        return idx, w

    def close(self):
        self._queue = []


class NumpyLoader(_LoaderBase):
    """client/numpy_loader.py:8-56: fresh host arrays (data, indices int64, weights float64)."""

    def __init__(self, meta_addr, topic="default", prefetch=1):
        super().__init__(meta_addr, topic, prefetch)

    def sample(self):
        if not self._queue:
            self._refill()
        slot, _ = self._queue.pop(0)
        with self._eng.lock:
            data, indices, weights = self._eng.shard.slot_read(slot, with_rows=True)
        return data, indices, weights


class TorchCudaLoader(_LoaderBase):
    """client/torch_cuda_loader.py:69-163: data = list of CUDA tensors lent until the next
    sample() (:152-163), indices on the host (np.int64), weights a CUDA float64 tensor.
    buffer_size = depth of the lending ring; prefetch defaults to buffer_size // 2 batches drawn
    per launch group (the reference keeps buffer_size batches in flight)."""

    def __init__(self, meta_addr, topic="default", buffer_size=8, num_procs=6, prefetch=None):
        eng = _engine(meta_addr)
        buffer_size = max(2, min(int(buffer_size), eng.max_ring))
        g = max(1, buffer_size // 2) if prefetch is None else int(prefetch)
        super().__init__(meta_addr, topic, min(g, eng.max_ring // 2))
        self.buffer_size = buffer_size
        self.num_procs = num_procs                            # accepted, unused: no gather processes are needed
        self._views = {}                                      # slot -> cached zero-copy torch views

    def sample(self):
        if not self._queue:
            self._refill()
        slot, indices = self._queue.pop(0)
        v = self._views.get(slot)
        if v is None:
            data, _, weights, _ = self._eng.shard.slot_tensors(slot)
            v = self._views[slot] = (data, weights)
        return v[0], indices, v[1]
