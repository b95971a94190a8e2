"""In-process twin: the method set of reth.buffer.PrioritizedBuffer
(reth/reth/buffer/prioritized_buffer.py:8-74) on a GPU ReplayShard.

Differences of the reference twin from the service that are kept here:
  * append(data, weight) stores the raw weight, untransformed (:29-33);
  * append_batch(data, weights=None) stores raw 1.0 (:36-37), else (w+1e-6)^alpha (:39);
  * alpha/beta advance once per sample(), after the indices are drawn and before the
    importance weights are computed (:46-50);
  * min() is `_min[0] if _min[0] != 0 else 1` (reth/reth/buffer/sumtree.py:112-113).
The row struct is detected from the first transition (reth/reth/buffer/buffer.py:25-33).
"""
import numpy as np

from .sampler import Schedule
from .shard import MIN_INPROCESS, ReplayShard


class PrioritizedBuffer:
    def __init__(self, capacity=50000, alpha=0.6, beta=0.4, struct=None, max_batch=1024, device=0,
                 device_output=False):
        self.alpha = Schedule.from_str(alpha)
        self._alpha_str = alpha
        self.beta = Schedule.from_str(beta)
        self._beta_str = beta
        self._eps = 1e-6
        self._capacity = int(capacity)
        self._struct = None
        self._max_batch = int(max_batch)
        self._device = device
        self._device_output = device_output
        self._shard = None
        self._slot = 0
        if struct is not None:
            self._create([(np.dtype(dt), tuple(sh)) for dt, sh in struct])

    def _create(self, struct):
        self._struct = [(np.dtype(dt).name, tuple(sh)) for dt, sh in struct]
        self._shard = ReplayShard(self._capacity, struct, max_batch=self._max_batch, ring_depth=2,
                                  device=self._device, min_mode=MIN_INPROCESS)

    def _detect(self, row):
        try:
            cols = [np.asarray(c) for c in row]
        except Exception as e:                                # buffer.py:31-32
            raise Exception("Invalid transition data") from e
        self._create([(c.dtype, c.shape) for c in cols])

    def append(self, data, weight=None):
        if weight is None:
            weight = 1
        if self._shard is None:
            self._detect(data)
        cols = [np.asarray(c, dtype=dt)[None] for c, (dt, _) in zip(data, self._shard.struct)]
        self._shard.append(cols, np.asarray([weight], dtype=np.float64), transform=False)

    def append_batch(self, data, weights=None):
        if self._shard is None:
            self._detect([col[0] for col in data])            # buffer.py:58-60
        n = len(data[0])
        assert n <= self.capacity                             # buffer.py:66
        if weights is None:
            self._shard.append(list(data), None)
        else:
            assert len(weights) == n                          # prioritized_buffer.py:38
            self._shard.append(list(data), np.asarray(weights), alpha=self.alpha.value(), transform=True)

    def sample(self, batch_size, uniforms=None):
        assert batch_size <= self.size                        # prioritized_buffer.py:45
        if batch_size > self._max_batch:
            raise ValueError(f"batch_size {batch_size} > max_batch {self._max_batch} given at construction")
        self.alpha.step()
        beta = self.beta.step()                               # indices do not depend on beta
        slot = self._slot
        self._slot ^= 1
        self._shard.sample(batch_size, beta=beta, uniforms=uniforms, slot=slot)
        self._shard.gather(slot)
        if self._device_output:
            data, idx, w, _ = self._shard.slot_tensors(slot)
            return data, idx, w
        return self._shard.slot_read(slot, with_rows=True)

    def update_priorities(self, indices, weights):
        self._shard.update(np.asarray(indices), np.asarray(weights), alpha=self.alpha.value(), transform=True)

    def clear(self):
        self.alpha = Schedule.from_str(self._alpha_str)
        self.beta = Schedule.from_str(self._beta_str)
        if self._shard is not None:
            self._shard.clear()

    @property
    def capacity(self):
        return self._capacity

    @property
    def size(self):
        return 0 if self._shard is None else self._shard.size

    @property
    def struct(self):
        return self._struct
