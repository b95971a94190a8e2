"""Host-side scalar logic of the sampler plugin surface.

Schedule  <- reth_buffer/reth_buffer/utils/schedule.py:4-52 (constant, "start,end,steps"
             linear, "exp,start,end,steps"); advanced once per update_priorities(step=True)
             in the service (sampler/per_sampler.py:30-32) and once per sample() in the
             in-process twin (reth/reth/buffer/prioritized_buffer.py:48-49).
FIFOPolicy<- cache_policy/fifo_policy.py:6-18 (host mirror of the cursor the shard keeps).
PERSampler<- sampler/per_sampler.py:5-35 behind BaseSampler (sampler/base_sampler.py:4-21),
             backed by a ReplayShard (tree-only) instead of NumbaSumTree.
"""
import math
from abc import ABC, abstractmethod

import numpy as np


class Schedule:
    def __init__(self, func, max_steps=1):
        self.func = func
        self.max_steps = max_steps
        self.cur_step = 0

    @staticmethod
    def from_str(spec):
        if isinstance(spec, (int, float)):
            return Schedule(lambda _step: spec)
        fields = spec.split(",")
        if len(fields) == 3:
            kind, (lo, hi, steps) = "linear", fields
        elif len(fields) == 4:
            kind, lo, hi, steps = fields
        else:
            raise Exception(f"Invalid schedule string {spec}")
        lo, hi, steps = float(lo), float(hi), int(steps)
        if kind == "linear":
            return Schedule(lambda step: lo + (hi - lo) * step / steps, steps)
        if kind == "exp":
            return Schedule(lambda step: hi - (hi - lo) * math.exp(-1 * step / steps), steps)
        raise Exception(f"Invalid schedule method {kind}")

    def step(self):
        if self.cur_step < self.max_steps:
            self.cur_step += 1
        return self.func(self.cur_step)

    def value(self, step=None):
        step = self.cur_step if step is None else min(step, self.max_steps)
        return self.func(step)


class BasePolicy(ABC):
    @abstractmethod
    def get_indices(self, batch_size):
        ...


class FIFOPolicy(BasePolicy):
    def __init__(self, capacity):
        self.capacity = capacity
        self.tail = 0

    def get_indices(self, batch_size):
        assert batch_size <= self.capacity
        res = ((self.tail + np.arange(batch_size, dtype=np.int64)) % self.capacity).astype("i4")
        self.tail = int((self.tail + batch_size) % self.capacity)
        return res


class BaseSampler(ABC):
    def ready_sample(self, batch_size):
        return True

    def on_step(self):
        pass

    @abstractmethod
    def sample(self, batch_size):
        ...

    @abstractmethod
    def update(self, indices, weights):
        ...

    @abstractmethod
    def clear(self):
        ...


class PERSampler(BaseSampler):
    """GPU PERSampler plugin: same constructor/methods as sampler/per_sampler.py:5-35.

    sample() returns host (indices int64, weights float64) like the reference; the
    device copies stay in batch slot `self.last_slot` of `self.shard`.
    """

    def __init__(self, capacity, alpha=0.6, beta=0.4, shard=None, max_batch=8192, device=0, seed=0):
        from .shard import ReplayShard
        self.shard = shard if shard is not None else ReplayShard(
            capacity, [], max_batch=max_batch, store_rows=False, device=device)
        self.alpha = Schedule.from_str(alpha)
        self._alpha_str = alpha
        self.beta = Schedule.from_str(beta)
        self._beta_str = beta
        self._eps = 1e-6
        self.seed = seed
        self.last_slot = 0
        self._next_slot = 0

    def clear(self):
        self.shard.clear()
        self.alpha = Schedule.from_str(self._alpha_str)
        self.beta = Schedule.from_str(self._beta_str)

    def sample(self, batch_size, uniforms=None):
        slot = self._next_slot
        self._next_slot = (slot + 1) % self.shard.ring_depth
        self.shard.sample(batch_size, beta=self.beta.value(), uniforms=uniforms, seed=self.seed, slot=slot)
        self.last_slot = slot
        _, indices, weights = self.shard.slot_read(slot)
        return indices, weights

    def on_step(self):
        self.alpha.step()
        self.beta.step()

    def update(self, indices, weights):
        self.shard.update(indices, weights, alpha=self.alpha.value(), transform=True)


class UniformSampler(BaseSampler):
    """sampler/uniform_sampler.py:6-25 on a ReplayShard: the index log and the draw live on the
    device; sample() returns host (indices int64, weights int64 ones) like the reference."""

    def __init__(self, capacity, shard=None, max_batch=8192, device=0, seed=0):
        from .shard import ReplayShard
        self.shard = shard if shard is not None else ReplayShard(capacity, [], max_batch=max_batch,
                                                                 store_rows=False, device=device)
        self.capacity = capacity
        self.seed = seed
        self.last_slot = 0

    @property
    def tail(self):
        return self.shard.sampler_state()["uniform_tail"]

    def sample(self, batch_size, draws=None):
        slot = self.last_slot = (self.last_slot + 1) % self.shard.ring_depth
        self.shard.sample_uniform(batch_size, slot, 1, draws=draws, seed=self.seed)
        _, indices, _ = self.shard.slot_read(slot)
        return indices, np.ones(batch_size, dtype="i8")

    def update(self, indices, weights):
        tail = self.tail
        self.shard.uniform_note(indices)                 # logs up to capacity, a full log ignores the message (:18-19)
        if tail < self.capacity < tail + len(indices):
            # the reference writes element by element and runs off the end of its log when a message CROSSES the
            # capacity (uniform_sampler.py:20-22: indices[capacity] = idx): same state (log filled), same exception
            raise IndexError(f"index {self.capacity} is out of bounds for axis 0 with size {self.capacity}")

    def clear(self):
        self.shard.clear()


class FIFOSampler(BaseSampler):
    """sampler/fifo_sampler.py:8-29 on a ReplayShard (an on-policy queue: every entry is handed
    out once, oldest first)."""

    def __init__(self, capacity, shard=None, max_batch=8192, device=0):
        from .shard import ReplayShard
        self.shard = shard if shard is not None else ReplayShard(capacity, [], max_batch=max_batch,
                                                                 store_rows=False, device=device)
        self.last_slot = 0

    def __len__(self):
        return self.shard.sampler_state()["fifo_len"]

    def ready_sample(self, batch_size):
        return len(self) > batch_size

    def sample(self, batch_size):
        slot = self.last_slot = (self.last_slot + 1) % self.shard.ring_depth
        self.shard.sample_fifo(batch_size, slot, 1)
        _, indices, weights = self.shard.slot_read(slot)
        return indices, weights

    def update(self, indices, weights):
        self.shard.fifo_push(indices, weights)

    def clear(self):
        self.shard.clear()
