"""Host logic of the frame-de-duplicated layout and the n-step oracle: the closed form of NStepAdder against
tapes recorded from the reference's own class (tests/golden/make_nstep_golden.py), FrameRefWriter's stacking rule against
FrameStack's (reth/reth/env/util.py:179-209) with a stand-in shard."""
import os
from collections import deque

import numpy as np
import pytest

from oracle import per_oracle as po
from reth_b200.frames import FrameRefWriter

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "nstep_adder.npz"))
TAPES = ["g99_s3", "g9_s5", "g1_s1", "f32_g99_s3", "f32_g97_s4"]


def tape_columns(name):
    """Raw rows of a recorded tape as columns [s0, a, r, s1, done] and what the reference emitted per push."""
    gamma, step, n = GOLD[name + "_cfg"]
    cin, r, want = GOLD[name + "_in"], GOLD[name + "_r"], GOLD[name + "_out"]
    done = cin[:, 3].astype(np.float32) if r.dtype == np.float32 else cin[:, 3].astype(np.int64)
    return float(gamma), int(step), int(n), [cin[:, 0].copy(), cin[:, 1].copy(), r.copy(), cin[:, 2].copy(), done], want


@pytest.mark.parametrize("name", TAPES)
def test_nstep_closed_form_matches_reference_tape(name):
    """oracle.nstep_fold (the closed form the device kernel implements) against tapes recorded from the
    reference's own NStepAdder: float64 rewards, and float32 0-d arrays as worker.py feeds it."""
    gamma, step, n, cols, want = tape_columns(name)
    out, push = po.nstep_fold(cols, step, gamma)
    assert (want[:step, 1] < 0).all() and (want[step:, 1] >= 0).all()       # push i emits row i - step
    w = want[step:]
    assert np.array_equal(push, w[:, 0].astype(np.int64))
    assert np.array_equal(out[0], w[:, 1].astype(np.int64)) and np.array_equal(out[1], w[:, 2].astype(np.int64))
    assert np.array_equal(out[3], w[:, 4].astype(np.int64))
    assert np.array_equal(out[4].astype(np.int64), w[:, 5].astype(np.int64))
    assert np.array_equal(out[2].astype(np.float64), w[:, 3])               # same float operations in the same order


class _FakeShard:
    """Records what FrameRefWriter appends; frame slots wrap like the device ring."""

    def __init__(self, k, frame_shape, fcap):
        self.frame_shape, self.frame_capacity = frame_shape, fcap
        self.frame_cols = {0: k, 3: k}
        self.struct = [(np.dtype(np.int32), (k,)), (np.dtype(np.int64), ()), (np.dtype(np.float32), ()),
                       (np.dtype(np.int32), (k,)), (np.dtype(np.float32), ())]
        self.ring = np.zeros((fcap,) + frame_shape, np.uint8)
        self.tail = 0
        self.rows = []

    def append_frames(self, frames, stream=None):
        first = self.tail
        for f in frames:
            self.ring[self.tail] = f
            self.tail = (self.tail + 1) % self.frame_capacity
        return first

    def append(self, cols, prio, alpha=0.6, transform=True, stream=None):
        first = len(self.rows)
        for i in range(len(cols[0])):
            self.rows.append([c[i] for c in cols])
        return first


def test_frame_ref_writer_reproduces_framestack():
    k, shape = 4, (6, 5)
    rng = np.random.default_rng(3)
    sh = _FakeShard(k, shape, fcap=4096)
    w = FrameRefWriter(sh, k=k, frame_cols=(0, 3))
    want = []                                               # (s0 stack, a, r, s1 stack, done) as FrameStack would give
    stacks = {}
    for env in range(3):
        f = rng.integers(0, 256, shape, dtype=np.uint8)
        stacks[env] = deque([f] * k, maxlen=k)
        w.reset(env, f)
    for t in range(200):
        env = int(rng.integers(0, 3))
        f = rng.integers(0, 256, shape, dtype=np.uint8)
        a, r, done = int(rng.integers(0, 6)), float(rng.standard_normal()), float(rng.random() < 0.05)
        s0 = np.stack(stacks[env])
        stacks[env].append(f)
        want.append((s0, a, np.float32(r), np.stack(stacks[env]), np.float32(done)))
        w.step(env, f, [a, r, done], prio=1.0)
        if done:
            f = rng.integers(0, 256, shape, dtype=np.uint8)
            stacks[env] = deque([f] * k, maxlen=k)
            w.reset(env, f)
        if t % 17 == 0:
            w.flush()
    w.flush()
    assert w.flush() is None
    assert len(sh.rows) == len(want)
    for row, (s0, a, r, s1, done) in zip(sh.rows, want):
        assert row[0].dtype == np.int32 and row[0].shape == (k,)
        assert np.array_equal(sh.ring[row[0]], s0) and np.array_equal(sh.ring[row[3]], s1)
        assert row[1] == a and row[2] == r and row[4] == done
