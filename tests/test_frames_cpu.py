"""Host logic of the frame-de-duplicated layout: NStepAdder against a tape recorded from the
reference's own class (tests/golden/make_nstep_golden.py), FrameRefWriter's stacking rule against
FrameStack's (reth/reth/env/util.py:179-209) with a stand-in shard."""
import os
from collections import deque

import numpy as np
import pytest

from reth_b200.frames import FrameRefWriter, NStepAdder

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "nstep_adder.npz"))


@pytest.mark.parametrize("name", ["g99_s3", "g9_s5", "g1_s1"])
def test_nstep_adder_matches_reference_tape(name):
    gamma, step, n = GOLD[name + "_cfg"]
    adder = NStepAdder(float(gamma), int(step))
    cols, r, want = GOLD[name + "_in"], GOLD[name + "_r"], GOLD[name + "_out"]
    for i in range(int(n)):
        s0, a, s1, done = (int(x) for x in cols[i])
        res = adder.push(s0, a, float(r[i]), s1, bool(done))
        if want[i][1] < 0:
            assert res is None
        else:
            assert res is not None
            assert [res[0], res[1], res[3], int(res[4])] == [int(want[i][1]), int(want[i][2]), int(want[i][4]), int(want[i][5])]
            assert res[2] == want[i][3]                      # same float operations in the same order


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
