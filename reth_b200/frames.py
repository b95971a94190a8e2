"""Host-side bookkeeping for frame-de-duplicated transitions.

The reference's Atari pipeline stacks the last k frames into every observation
(FrameStack, reth/reth/env/util.py:179-209: reset() repeats the first frame k times, step() pushes
one new frame, the observation is the concatenation) and then stores s0 AND s1 of every transition,
so each frame is kept 2k = 8 times (its LazyFrames remedy is commented out, env/util.py:207).
Here a shard built with `frames=` keeps single frames once in a frame ring and a transition stores
k int32 frame slots for s0 and for s1; the gather rebuilds the stacks on the device.

FrameRefWriter turns per-environment step streams into batched `append_frames` + `append` calls
with exactly FrameStack's stacking rule.  n-step returns are assembled on the device at append time
(reth_replay_append_nstep, csrc/nstep.cuh: with frame references the n-step s1 is just the k slots of a
later row); an actor that still folds them itself uses the reference's own NStepAdder unchanged."""
from collections import deque

import numpy as np


class FrameRefWriter:
    """Collects steps of several environments and appends them to a frame-de-duplicated shard.

        w = FrameRefWriter(shard, k=4, frame_cols=(0, 3))
        w.reset(env_id, first_frame)                       # FrameStack.reset: k copies of the frame
        w.step(env_id, next_frame, [a, r, done], prio)      # one transition (s0 = stack before, s1 = after)
        w.flush()                                           # one append_frames + one append

    A transition's row is [s0 refs, *scalars before s1, s1 refs, *the rest] following the shard's
    struct: the two columns named by frame_cols get the references, the other columns get the
    values passed to step() in order.  Frames are numbered in arrival order, so references made
    before a flush are resolved when the ring slots are known."""

    def __init__(self, shard, k=4, frame_cols=(0, 3), alpha=0.6, transform=True):
        self.shard, self.k = shard, int(k)
        self.frame_cols = tuple(frame_cols)
        assert all(shard.frame_cols.get(c) == self.k for c in self.frame_cols), "frame columns must hold k references"
        self.other_cols = [c for c in range(len(shard.struct)) if c not in self.frame_cols]
        self.alpha, self.transform = alpha, transform
        self._stacks = {}            # env id -> deque of frame tokens (>= 0: ring slot, < 0: pending frame -(j+1))
        self._frames = []            # pending frames
        self._rows = []              # pending transitions: (s0 tokens, s1 tokens, other values, priority)

    def _push_frame(self, frame):
        self._frames.append(np.ascontiguousarray(frame, dtype=np.uint8).reshape(self.shard.frame_shape))
        return -len(self._frames)

    def reset(self, env, frame):
        tok = self._push_frame(frame)
        self._stacks[env] = deque([tok] * self.k, maxlen=self.k)            # env/util.py:194-198

    def step(self, env, next_frame, others, prio=None):
        st = self._stacks[env]
        s0 = tuple(st)
        st.append(self._push_frame(next_frame))                             # env/util.py:200-203
        self._rows.append((s0, tuple(st), tuple(others), prio))

    def flush(self, stream=None):
        """-> first transition slot of the append, or None when nothing was pending."""
        if not self._frames:
            return None
        fcap = self.shard.frame_capacity
        first = self.shard.append_frames(np.stack(self._frames), stream=stream)

        def slot(tok):
            return tok if tok >= 0 else (first + (-tok - 1)) % fcap
        for env, st in self._stacks.items():
            self._stacks[env] = deque([slot(t) for t in st], maxlen=self.k)
        out = None
        if self._rows:
            n = len(self._rows)
            cols = [None] * len(self.shard.struct)
            for j, c in enumerate(self.frame_cols[:2]):
                cols[c] = np.array([[slot(t) for t in row[j]] for row in self._rows], dtype=np.int32).reshape(n, self.k)
            for j, c in enumerate(self.other_cols):
                dt, sh = self.shard.struct[c]
                cols[c] = np.array([row[2][j] for row in self._rows], dtype=dt).reshape((n,) + sh)
            prios = [row[3] for row in self._rows]
            prio = None if any(p is None for p in prios) else np.asarray(prios, dtype=np.float32)
            out = self.shard.append(cols, prio, alpha=self.alpha, transform=self.transform, stream=stream)
        self._frames, self._rows = [], []
        return out
