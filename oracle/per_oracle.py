"""ctypes face of oracle/per_oracle.c -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.  The product package (reth_b200/) must
never import it.  Parity status: pinned against the reference's own njit code
through tests/golden/ (see per_oracle.c header).

Class names mirror the reference objects they restate:
  OracleSumTree   <- reth_buffer/reth_buffer/utils/sumtree.py:82-113 (NumbaSumTree)
  OraclePERSampler<- reth_buffer/reth_buffer/sampler/per_sampler.py:5-35
  OracleRing      <- reth/reth/buffer/buffer.py:4-93 + cache_policy/fifo_policy.py:6-18
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libper_oracle.so")


def build(force=False):
    """Compile per_oracle.c with gcc (also called by __graft_entry__.build())."""
    src = os.path.join(_HERE, "per_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(
            ["gcc", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math",
             "-fvisibility=hidden", "-o", _LIB_PATH, src, "-lm"])
    return _LIB_PATH


_lib = None
_i64p = C.POINTER(C.c_int64)
_f64p = C.POINTER(C.c_double)
_f32p = C.POINTER(C.c_float)
_u8p = C.POINTER(C.c_uint8)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(_LIB_PATH)
    vp = C.c_void_p
    sig = {
        "po_tree_create": (vp, [C.c_int64]),
        "po_tree_destroy": (None, [vp]),
        "po_tree_clear": (None, [vp]),
        "po_tree_sum_ptr": (_f64p, [vp]),
        "po_tree_min_ptr": (_f64p, [vp]),
        "po_tree_val_ptr": (_f64p, [vp]),
        "po_find_index": (C.c_int64, [vp, C.c_double]),
        "po_tree_update": (None, [vp, vp, vp, C.c_int64]),
        "po_tree_sample": (None, [vp, C.c_int64, vp, vp, vp]),
        "po_tree_total": (C.c_double, [vp]),
        "po_tree_minimum": (C.c_double, [vp, C.c_int]),
        "po_normalize_f32": (None, [vp, C.c_int64, C.c_double, vp]),
        "po_normalize_f64": (None, [vp, C.c_int64, C.c_double, vp]),
        "po_is_weights": (None, [vp, C.c_int64, C.c_double, C.c_double, vp]),
        "po_ring_create": (vp, [C.c_int64, C.c_int32, vp]),
        "po_ring_destroy": (None, [vp]),
        "po_ring_col_ptr": (_u8p, [vp, C.c_int]),
        "po_ring_tail": (C.c_int64, [vp]),
        "po_ring_size": (C.c_int64, [vp]),
        "po_ring_reset": (None, [vp]),
        "po_fifo_indices": (None, [C.c_int64, vp, C.c_int64, vp]),
        "po_ring_append": (None, [vp, vp, C.c_int64, vp]),
        "po_ring_select": (None, [vp, vp, C.c_int64, C.c_int64, vp]),
        "po_step_loop": (C.c_double, [vp, vp, C.c_int64, C.c_int64, C.c_int64, vp, vp,
                                      C.c_double, C.c_double, C.c_int, vp, vp, vp, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class OracleSumTree:
    """utils/sumtree.py:82-113; arrays are exposed as numpy views (_sum/_min/_val)."""

    def __init__(self, capacity):
        self.capacity = int(capacity)
        self._h = lib().po_tree_create(self.capacity)
        L = lib()
        n = self.capacity
        self._sum = np.ctypeslib.as_array(L.po_tree_sum_ptr(self._h), shape=(n,))
        self._min = np.ctypeslib.as_array(L.po_tree_min_ptr(self._h), shape=(n,))
        self._val = np.ctypeslib.as_array(L.po_tree_val_ptr(self._h), shape=(n,))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().po_tree_destroy(self._h)
            self._h = None

    def clear(self):
        lib().po_tree_clear(self._h)

    def update(self, indices, weights):
        idx = np.ascontiguousarray(np.atleast_1d(indices), dtype=np.int64)
        w = np.ascontiguousarray(np.atleast_1d(weights), dtype=np.float64)
        assert len(idx) == len(w)
        lib().po_tree_update(self._h, _ptr(idx), _ptr(w), len(idx))

    def sum(self):
        return lib().po_tree_total(self._h)

    def min(self, mode=0):
        return lib().po_tree_minimum(self._h, mode)

    def find_index(self, weight):
        return lib().po_find_index(self._h, float(weight))

    def sample(self, batch_size, u):
        """_numba_sample (sumtree.py:70-79) with injected uniforms u[batch]."""
        u = np.ascontiguousarray(u, dtype=np.float64)
        assert len(u) == batch_size
        idx = np.empty(batch_size, np.int64)
        val = np.empty(batch_size, np.float64)
        lib().po_tree_sample(self._h, batch_size, _ptr(u), _ptr(idx), _ptr(val))
        return idx, val


def normalize_weights(weights, alpha):
    """per_sampler.py:16-17 in the dtype of the incoming array; returns float64."""
    w = np.ascontiguousarray(np.atleast_1d(weights))
    out = np.empty(len(w), np.float64)
    if w.dtype == np.float32:
        lib().po_normalize_f32(_ptr(w), len(w), float(alpha), _ptr(out))
    else:
        w = w.astype(np.float64)
        lib().po_normalize_f64(_ptr(w), len(w), float(alpha), _ptr(out))
    return out


def is_weights(val, pmin, beta):
    """per_sampler.py:26."""
    val = np.ascontiguousarray(val, dtype=np.float64)
    out = np.empty(len(val), np.float64)
    lib().po_is_weights(_ptr(val), len(val), float(pmin), float(beta), _ptr(out))
    return out


class OraclePERSampler:
    """per_sampler.py:5-35 with constant alpha/beta supplied per call (the
    Schedule, utils/schedule.py, is host scalar logic shared with the product)."""

    def __init__(self, capacity, min_mode=0):
        self.sumtree = OracleSumTree(capacity)
        self.min_mode = min_mode

    def update(self, indices, weights, alpha, transform=True):
        w = normalize_weights(weights, alpha) if transform else np.asarray(weights, np.float64)
        self.sumtree.update(indices, w)

    def sample(self, batch_size, u, beta):
        idx, val = self.sumtree.sample(batch_size, u)
        return idx, is_weights(val, self.sumtree.min(self.min_mode), beta), val


class OracleRing:
    """Columnar ring (buffer.py:57-93) with FIFO slots (fifo_policy.py:11-18)."""

    def __init__(self, capacity, struct):
        """struct: list of (dtype, per-row shape)."""
        self.capacity = int(capacity)
        self.struct = [(np.dtype(dt), tuple(sh)) for dt, sh in struct]
        self.row_bytes = np.array(
            [dt.itemsize * int(np.prod(sh, dtype=np.int64)) for dt, sh in self.struct], np.int64)
        self._h = lib().po_ring_create(self.capacity, len(self.struct), _ptr(self.row_bytes))
        if not self._h:
            raise MemoryError("po_ring_create failed")

    def __del__(self):
        if getattr(self, "_h", None):
            lib().po_ring_destroy(self._h)
            self._h = None

    @property
    def tail(self):
        return lib().po_ring_tail(self._h)

    @property
    def size(self):
        return lib().po_ring_size(self._h)

    def column(self, c):
        dt, sh = self.struct[c]
        raw = np.ctypeslib.as_array(lib().po_ring_col_ptr(self._h, c),
                                    shape=(self.capacity * int(self.row_bytes[c]),))
        return raw.view(dt).reshape((self.capacity,) + sh)

    def append_batch(self, cols):
        n = len(cols[0])
        assert n <= self.capacity
        arrs = [np.ascontiguousarray(c, dtype=dt) for c, (dt, _) in zip(cols, self.struct)]
        ptrs = (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
        idx = np.empty(n, np.int64)
        lib().po_ring_append(self._h, ptrs, n, _ptr(idx))
        return idx

    def select(self, indices, slot_mask=-1):
        idx = np.ascontiguousarray(indices, dtype=np.int64)
        n = len(idx)
        outs = [np.empty((n,) + sh, dt) for dt, sh in self.struct]
        ptrs = (C.c_void_p * len(outs))(*[a.ctypes.data for a in outs])
        lib().po_ring_select(self._h, _ptr(idx), n, slot_mask, ptrs)
        return outs


def global_sample(trees, shard, batch, u, beta, min_mode=0):
    """Mode-G oracle: n reference trees + the range rule.  Shard `shard` serves the
    targets of the global cdf inside [prefix, prefix + total_shard); indices come from the
    reference descent (_numba_find_index, utils/sumtree.py:34-58) on the LOCAL target, weights use
    the global minimum.  -> (first, indices, weights, values)."""
    totals = [t.sum() for t in trees]
    total, prefix = 0.0, []
    for t in totals:
        prefix.append(total)
        total = total + t
    mins = [t._min[0] for t, s in zip(trees, totals) if s != 0.0 and t._min[0] != 0.0]
    pmin = min(mins) if mins else 1.0
    seg = total / float(batch)
    u = np.asarray(u, np.float64)
    first, idx, val = None, [], []
    for i in range(batch):
        tgt = (float(i) + u[i]) * seg
        lo_ok = (not prefix[shard] > 0.0) or tgt >= prefix[shard]
        hi_ok = shard == len(trees) - 1 or not tgt >= prefix[shard] + totals[shard]
        if lo_ok and hi_ok:
            if first is None:
                first = i
            j = trees[shard].find_index(tgt - prefix[shard])
            idx.append(j)
            val.append(trees[shard]._val[j])
    idx = np.asarray(idx, np.int64)
    val = np.asarray(val, np.float64)
    return (first if first is not None else 0), idx, is_weights(val, pmin, beta), val


def fifo_indices(capacity, tail, n):
    """fifo_policy.py:11-18 -> (indices, new_tail)."""
    t = C.c_int64(tail)
    out = np.empty(n, np.int64)
    lib().po_fifo_indices(capacity, C.byref(t), n, _ptr(out))
    return out, t.value


def step_loop(tree, ring, slot_mask, steps, batch, u, new_prio, alpha, beta, min_mode=0):
    """Timed CPU-baseline loop (po_step_loop). ring may be None (tree-only)."""
    u = np.ascontiguousarray(u, np.float64)
    new_prio = np.ascontiguousarray(new_prio, np.float32)
    assert u.size >= steps * batch and new_prio.size >= steps * batch
    idx = np.empty(batch, np.int64)
    val = np.empty(batch, np.float64)
    w = np.empty(batch, np.float64)
    p = np.empty(batch, np.float64)
    if ring is not None:
        outs = [np.empty((batch,) + sh, dt) for dt, sh in ring.struct]
        ptrs = (C.c_void_p * len(outs))(*[a.ctypes.data for a in outs])
        rh = ring._h
    else:
        outs, ptrs, rh = [], None, None
    acc = lib().po_step_loop(tree._h, rh, slot_mask, steps, batch, _ptr(u), _ptr(new_prio),
                             float(alpha), float(beta), min_mode, _ptr(idx), _ptr(val), _ptr(w),
                             _ptr(p), ptrs)
    return acc, idx, w, outs


class OracleUniformSampler:
    """sampler/uniform_sampler.py:6-25: a log of the first `capacity` notified indices; sample =
    log[draws], draws in [0, tail) (the reference draws them with np.random.choice(tail, batch)),
    weights are int64 ones.  The reference raises IndexError when one message crosses the end of
    the log (:21-22 writes indices[capacity]); the restatement (and the engine) drop the excess."""

    def __init__(self, capacity):
        self.indices = np.zeros(capacity, dtype="i8")
        self.capacity = capacity
        self.tail = 0

    def update(self, indices, weights=None):
        if self.tail == self.capacity:                       # :18-19
            return
        idx = np.asarray(indices, dtype=np.int64)
        m = min(len(idx), self.capacity - self.tail)
        self.indices[self.tail:self.tail + m] = idx[:m]
        self.tail += m

    def sample(self, draws):
        draws = np.asarray(draws, dtype=np.int64)
        assert np.all((draws >= 0) & (draws < self.tail))
        return self.indices[draws], np.ones(len(draws), dtype="i8")

    def clear(self):
        self.indices[:] = 0
        self.tail = 0


class OracleFIFOSampler:
    """sampler/fifo_sampler.py:8-29: deque(maxlen=capacity); update appends on the left (a full
    queue drops its oldest entry), sample pops the oldest entries from the right."""

    def __init__(self, capacity):
        from collections import deque
        self.buffer = deque(maxlen=capacity)

    def ready_sample(self, batch_size):
        return len(self.buffer) > batch_size                 # :17-18

    def update(self, indices, weights):
        for i, idx in enumerate(indices):
            self.buffer.appendleft((int(idx), float(weights[i])))

    def sample(self, batch_size):
        out = [self.buffer.pop() for _ in range(batch_size)]
        return np.array([o[0] for o in out], np.int64), np.array([o[1] for o in out], np.float64)

    def clear(self):
        self.buffer.clear()


def nstep_fold(cols, step, gamma, r_col=2, s1_col=3, done_col=4):
    """Closed form of what NStepAdder.push emits over one environment stream
    (reth/reth/utils/nstep_adder.py:11-28, fed as in test/apex-dqn/worker.py:35,53).

    cols: the stream's raw 1-step rows, one array per column.  Row t leaves the adder at push t+step
    (:13-14) after the pushes t+1 .. t+step-1 have touched it (:16-24): push t+j adds gamma^j * r_{t+j} to
    its reward and replaces its s1, unless one of the rows t .. t+j-1 has its done flag set (the scan runs
    from the newest row and stops at the first done row, :18-19).  gamma^j is the running product
    t_gamma *= gamma (:22); a float32 reward column accumulates in float32 with the factor rounded to
    float32 (numpy on the 0-d float32 array of worker.py:49), a float64 one in float64.
    -> (emitted columns for rows 0 .. L-step-1, push index that emits each row = t + step)."""
    L = len(cols[r_col])
    E = max(0, L - step)
    out = [np.array(c[:E], copy=True) for c in cols]
    r = np.asarray(cols[r_col])
    done = np.asarray(cols[done_col]).reshape(L, -1).any(axis=1) if L else np.zeros(0, bool)
    f32 = r.dtype == np.float32
    for t in range(E):
        acc = r[t]
        g = float(gamma)
        last = t
        for j in range(1, step):
            if done[t + j - 1]:
                break
            if f32:
                acc = np.float32(acc + np.float32(np.float32(g) * r[t + j]))
            else:
                acc = acc + g * r[t + j]
            g *= float(gamma)
            last = t + j
        out[r_col][t] = acc
        out[s1_col][t] = cols[s1_col][last]
    return out, np.arange(E) + step
