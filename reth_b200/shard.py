"""ReplayShard: Python face of one C-ABI replay shard (include/reth_replay.h).

Thin by design: argument marshalling (numpy = host pointers, torch CUDA tensors /
__cuda_array_interface__ = device pointers), error translation, and the tensor
hand-off to the learner.  All arithmetic runs in the CUDA library.
"""
import ctypes as C

import numpy as np

from . import _cabi
from ._cabi import DEVICE, F32, F64, HOST, I32, I64, MIN_INPROCESS, MIN_SERVICE, ReplayError  # noqa: F401

_TYPESTR = {np.dtype(np.uint8): "|u1", np.dtype(np.int8): "|i1", np.dtype(np.bool_): "|b1"}


def _typestr(dt):
    dt = np.dtype(dt)
    return _TYPESTR.get(dt, dt.str)


class _DevArray:
    """Borrowed device memory exposed through __cuda_array_interface__ (v3)."""

    def __init__(self, ptr, shape, dtype, owner):
        self._owner = owner
        self.__cuda_array_interface__ = {
            "shape": tuple(int(s) for s in shape), "typestr": _typestr(dtype),
            "data": (int(ptr), False), "version": 3, "strides": None, "stream": None,
        }


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _ptr_loc(x):
    """-> (address, loc, numpy dtype, nelems, keepalive)"""
    if x is None:
        return None, HOST, None, 0, None
    if _is_torch(x):
        if not x.is_contiguous():
            x = x.contiguous()
        dt = np.dtype(str(x.dtype).replace("torch.", ""))
        return x.data_ptr(), (DEVICE if x.is_cuda else HOST), dt, x.numel(), x
    if hasattr(x, "__cuda_array_interface__"):
        cai = x.__cuda_array_interface__
        if cai.get("strides"):
            raise ValueError("device array must be C-contiguous")
        n = int(np.prod(cai["shape"], dtype=np.int64))
        return cai["data"][0], DEVICE, np.dtype(cai["typestr"]), n, x
    a = np.ascontiguousarray(x)
    return a.ctypes.data, HOST, a.dtype, a.size, a


def _has_cuda_tensor(keep):
    stack = [keep]
    while stack:
        x = stack.pop()
        if isinstance(x, (list, tuple)):
            stack.extend(x)
        elif x is not None and _is_torch(x) and x.is_cuda:
            return True
    return False


def _stream_handle(stream, device):
    if stream is None:
        import sys
        torch = sys.modules.get("torch")
        if torch is not None and torch.cuda.is_available() and torch.cuda.is_initialized():
            return int(torch.cuda.current_stream(device).cuda_stream)
        return 0
    if hasattr(stream, "cuda_stream"):
        return int(stream.cuda_stream)
    return int(stream)


class ShardPipe:
    """Asynchronous loader pipeline of a shard (reth_replay_pipe_*, include/reth_replay.h):
    groups of `group` batches are drawn + gathered ahead on the shard's own streams, indices and
    weights land in pinned host rings, priority write-backs go through a pinned staging ring."""

    def __init__(self, shard, batch, group, seed=0):
        self.shard = shard
        self._lib, self._h = shard._lib, shard._h
        info = _cabi.PipeInfo()
        shard._ck(self._lib.reth_replay_pipe_open(self._h, int(batch), int(group), int(seed) & (2**64 - 1),
                                                  C.byref(info)))
        self.batch, self.group, self.n_groups = info.batch, info.group, info.n_groups
        self.upd_entries, E = info.upd_entries, info.entry_elems
        self.entry_elems = E

        def host(addr, nbytes):
            return np.frombuffer((C.c_uint8 * nbytes).from_address(addr), np.uint8)
        self.idx_host = host(info.idx_host, self.n_groups * E * 8).view(np.int64).reshape(self.n_groups, group, batch)
        self.w_host = host(info.w_host, self.n_groups * E * 8).view(np.float64).reshape(self.n_groups, group, batch)
        raw = host(info.upd_host, self.upd_entries * E * 16).reshape(self.upd_entries, E * 16)
        self.upd_idx = raw[:, :E * 8].view(np.int64)              # [entry][E]
        self.upd_p32 = raw[:, E * 8:E * 12].view(np.float32)      # [entry][E]
        self.upd_p64 = raw[:, E * 8:].view(np.float64)            # [entry][E]
        self.issued = self.waited = 0
        self.lend = 1                                             # delivered groups that stay lent
        self._g = C.c_int32()
        self._e = C.c_int32()
        p = C.c_void_p()
        shard._ck(self._lib.reth_replay_pipe_stream(self._h, C.byref(p)))
        self.stream = p.value

    @property
    def in_flight(self):
        return self.issued - self.waited

    def can_issue(self):
        return self.in_flight + min(self.waited, self.lend) < self.n_groups

    def hold(self, groups):
        """Keep the newest `groups` delivered groups lent (reth_replay_pipe_hold); before the first issue."""
        self.shard._ck(self._lib.reth_replay_pipe_hold(self._h, int(groups)))
        self.lend = int(groups)

    def advance(self, beta):
        """Hand the oldest lent group back and issue a new group into its slots; never blocks."""
        self.shard._ck(self._lib.reth_replay_pipe_advance(self._h, float(beta)))
        self.issued += 1

    def ready(self):
        """True when wait() would not block."""
        self.shard._ck(self._lib.reth_replay_pipe_ready(self._h, C.byref(self._g)))
        return bool(self._g.value)

    def issue(self, beta):
        self.shard._ck(self._lib.reth_replay_pipe_issue(self._h, float(beta)))
        self.issued += 1

    def wait(self):
        self.shard._ck(self._lib.reth_replay_pipe_wait(self._h, C.byref(self._g)))
        self.waited += 1
        return self._g.value

    def next(self, beta):
        """wait() for the oldest group, then issue() a new one: one C call."""
        self.shard._ck(self._lib.reth_replay_pipe_next(self._h, float(beta), C.byref(self._g)))
        self.waited += 1
        self.issued += 1
        return self._g.value

    def update_begin(self):
        self.shard._ck(self._lib.reth_replay_pipe_update_begin(self._h, C.byref(self._e)))
        return self._e.value

    def update_commit(self, entry, n, f32, alpha, transform=True):
        self.shard._ck(self._lib.reth_replay_pipe_update_commit(self._h, entry, n, F32 if f32 else F64,
                                                                float(alpha), int(bool(transform))))

    def update_async(self, indices, prio, alpha, transform=True):
        idx = np.ascontiguousarray(indices, dtype=np.int64)
        w = np.ascontiguousarray(prio)
        if w.dtype != np.float32:
            w = w.astype(np.float64, copy=False)
        self.shard._ck(self._lib.reth_replay_update_async(self._h, idx.ctypes.data, w.ctypes.data,
                                                          F32 if w.dtype == np.float32 else F64, idx.size,
                                                          float(alpha), int(bool(transform))))

    def drain(self):
        self.shard._ck(self._lib.reth_replay_pipe_drain(self._h))

    def close(self):
        if self._h is not None and self.shard._h is not None:
            self._lib.reth_replay_pipe_close(self._h)
        self._h = None


class IpcRing:
    """A peer's batch ring mapped into this process (reth_ipc_open): same GPU in another process
    (loader hand-off) or another GPU (cross-shard batch assembly over NVLink)."""

    def __init__(self, blob):
        self._lib = _cabi.lib()
        self.ex = _cabi.IpcExport.from_buffer_copy(blob)
        self._v = C.c_void_p()
        rc = self._lib.reth_ipc_open(C.byref(self.ex), C.byref(self._v))
        if rc != 0:
            self._v = None
            _cabi.check(None, rc)

    def close(self):
        if self._v is not None:
            self._lib.reth_ipc_close(self._v)
            self._v = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ReplayShard:
    """One replay shard on one GPU.

    struct: list of (dtype, per-row shape), the row layout the reference discovers
    from its first row (client/torch_cuda_loader.py:106-112).
    """

    def __init__(self, capacity, struct, max_batch, ring_depth=2, device=0, min_mode=MIN_SERVICE,
                 store_rows=True, update_cluster=0, gather_mode=0, out_cast=None, frames=None):
        """frames: {"shape": per-frame shape, "capacity": frames in the frame ring, "columns": [column
        indices]} turns those columns of `struct` -- uint8 stacks of shape (K,) + frame shape -- into
        frame-reference columns: a transition stores K int32 frame slots (append_frames() returns the
        slots), the gather hands the learner the K-frame stack.  Frames are stored once (8x less HBM per
        Atari transition than re-storing 4-stacks for s0 and s1)."""
        self._h = None
        self._lib = _cabi.lib()
        self.capacity = int(capacity)
        self.struct = [(np.dtype(dt), tuple(int(s) for s in sh)) for dt, sh in (struct or [])]
        self.frame_cols, self.frame_shape, self.frame_capacity = {}, None, 0
        logical = list(self.struct)
        if frames:
            self.frame_shape = tuple(int(x) for x in frames["shape"])
            self.frame_capacity = int(frames["capacity"])
            for c in frames["columns"]:
                dt, sh = self.struct[c]
                if dt != np.dtype(np.uint8) or sh[1:] != self.frame_shape or len(sh) != len(self.frame_shape) + 1:
                    raise ValueError(f"column {c}: a frame column is uint8 of shape (K,) + {self.frame_shape}")
                self.frame_cols[c] = sh[0]
                self.struct[c] = (np.dtype(np.int32), (sh[0],))       # what the ring stores
        self.row_bytes = [dt.itemsize * int(np.prod(sh, dtype=np.int64)) for dt, sh in self.struct]
        self.max_batch = int(max_batch)
        self.ring_depth = int(ring_depth)
        self.device = int(device)
        self.store_rows = bool(store_rows) and len(self.struct) > 0
        cfg = _cabi.Config()
        cfg.capacity = self.capacity
        cfg.n_columns = len(self.struct)
        for i, rb in enumerate(self.row_bytes):
            cfg.row_bytes[i] = rb
        cfg.max_batch = self.max_batch
        cfg.ring_depth = self.ring_depth
        cfg.device = self.device
        cfg.min_mode = int(min_mode)
        cfg.update_cluster = int(update_cluster)
        cfg.store_rows = int(self.store_rows)
        # out_cast[c]: None | "f32" | "f32/255": the gather widens a uint8 column to float32 for the learner
        codes = {None: 0, 0: 0, "f32": 1, 1: 1, "f32/255": 2, 2: 2}
        self.out_cast = [codes[x] for x in (out_cast or [None] * len(self.struct))]
        if len(self.out_cast) != len(self.struct):
            raise ValueError("out_cast needs one entry per column")
        for i, code in enumerate(self.out_cast):
            if code and logical[i][0] != np.dtype(np.uint8):
                raise ValueError(f"column {i}: only uint8 columns can be widened to float32")
            cfg.out_cast[i] = code | (_cabi.COL_FRAMES if i in self.frame_cols else 0)
        if self.frame_cols:
            cfg.frame_bytes = int(np.prod(self.frame_shape, dtype=np.int64))
            cfg.frame_capacity = self.frame_capacity
        # what a batch slot holds (struct is what the ring stores)
        self.out_struct = [((np.dtype(np.float32) if code else dt), sh)
                           for (dt, sh), code in zip(logical, self.out_cast)]
        cfg.gather_mode = int(gather_mode)      # 0 auto (= 1: LDG/STG.128, the faster one on B200), 1 LDG/STG, 2 TMA bulk
        h = C.c_void_p()
        rc = self._lib.reth_replay_create(C.byref(cfg), C.byref(h))
        if rc != 0:
            _cabi.check(None, rc)
        self._h = h

    # -- lifetime ---------------------------------------------------------
    def close(self):
        if self._h is not None:
            self._lib.reth_replay_destroy(self._h)            # synchronises the device first
            self._h = None
            self.__dict__.pop("_held", None)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            _cabi.check(self._h, rc)

    def _hold(self, keep, stream):
        """Device tensors handed to a call that runs on a raw (non-torch) stream: keep them referenced until an event
        recorded behind the call on that stream has completed, so torch's caching allocator cannot hand their memory
        to somebody else while the library's kernels still read it.  (tensor.record_stream() would do the same, but
        it makes the allocator record an event on OUR stream when the tensor is freed -- after close() has destroyed
        the stream that aborts the process.)"""
        h = _stream_handle(stream, self.device)
        if not h or not _has_cuda_tensor(keep):
            return
        import torch
        if int(torch.cuda.current_stream(self.device).cuda_stream) == int(h):
            return                                            # the caller's own stream: torch orders the reuse itself
        ev = torch.cuda.Event()
        ev.record(torch.cuda.ExternalStream(int(h), device=self.device))
        held = self.__dict__.setdefault("_held", [])
        held.append((ev, keep))
        while held and held[0][0].query():
            held.pop(0)

    def _s(self, stream):
        return C.c_void_p(_stream_handle(stream, self.device))

    # -- marshalling ------------------------------------------------------
    def _prio(self, prio, n):
        if prio is None:
            return None, F32, HOST, None
        p, loc, dt, cnt, keep = _ptr_loc(prio)
        if dt not in (np.dtype(np.float32), np.dtype(np.float64)):
            if loc == DEVICE:
                raise ValueError("device priorities must be float32 or float64")
            keep = np.ascontiguousarray(keep, dtype=np.float64)   # ints etc. -> float64 like numpy would
            p, dt = keep.ctypes.data, keep.dtype
        if cnt != n:
            raise AssertionError("len(weights) != number of rows")   # client/client.py:25,38
        return p, (F32 if dt == np.dtype(np.float32) else F64), loc, keep

    def _cols(self, cols, n):
        if not self.store_rows:
            return None, HOST, []
        if not isinstance(cols, (list, tuple)) or len(cols) != len(self.struct):
            raise AssertionError("data must be a list with one array per column")   # client.py:22
        ptrs = (C.c_void_p * len(cols))()
        keep, locs = [], set()
        for i, (col, (dt, sh)) in enumerate(zip(cols, self.struct)):
            if not _is_torch(col) and not hasattr(col, "__cuda_array_interface__"):
                col = np.ascontiguousarray(col, dtype=dt)
            p, loc, cdt, cnt, k = _ptr_loc(col)
            if cdt.itemsize * cnt != n * self.row_bytes[i]:
                raise AssertionError(f"column {i}: expected {n} rows of {self.row_bytes[i]} bytes")
            ptrs[i] = p
            keep.append(k)
            locs.add(loc)
        if len(locs) > 1:
            raise ValueError("all columns of one append must live on the same side (host or device)")
        return ptrs, locs.pop(), keep

    # -- write side -------------------------------------------------------
    def clear(self, stream=None):
        self._ck(self._lib.reth_replay_clear(self._h, self._s(stream)))

    def resize_batch_ring(self, max_batch, ring_depth):
        """Re-shape the learner-batch ring (reth_replay_batch_ring_resize); all slots are emptied."""
        self._ck(self._lib.reth_replay_batch_ring_resize(self._h, int(max_batch), int(ring_depth)))
        self.max_batch, self.ring_depth = int(max_batch), int(ring_depth)

    def append(self, cols, prio=None, alpha=0.6, transform=True, stream=None):
        """Client.append / PrioritizedBuffer.append_batch; returns the first FIFO slot."""
        n = len(prio) if prio is not None else len(cols[0])
        ptrs, cloc, keep = self._cols(cols, n)
        p, pdt, ploc, pkeep = self._prio(prio, n)
        first = C.c_int64(-1)
        self._ck(self._lib.reth_replay_append(self._h, ptrs, cloc, p, pdt, ploc, n, float(alpha),
                                              int(bool(transform)), C.byref(first), self._s(stream)))
        self._keep = (keep, pkeep)
        self._hold(self._keep, stream)
        return first.value

    def nstep_open(self, n_streams, step, gamma, r_col=2, s1_col=3, done_col=4, max_rows=4096):
        """n-step assembly at append time (reth_replay_nstep_open): replaces NStepAdder on the actor."""
        self._ck(self._lib.reth_replay_nstep_open(self._h, int(n_streams), int(step), float(gamma), int(r_col),
                                                  int(s1_col), int(done_col), int(max_rows)))

    def nstep_reset(self, stream_id=-1):
        self._ck(self._lib.reth_replay_nstep_reset(self._h, int(stream_id)))

    def append_nstep(self, stream_id, cols, prio=None, alpha=0.6, transform=True, stream=None):
        """Raw 1-step rows of one environment stream -> finished n-step transitions in the ring.
        prio[p] belongs to the row that push p sends out of the adder.  -> (first slot, rows stored)."""
        n = len(prio) if prio is not None else len(cols[0])
        ptrs, cloc, keep = self._cols(cols, n)
        p, pdt, ploc, pkeep = self._prio(prio, n)
        first, emitted = C.c_int64(-1), C.c_int64(0)
        self._ck(self._lib.reth_replay_append_nstep(self._h, int(stream_id), ptrs, cloc, p, pdt, ploc, n, float(alpha),
                                                    int(bool(transform)), C.byref(first), C.byref(emitted), self._s(stream)))
        self._keep = (keep, pkeep)
        return first.value, emitted.value

    def append_frames(self, frames, stream=None):
        """n frames -> the frame ring; returns the first frame slot (slots are first, first+1, ... mod
        frame_capacity).  Write these slots into the frame-reference columns of later appends."""
        p, loc, dt, cnt, keep = _ptr_loc(frames)
        fb = int(np.prod(self.frame_shape, dtype=np.int64))
        if dt != np.dtype(np.uint8) or cnt % fb:
            raise ValueError("frames must be uint8 with the shard's frame shape")
        first = C.c_int64()
        self._ck(self._lib.reth_replay_append_frames(self._h, p, loc, cnt // fb, C.byref(first), self._s(stream)))
        self._keep = keep
        return first.value

    def frame_read(self, first, n, stream=None):
        out = np.empty((int(n),) + self.frame_shape, np.uint8)
        self._ck(self._lib.reth_replay_frame_read(self._h, int(first), int(n), out.ctypes.data, self._s(stream)))
        return out

    def append_at(self, slots, cols, prio=None, alpha=0.6, transform=True, stream=None):
        sp, sloc, sdt, n, skeep = _ptr_loc(slots)
        if sdt != np.dtype(np.int64):
            if sloc == DEVICE:
                raise ValueError("device slots must be int64")
            skeep = np.ascontiguousarray(skeep, dtype=np.int64)
            sp = skeep.ctypes.data
        ptrs, cloc, keep = self._cols(cols, n)
        p, pdt, ploc, pkeep = self._prio(prio, n)
        self._ck(self._lib.reth_replay_append_at(self._h, sp, sloc, ptrs, cloc, p, pdt, ploc, n,
                                                 float(alpha), int(bool(transform)), self._s(stream)))
        self._keep = (skeep, keep, pkeep)
        self._hold(self._keep, stream)

    def update(self, indices, prio, alpha=0.6, transform=True, stream=None):
        """Client.update_priorities -> PERSampler.update."""
        ip, iloc, idt, n, ikeep = _ptr_loc(indices)
        if idt not in (np.dtype(np.int32), np.dtype(np.int64)):
            if iloc == DEVICE:
                raise ValueError("device indices must be int32 or int64")
            ikeep = np.ascontiguousarray(ikeep, dtype=np.int64)
            ip, idt = ikeep.ctypes.data, ikeep.dtype
        p, pdt, ploc, pkeep = self._prio(prio, n)
        if p is None:
            raise AssertionError("update needs priorities")
        self._ck(self._lib.reth_replay_update(self._h, ip, I32 if idt.itemsize == 4 else I64, iloc,
                                              p, pdt, ploc, n, float(alpha), int(bool(transform)),
                                              self._s(stream)))
        self._keep = (ikeep, pkeep)
        self._hold(self._keep, stream)

    def update_slot(self, slot, prio, alpha=0.6, transform=True, stream=None):
        p, pdt, ploc, pkeep = self._prio(prio, self.slot_batch(slot))
        self._ck(self._lib.reth_replay_update_slot(self._h, slot, p, pdt, ploc, float(alpha),
                                                   int(bool(transform)), self._s(stream)))
        self._keep = pkeep

    # -- read side --------------------------------------------------------
    def sample(self, batch, beta=0.4, uniforms=None, seed=0, slot=0, stream=None):
        up, uloc, udt, cnt, ukeep = _ptr_loc(uniforms)
        if uniforms is not None:
            if udt != np.dtype(np.float64):
                if uloc == DEVICE:
                    raise ValueError("device uniforms must be float64")
                ukeep = np.ascontiguousarray(ukeep, dtype=np.float64)
                up = ukeep.ctypes.data
            assert cnt == batch
        self._ck(self._lib.reth_replay_sample(self._h, int(batch), up, uloc, int(seed) & (2**64 - 1),
                                              float(beta), int(slot), self._s(stream)))
        self._keep = ukeep

    def sample_global(self, global_batch, all_roots, n_shards, shard, beta=0.4, uniforms=None,
                      seed=0, slot=0, stream=None):
        rp, rloc, rdt, rcnt, rkeep = _ptr_loc(all_roots)
        assert rloc == DEVICE and rdt == np.dtype(np.float64) and rcnt >= 2 * n_shards
        up, uloc, udt, cnt, ukeep = _ptr_loc(uniforms)
        if uniforms is not None and udt != np.dtype(np.float64):
            ukeep = np.ascontiguousarray(ukeep, dtype=np.float64)
            up = ukeep.ctypes.data
        self._ck(self._lib.reth_replay_sample_global(self._h, int(global_batch), rp, int(n_shards),
                                                     int(shard), up, uloc, int(seed) & (2**64 - 1),
                                                     float(beta), int(slot), self._s(stream)))
        self._keep = (rkeep, ukeep)

    def sample_global_multi(self, global_batch, all_roots, n_shards, shard, first_slot, n_slots, beta=0.4,
                            uniforms_dev=None, seed=0, stream=None):
        """Mode G prefetch group: n_slots global stratified draws in one launch; this shard's shares are packed
        back to back from slot first_slot on (gather_multi / update_slots over the same slots handle the packed
        batch; slot_quota gives the per-draw {first, count, offset})."""
        rp, rloc, rdt, rcnt, rkeep = _ptr_loc(all_roots)
        assert rloc == DEVICE and rdt == np.dtype(np.float64) and rcnt >= 2 * n_shards
        up, uloc, udt, cnt, ukeep = _ptr_loc(uniforms_dev)
        if uniforms_dev is not None:
            assert uloc == DEVICE and udt == np.dtype(np.float64) and cnt >= global_batch * n_slots
        self._ck(self._lib.reth_replay_sample_global_multi(self._h, int(global_batch), rp, int(n_shards), int(shard),
                                                           int(first_slot), int(n_slots), up,
                                                           int(seed) & (2**64 - 1), float(beta), self._s(stream)))
        self._keep = (rkeep, ukeep)

    def slot_quota(self, first_slot, n_slots, stream=None):
        """-> int64[n_slots, 3]: {first, count, packed offset} of each draw of a packed Mode G group."""
        out = (C.c_int64 * (4 * n_slots))()
        self._ck(self._lib.reth_replay_slot_quota(self._h, int(first_slot), int(n_slots), out, self._s(stream)))
        return np.array(out, dtype=np.int64).reshape(n_slots, 4)[:, :3]

    def sample_multi(self, batch, first_slot, n_slots, beta=0.4, uniforms_dev=None, seed=0, stream=None):
        """n_slots independent stratified draws in one launch (prefetch group)."""
        up, uloc, udt, cnt, ukeep = _ptr_loc(uniforms_dev)
        if uniforms_dev is not None:
            assert uloc == DEVICE and udt == np.dtype(np.float64) and cnt >= batch * n_slots
        self._ck(self._lib.reth_replay_sample_multi(self._h, int(batch), int(first_slot), int(n_slots), up,
                                                    int(seed) & (2**64 - 1), float(beta), self._s(stream)))
        self._keep = ukeep

    def gather_multi(self, first_slot, n_slots, stream=None):
        self._ck(self._lib.reth_replay_gather_multi(self._h, int(first_slot), int(n_slots), self._s(stream)))

    def update_slots(self, first_slot, n_slots, prio, alpha=0.6, transform=True, stream=None):
        p, pdt, ploc, pkeep = self._prio(prio, self.slot_batch(first_slot) * n_slots)
        self._ck(self._lib.reth_replay_update_slots(self._h, int(first_slot), int(n_slots), p, pdt, ploc,
                                                    float(alpha), int(bool(transform)), self._s(stream)))
        self._keep = pkeep

    def gather(self, slot=0, stream=None):
        self._ck(self._lib.reth_replay_gather(self._h, int(slot), self._s(stream)))

    def slot_batch(self, slot):
        b = C.c_int32()
        self._ck(self._lib.reth_replay_slot_ptrs(self._h, slot, None, None, None, None, C.byref(b)))
        return b.value

    def slot_arrays(self, slot, rows=None):
        """Device views of a batch slot (borrowed until the slot is re-used):
        (columns, indices, weights, priorities) as __cuda_array_interface__ objects."""
        cols = (C.c_void_p * max(1, len(self.struct)))()
        ip, wp, pp, b = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_int32()
        self._ck(self._lib.reth_replay_slot_ptrs(self._h, slot, cols, C.byref(ip), C.byref(wp),
                                                 C.byref(pp), C.byref(b)))
        n = b.value if rows is None else int(rows)
        data = []
        if self.store_rows:
            data = [_DevArray(cols[i], (n,) + sh, dt, self) for i, (dt, sh) in enumerate(self.out_struct)]
        return (data, _DevArray(ip.value, (n,), np.int64, self), _DevArray(wp.value, (n,), np.float64, self),
                _DevArray(pp.value, (n,), np.float64, self))

    def slot_tensors(self, slot, rows=None):
        """Same as slot_arrays but as torch CUDA tensors (zero-copy hand-off to the learner)."""
        import torch
        dev = torch.device("cuda", self.device)
        data, idx, w, p = self.slot_arrays(slot, rows)
        t = lambda a: torch.as_tensor(a, device=dev)  # noqa: E731
        return [t(a) for a in data], t(idx), t(w), t(p)

    def slot_read(self, slot, with_rows=False, rows=None, stream=None):
        """Host copies of a slot: (columns or None, indices int64, weights float64). Synchronises."""
        n = self.slot_batch(slot)
        idx = np.empty(n, np.int64)
        w = np.empty(n, np.float64)
        ptrs, outs = None, None
        if with_rows and self.store_rows:
            outs = [np.empty((n,) + sh, dt) for dt, sh in self.out_struct]
            ptrs = (C.c_void_p * len(outs))(*[o.ctypes.data for o in outs])
        self._ck(self._lib.reth_replay_slot_read(self._h, slot, idx.ctypes.data, w.ctypes.data, ptrs,
                                                 self._s(stream)))
        if rows is not None:
            idx, w = idx[:rows], w[:rows]
            outs = [o[:rows] for o in outs] if outs is not None else None
        return outs, idx, w

    def group_read(self, first_slot, n_slots, with_weights=False, stream=None):
        """Host indices (and weights) of n_slots consecutive slots: one D2H copy, one sync."""
        n = self.slot_batch(first_slot) * n_slots
        idx = np.empty(n, np.int64)
        w = np.empty(n, np.float64) if with_weights else None
        self._ck(self._lib.reth_replay_group_read(self._h, int(first_slot), int(n_slots), idx.ctypes.data,
                                                  w.ctypes.data if with_weights else None, self._s(stream)))
        return idx, w

    def slot_meta(self, slot, stream=None):
        fc = (C.c_int64 * 2)()
        self._ck(self._lib.reth_replay_slot_meta(self._h, slot, fc, self._s(stream)))
        return fc[0], fc[1]

    # -- non-PER sampler plugins (uniform_sampler.py, fifo_sampler.py) --------
    def _idx_arg(self, indices):
        ip, iloc, idt, n, keep = _ptr_loc(indices)
        if idt not in (np.dtype(np.int32), np.dtype(np.int64)):
            if iloc == DEVICE:
                raise ValueError("device indices must be int32 or int64")
            keep = np.ascontiguousarray(keep, dtype=np.int64)
            ip, idt = keep.ctypes.data, keep.dtype
        return ip, (I32 if idt.itemsize == 4 else I64), iloc, n, keep

    def uniform_note(self, indices=None, first_slot=0, n=None, stream=None):
        """UniformSampler.update: indices, or (indices=None) the FIFO slots first_slot.. of an append."""
        if indices is None:
            self._ck(self._lib.reth_replay_uniform_note(self._h, None, I64, HOST, int(first_slot), int(n), self._s(stream)))
            return
        ip, idt, iloc, cnt, keep = self._idx_arg(indices)
        self._ck(self._lib.reth_replay_uniform_note(self._h, ip, idt, iloc, 0, cnt, self._s(stream)))
        self._keep = keep

    def sample_uniform(self, batch, first_slot=0, n_slots=1, draws=None, seed=0, stream=None):
        dp, dloc, keep = None, HOST, None
        if draws is not None:
            dp, _, dloc, cnt, keep = self._idx_arg(draws if _is_torch(draws) else np.ascontiguousarray(draws, dtype=np.int64))
            assert cnt == batch * n_slots
        self._ck(self._lib.reth_replay_sample_uniform(self._h, int(batch), int(first_slot), int(n_slots), dp, dloc,
                                                      int(seed) & (2**64 - 1), self._s(stream)))
        self._keep = keep

    def fifo_push(self, indices, weights, first_slot=0, stream=None):
        """FIFOSampler.update: (index, weight) pairs; indices=None: the FIFO slots first_slot.. of an append."""
        p, pdt, ploc, pkeep = self._prio(weights, len(weights))
        if indices is None:
            self._ck(self._lib.reth_replay_fifo_push(self._h, None, I64, HOST, int(first_slot), p, pdt, ploc,
                                                     len(weights), self._s(stream)))
        else:
            ip, idt, iloc, cnt, keep = self._idx_arg(indices)
            assert cnt == len(weights)
            self._ck(self._lib.reth_replay_fifo_push(self._h, ip, idt, iloc, 0, p, pdt, ploc, cnt, self._s(stream)))
            self._keep = keep
        self._keep2 = pkeep

    def sample_fifo(self, batch, first_slot=0, n_slots=1, stream=None):
        self._ck(self._lib.reth_replay_sample_fifo(self._h, int(batch), int(first_slot), int(n_slots), self._s(stream)))

    def sampler_state(self):
        out = (C.c_int64 * 4)()
        self._ck(self._lib.reth_replay_sampler_state(self._h, out))
        return {"uniform_tail": out[0], "fifo_len": out[1], "fifo_head": out[2]}

    def pipe_sampler(self, kind):
        self._ck(self._lib.reth_replay_pipe_sampler(self._h, {"per": 0, "uniform": 1, "fifo": 2}[kind]))

    def ipc_export(self):
        """bytes descriptor of the batch ring (CUDA IPC handles) for IpcRing in another process."""
        ex = _cabi.IpcExport()
        self._ck(self._lib.reth_replay_ipc_export(self._h, C.byref(ex)))
        return bytes(ex)

    # ---- shard roots over NVLink (Mode G without a host-driven collective)
    def board_export(self):
        """bytes descriptor of this shard's roots board, for board_connect() on every shard of the group."""
        ex = _cabi.BoardExport()
        self._ck(self._lib.reth_replay_board_export(self._h, C.byref(ex)))
        return bytes(ex)

    def board_connect(self, blobs, shard):
        """blobs: board_export() of every shard in rank order (own entry included)."""
        arr = (_cabi.BoardExport * len(blobs))(*[_cabi.BoardExport.from_buffer_copy(b) for b in blobs])
        self._ck(self._lib.reth_replay_board_connect(self._h, arr, len(blobs), int(shard)))

    def roots_exchange(self, phases=3, stream=None):
        """ONE kernel on `stream`: store this shard's root {total, min} into every peer's board over NVLink
        (phase 1) and wait until all roots of this exchange arrived in the own board (phase 2); the snapshot
        is all_roots_dev().  Collective: every shard issues the same sequence."""
        self._ck(self._lib.reth_replay_roots_exchange(self._h, int(phases), self._s(stream)))

    def all_roots_dev(self, n_shards):
        p = C.c_void_p()
        self._ck(self._lib.reth_replay_all_roots_dev(self._h, C.byref(p)))
        return _DevArray(p.value, (2 * int(n_shards),), np.float64, self)

    def gather_to(self, slot, ring, dst_slot, stream=None):
        """After sample_global(slot): gather this shard's rows, indices and weights straight into rows
        [first, first+count) of batch slot `dst_slot` of a peer's ring (NVLink stores from the kernel);
        ring=None: this shard's own ring."""
        self._ck(self._lib.reth_replay_gather_to(self._h, int(slot), ring._v if ring is not None else None,
                                                 int(dst_slot), self._s(stream)))

    def gather_from(self, dst_slot, index_shard, src_slot, stream=None):
        """Rows, indices and weights of the batch `index_shard` (a tree-only shard over the same slots:
        another sampler topic) drew into its slot src_slot -> this shard's batch slot dst_slot."""
        self._ck(self._lib.reth_replay_gather_from(self._h, int(dst_slot), index_shard._h, int(src_slot), self._s(stream)))

    def pipe_reserve(self, n_slots):
        """Keep the last n_slots batch slots out of the loader pipe (other sampler topics use them)."""
        self._ck(self._lib.reth_replay_pipe_reserve(self._h, int(n_slots)))

    def open_pipe(self, batch, group, seed=0):
        return ShardPipe(self, batch, group, seed)

    @property
    def pipe_stream(self):
        """The shard's own ordering stream (tree updates + sampling) as a raw cudaStream_t."""
        p = C.c_void_p()
        self._ck(self._lib.reth_replay_pipe_stream(self._h, C.byref(p)))
        return p.value

    # -- parity / debug ---------------------------------------------------
    def tree_arrays(self, stream=None):
        """(_sum, _min, _val) in the reference's heap order (utils/sumtree.py:85-88)."""
        s, m, v = (np.empty(self.capacity, np.float64) for _ in range(3))
        self._ck(self._lib.reth_replay_tree_read(self._h, s.ctypes.data, m.ctypes.data, v.ctypes.data,
                                                 self._s(stream)))
        return s, m, v

    def root(self, stream=None):
        t, m = C.c_double(), C.c_double()
        self._ck(self._lib.reth_replay_tree_root(self._h, C.byref(t), C.byref(m), self._s(stream)))
        return t.value, m.value

    def root_dev(self):
        p = C.c_void_p()
        self._ck(self._lib.reth_replay_root_dev(self._h, C.byref(p)))
        return _DevArray(p.value, (2,), np.float64, self)

    def ring_read(self, column, first_slot, n, stream=None):
        dt, sh = self.struct[column]
        out = np.empty((n,) + sh, dt)
        self._ck(self._lib.reth_replay_ring_read(self._h, column, first_slot, n, out.ctypes.data,
                                                 self._s(stream)))
        return out

    def stats(self):
        st = _cabi.Stats()
        self._ck(self._lib.reth_replay_stats_get(self._h, C.byref(st)))
        return {n: getattr(st, n) for n, _ in st._fields_}

    @property
    def size(self):
        return self.stats()["size"]

    @property
    def tail(self):
        return self.stats()["tail"]

    def step_loop(self, steps, batch, prio_dev, uniforms_dev=None, seed=0, alpha=0.6, beta=0.4,
                  gather=True, serial=False, time_gather=False, sample=True, update=True, group=1,
                  one_gather_stream=False, global_mode=False, stream=None, ingest=None):
        """ingest = (device columns, device float32 priorities): after every group of `group` learner steps those
        rows are appended without draining the loop (reth_replay_step_loop_ingest, BASELINE config 4)."""
        pp, ploc, pdt, pcnt, pk = _ptr_loc(prio_dev)
        assert ploc == DEVICE and pdt == np.dtype(np.float32) and pcnt >= steps * batch
        up, uloc, udt, ucnt, uk = _ptr_loc(uniforms_dev)
        if uniforms_dev is not None:
            assert uloc == DEVICE and udt == np.dtype(np.float64) and ucnt >= steps * batch
        flags = (0 if gather else 1) | (2 if serial else 0) | (4 if time_gather else 0) | (0 if sample else 8) | (0 if update else 16) | (32 if one_gather_stream else 0) | (64 if global_mode else 0) | ((int(group) & 0xff) << 8)
        if ingest is not None:
            icols, iprio = ingest
            ip, iloc, idt, n, ik = _ptr_loc(iprio)
            assert iloc == DEVICE and idt == np.dtype(np.float32)
            ptrs, cloc, ckeep = self._cols(icols, n)
            assert cloc == DEVICE
            self._ck(self._lib.reth_replay_step_loop_ingest(self._h, int(steps), int(batch), pp, up,
                                                            int(seed) & (2**64 - 1), float(alpha), float(beta), flags,
                                                            ptrs, ip, n, self._s(stream)))
            self._keep = (pk, uk, ik, ckeep)
            return
        self._ck(self._lib.reth_replay_step_loop(self._h, int(steps), int(batch), pp, up,
                                                 int(seed) & (2**64 - 1), float(alpha), float(beta),
                                                 flags, self._s(stream)))
        self._keep = (pk, uk)

    def gather_time(self):
        """(total_ms, launches) of the event-bracketed gathers since the last call."""
        ms, n = C.c_double(), C.c_int64()
        self._ck(self._lib.reth_replay_gather_time(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def debug_stamps(self):
        out = (C.c_int64 * 32)()
        self._ck(self._lib.reth_replay_debug_stamps(self._h, out))
        return list(out)
