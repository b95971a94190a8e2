"""Host logic of the in-process engine (reth_b200/service._Engine) on the CPU, around a stand-in shard that records
the calls it receives: what is pinned is the ORDER of shard calls and the bookkeeping (the arithmetic of those calls
is the GPU tests' business):
  * staged priority write-backs land before an append that may recycle their slots;
  * n-step assembly behind the service surface: one stream per producer, least recently used stream recycled;
  * `cnt` counts what the reference's sampler loop counts (server/sampler_loop.py:20,36), the schedules step once
    per update message (per_sampler.py:30-32)."""
import numpy as np
import pytest


class _Pipe:
    """ShardPipe's staging surface: one entry of group * batch indices / priorities at a time."""

    def __init__(self, shard, batch, group):
        self.shard, self.B, self.G = shard, batch, group
        self.upd_idx = np.zeros((4, batch * group), np.int64)
        self.upd_p32 = np.zeros((4, batch * group), np.float32)
        self.upd_p64 = np.zeros((4, batch * group), np.float64)
        self.n_groups, self._e = 3, -1

    def update_begin(self):
        self._e = (self._e + 1) % 4
        return self._e

    def update_commit(self, e, n, f32, alpha):
        self.shard.calls.append(("commit", [int(i) for i in self.upd_idx[e, :n:self.B]], n, bool(f32)))

    def drain(self):
        pass

    def close(self):
        self.shard.calls.append(("pipe_close",))


class _Shard:
    made = []

    def __init__(self, capacity, struct, **kw):
        self.capacity, self.struct, self.kw = capacity, struct, kw
        self.calls, self.tail, self.pipe_stream, self.ring_depth = [], 0, 1234, kw.get("ring_depth", 4)
        self._nstep = None
        _Shard.made.append(self)

    def pipe_sampler(self, kind):
        self.calls.append(("pipe_sampler", kind))

    def append(self, cols, prio, alpha=0.6, transform=True, stream=None):
        assert stream == self.pipe_stream
        first, n = self.tail, len(prio)
        self.tail = (self.tail + n) % self.capacity
        self.calls.append(("append", first, n))
        return first

    def update(self, idx, prio, alpha=0.6, transform=True, stream=None):
        self.calls.append(("update", [int(i) for i in idx]))

    def nstep_open(self, streams, step, gamma, **kw):
        self._nstep = (streams, step, gamma, kw)
        self.calls.append(("nstep_open", streams, step))

    def nstep_reset(self, sid):
        self.calls.append(("nstep_reset", sid))

    def append_nstep(self, sid, cols, prio, alpha=0.6, transform=True, stream=None):
        n = len(prio)
        self.calls.append(("append_nstep", sid, n))
        return self.tail, max(0, n - 1)

    def open_pipe(self, batch, group, seed=0):
        self.calls.append(("open_pipe", group))
        return _Pipe(self, batch, group)

    def close(self):
        pass


@pytest.fixture
def service(monkeypatch):
    from reth_b200 import service as svc
    _Shard.made = []
    monkeypatch.setattr(svc, "ReplayShard", _Shard)
    return svc


def _rows(n, first=0):
    ids = first + np.arange(n)
    return [ids.astype(np.float32), ids, np.ones(n, np.float32), ids.astype(np.float32) + 1, np.zeros(n, np.float32)]


def test_staged_write_backs_land_before_an_append(service):
    B = 4
    proc, addr = service.start_per(64, B, sample_start=B)
    eng = service._ENGINES[addr]
    client = service.Client(addr)
    client.append(_rows(16), np.ones(16, np.float32))
    sh = _Shard.made[0]
    assert sh.kw["max_batch"] == B and sh.calls[:2] == [("pipe_sampler", "per"), ("append", 0, 16)]
    assert eng.cnt == 16 and eng.stream == sh.pipe_stream
    eng.open_pipe(4, 1)                                       # a loader with prefetch group 4
    eng._ensure_pipe()
    # three learner batches staged (a group is four): nothing committed yet
    for k in range(3):
        client.update_priorities(np.arange(k * B, (k + 1) * B), np.ones(B, np.float32))
    assert not [c for c in sh.calls if c[0] == "commit"] and eng._upd_k == 3 and eng.cnt == 16 + 3 * B
    # an append may recycle those slots: the staged entry is committed first
    client.append(_rows(8, 100), np.ones(8, np.float32))
    assert sh.calls[-2:] == [("commit", [0, 4, 8], 3 * B, True), ("append", 16, 8)] and eng._upd_k == 0
    # a full group commits by itself; a write-back that is not a learner batch goes through on its own, behind it
    for k in range(4):
        client.update_priorities(np.arange(k * B, (k + 1) * B), np.ones(B, np.float32))
    assert sh.calls[-1] == ("commit", [0, 4, 8, 12], 4 * B, True)
    client.update_priorities(np.array([1, 2]), np.ones(2, np.float32))
    client.update_priorities(np.array([3]), np.ones(1))
    assert sh.calls[-2:] == [("update", [1, 2]), ("update", [3])]
    # a float64 message does not share a staging entry with float32 ones
    client.update_priorities(np.arange(B), np.ones(B, np.float32))
    client.update_priorities(np.arange(B), np.ones(B, np.float64))
    assert sh.calls[-1] == ("commit", [0], B, True) and eng._upd_k == 1 and eng._upd_f32 is False
    proc.terminate()
    assert not proc.is_alive() and ("pipe_close",) in sh.calls


def test_schedules_step_once_per_update_message(service):
    from reth_b200.sampler import PERSampler
    samplers = [{"sampler_cls": PERSampler, "num_procs": 1, "sample_start": 4,
                 "kwargs": {"alpha": 0.5, "beta": "0.4,1.0,10"}}]    # utils/schedule.py:13-32: linear, 10 steps
    proc, addr = service.start_server(64, 4, samplers=samplers)
    eng = service._ENGINES[addr]
    client = service.Client(addr)
    client.append(_rows(8), np.ones(8, np.float32))
    b0 = eng.sampler["beta"].value()
    for _ in range(5):
        client.update_priorities(np.arange(4), np.ones(4, np.float32))
    assert np.isclose(eng.sampler["beta"].value(), b0 + 5 * (1.0 - 0.4) / 10)
    client.update_priorities(np.arange(4), np.ones(4, np.float32), step=False)     # an append notice does not step
    assert np.isclose(eng.sampler["beta"].value(), b0 + 5 * (1.0 - 0.4) / 10)
    proc.terminate()


def test_nstep_streams_one_per_producer_least_recently_used_recycled(service):
    proc, addr = service.start_per(256, 4, sample_start=4, nstep={"step": 3, "gamma": 0.9, "streams": 2})
    eng = service._ENGINES[addr]
    a, b, c = (service.Client(addr) for _ in range(3))
    a.append(_rows(5), np.ones(5, np.float32))
    sh = _Shard.made[0]
    assert sh.calls[-2:] == [("nstep_open", 2, 3), ("append_nstep", 0, 5)]
    assert sh._nstep[3] == {"r_col": 2, "s1_col": 3, "done_col": 4, "max_rows": 4096}
    b.append(_rows(5), np.ones(5, np.float32))
    a.append(_rows(2), np.ones(2, np.float32))                # a is now the most recently used
    assert [x for x in sh.calls if x[0] == "append_nstep"] == [("append_nstep", 0, 5), ("append_nstep", 1, 5), ("append_nstep", 0, 2)]
    assert eng.cnt == 4 + 4 + 1                               # rows that left the adder (the stand-in emits n - 1)
    c.append(_rows(3), np.ones(3, np.float32))                # a third producer: b's stream starts over for it
    assert sh.calls[-2:] == [("nstep_reset", 1), ("append_nstep", 1, 3)]
    b.append(_rows(3), np.ones(3, np.float32))                # b is back: now a is the least recently used
    assert sh.calls[-2:] == [("nstep_reset", 0), ("append_nstep", 0, 3)]
    assert len([x for x in sh.calls if x[0] == "nstep_open"]) == 1
    # raw engine appends without a producer id are refused: the fold needs to know whose rows these are
    with pytest.raises(ValueError, match="name their producer"):
        eng.append(_rows(2), np.ones(2, np.float32))
    proc.terminate()
