"""n-step assembly on the device (reth_replay_append_nstep, csrc/nstep.cuh) through the C ABI: against the
tapes recorded from the reference's own NStepAdder (reth/reth/utils/nstep_adder.py:11-28) and against the
oracle's closed form on Atari-shaped rows, several interleaved environment streams, ragged call sizes, host
and device sources, a wrapping ring, and frame-reference rows."""
import numpy as np
import pytest

from oracle import per_oracle as po
from test_frames_cpu import TAPES, tape_columns

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def Shard():
    import __graft_entry__ as entry
    entry.build()
    from reth_b200 import ReplayShard
    return ReplayShard


def _ragged(n, rng, hi):
    at, out = 0, []
    while at < n:
        m = int(min(n - at, rng.integers(1, hi)))
        out.append((at, at + m))
        at += m
    return out


@pytest.mark.parametrize("name", TAPES)
@pytest.mark.parametrize("source", ["host", "device"])
def test_nstep_kernel_matches_reference_tape(Shard, name, source):
    gamma, step, n, cols, want = tape_columns(name)
    struct = [(c.dtype, ()) for c in cols]
    cap = 512
    sh = Shard(cap, struct, max_batch=8)
    sh.nstep_open(2, step, gamma, max_rows=16)                     # max_rows < some calls: staged in pieces
    rng = np.random.default_rng(5)
    prio = (rng.random(n) + 0.1).astype(np.float32)
    stored = 0
    for lo, hi in _ragged(n, rng, 41):
        part = [c[lo:hi] for c in cols]
        if source == "device":
            import torch
            part = [torch.from_numpy(np.ascontiguousarray(c)).cuda() for c in part]
        first, emitted = sh.append_nstep(1, part, prio[lo:hi], alpha=1.0, transform=False)
        assert first == stored % cap
        assert emitted == max(0, hi - step) - max(0, lo - step)   # push i emits row i - step
        stored += emitted
    w = want[step:]
    assert stored == len(w) == n - step
    got = [sh.ring_read(c, 0, stored) for c in range(5)]
    assert np.array_equal(got[0], w[:, 1].astype(np.int64)) and np.array_equal(got[1], w[:, 2].astype(np.int64))
    assert np.array_equal(got[3], w[:, 4].astype(np.int64))
    assert np.array_equal(got[4].astype(np.int64), w[:, 5].astype(np.int64))
    assert np.array_equal(got[2].astype(np.float64), w[:, 3])     # bit-exact rewards (float64 and float32 tapes)
    # the row leaving at push p carries prio[p]
    _, _, val = sh.tree_arrays()
    assert np.array_equal(val[:stored], prio[step:].astype(np.float64))
    assert not val[stored:].any()


def test_nstep_streams_wide_rows_and_wrap(Shard):
    """Three interleaved streams of Atari-shaped rows (28 224 B columns are cut into parts) into a ring that wraps;
    expected ring content = the oracle's closed form per stream, laid out in call order."""
    struct = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
    cap, step, gamma, n_streams, n = 96, 3, 0.99, 3, 70
    sh = Shard(cap, struct, max_batch=8)
    sh.nstep_open(n_streams, step, gamma, max_rows=32)
    rng = np.random.default_rng(9)
    streams = []
    for s in range(n_streams):
        streams.append([rng.integers(0, 256, (n, 4, 84, 84), dtype=np.uint8), rng.integers(0, 18, n),
                        rng.standard_normal(n).astype(np.float32), rng.integers(0, 256, (n, 4, 84, 84), dtype=np.uint8),
                        (rng.random(n) < 0.15).astype(np.float32)])
    folded = [po.nstep_fold(cols, step, gamma)[0] for cols in streams]
    ring = [np.zeros((cap,) + sh_, dt) for dt, sh_ in struct]
    tail, sent, out = 0, [0] * n_streams, [0] * n_streams
    while min(sent) < n:
        s = int(rng.integers(0, n_streams))
        if sent[s] >= n:
            continue
        m = int(min(n - sent[s], rng.integers(1, 25)))
        lo, hi = sent[s], sent[s] + m
        first, emitted = sh.append_nstep(s, [c[lo:hi] for c in streams[s]], None)
        assert first == tail
        for e in range(emitted):
            for c in range(5):
                ring[c][(tail + e) % cap] = folded[s][c][out[s] + e]
        tail = (tail + emitted) % cap
        out[s] += emitted
        sent[s] = hi
    assert sum(out) == n_streams * (n - step) > cap               # the ring wrapped
    for c in range(5):
        assert np.array_equal(sh.ring_read(c, 0, cap), ring[c]), c
    # a reset stream starts a fresh adder: its buffered rows never reach the ring
    sh.nstep_reset(0)
    first, emitted = sh.append_nstep(0, [c[:step] for c in streams[0]], None)
    assert emitted == 0
    assert sh.stats()["appended_rows"] == sum(out)


def test_nstep_frame_reference_rows(Shard):
    """De-duplicated layout: the n-step s1 is the k frame slots of a later row; the gather then rebuilds the stack."""
    k, shape, step, gamma, n = 4, (84, 84), 3, 0.9, 40
    struct = [(np.uint8, (k,) + shape), (np.int64, ()), (np.float64, ()), (np.uint8, (k,) + shape), (np.float32, ())]
    sh = Shard(64, struct, max_batch=16, frames={"shape": shape, "capacity": 256, "columns": [0, 3]})
    sh.nstep_open(1, step, gamma, max_rows=64)
    rng = np.random.default_rng(11)
    frames = rng.integers(0, 256, (n + k,) + shape, dtype=np.uint8)
    f0 = sh.append_frames(frames)
    refs = (f0 + np.arange(n)[:, None] + np.arange(k)[None, :]).astype(np.int32)       # s0 of step i = frames i..i+k-1
    cols = [refs, rng.integers(0, 6, n), rng.standard_normal(n), refs + 1, (rng.random(n) < 0.2).astype(np.float32)]
    want, _ = po.nstep_fold(cols, step, gamma)
    first, emitted = sh.append_nstep(0, cols, np.ones(n, np.float32), alpha=1.0, transform=False)
    assert (first, emitted) == (0, n - step)
    for c in range(5):
        assert np.array_equal(sh.ring_read(c, 0, emitted), want[c]), c
    u = (np.arange(16) + 0.5) / 16
    sh.sample(16, beta=0.4, uniforms=u, slot=0)
    sh.gather(0)
    data, idx, _ = sh.slot_read(0, with_rows=True)
    assert np.array_equal(data[3], frames[want[3][idx] - f0])       # the stack of the row folded in last
    assert np.array_equal(data[0], frames[want[0][idx] - f0])
    assert np.array_equal(data[2], want[2][idx])


def test_nstep_argument_errors(Shard):
    sh = Shard(64, [(np.int64, ()), (np.int64, ()), (np.float32, ()), (np.int64, ()), (np.float32, ())], max_batch=8)
    from reth_b200 import ReplayError
    with pytest.raises(ReplayError):
        sh.append_nstep(0, [np.zeros(2, np.int64)] * 5, None)      # not open
    with pytest.raises(ReplayError):
        sh.nstep_open(1, 3, 0.99, r_col=3)                         # reward and s1 must be different columns
    sh.nstep_open(2, 3, 0.99)
    with pytest.raises(ReplayError):
        sh.nstep_open(2, 3, 0.99)
    with pytest.raises(ReplayError):
        sh.append_nstep(2, [np.zeros(2, np.int64), np.zeros(2, np.int64), np.zeros(2, np.float32), np.zeros(2, np.int64),
                            np.zeros(2, np.float32)], None)


@pytest.mark.parametrize("layout", ["inproc", "process"])
def test_nstep_through_the_service_surface(layout):
    """start_per(..., nstep={"step", "gamma"}): producers send RAW 1-step rows through Client.append, every Client is
    one environment stream (as every reference worker owns one NStepAdder, test/apex-dqn/worker.py:35,53), and what the
    loaders sample are the n-step transitions the reference's adder would have emitted, in arrival order."""
    import __graft_entry__ as entry
    entry.build()
    from reth_b200 import service, service_mp
    mod = service if layout == "inproc" else service_mp
    step, gamma, n, B, cap = 3, 0.99, 120, 16, 1024
    rng = np.random.default_rng(31)
    streams = []
    for s in range(2):
        ids = 1000 * s + np.arange(n)
        streams.append([np.repeat(ids[:, None], 8, 1).astype(np.float32), rng.integers(0, 6, n), rng.standard_normal(n).astype(np.float32),
                        np.repeat((ids + 1)[:, None], 8, 1).astype(np.float32), (rng.random(n) < 0.1).astype(np.float32)])
    folded = [po.nstep_fold(cols, step, gamma)[0] for cols in streams]
    proc, addr = mod.start_per(cap, B, sample_start=B, nstep={"step": step, "gamma": gamma, "streams": 4})
    try:
        clients = [mod.Client(addr), mod.Client(addr)]
        loader = mod.NumpyLoader(addr)
        want = [[] for _ in range(5)]
        sent, out = [0, 0], [0, 0]
        while min(sent) < n:
            s = int(rng.integers(0, 2))
            if sent[s] >= n:
                continue
            m = int(min(n - sent[s], rng.integers(1, 20)))
            lo, hi = sent[s], sent[s] + m
            clients[s].append([c[lo:hi] for c in streams[s]], np.ones(m, np.float32))
            if layout == "process":
                loader.drain()                         # two connections: arrival order = send order only when serialised
            emitted = max(0, hi - step) - max(0, lo - step)
            for c in range(5):
                want[c].append(folded[s][c][out[s]:out[s] + emitted])
            out[s] += emitted
            sent[s] = hi
        want = [np.concatenate(w) for w in want]
        total = len(want[0])
        assert total == 2 * (n - step)
        if layout == "process":
            assert loader.drain()["cnt"] == total
            got = [loader.ring_read(c, 0, total) for c in range(5)]
        else:
            eng = service._ENGINES[addr]
            eng.drain()
            assert eng.cnt == total
            got = [eng.shard.ring_read(c, 0, total) for c in range(5)]
        for c in range(5):
            assert np.array_equal(got[c], want[c]), c
        data, indices, weights = loader.sample()
        assert np.array_equal(data[0], want[0][indices]) and np.array_equal(data[2], want[2][indices])
        assert np.array_equal(data[3], want[3][indices])
        loader.close()
    finally:
        proc.terminate()
        proc.join()


def test_nstep_service_streams_are_recycled():
    """More producers over time than n-step streams: the least recently used stream starts over for the newcomer (a
    restarted actor connects with a new id), its unfinished rows are dropped."""
    import __graft_entry__ as entry
    entry.build()
    from reth_b200 import service
    step, gamma, n = 3, 0.9, 10
    rng = np.random.default_rng(2)
    proc, addr = service.start_per(256, 8, sample_start=8, nstep={"step": step, "gamma": gamma, "streams": 2})
    try:
        eng = service._ENGINES[addr]
        want = [[] for _ in range(5)]
        for who in range(4):                                  # four producers, two streams
            ids = 100 * who + np.arange(n)
            cols = [ids.astype(np.float32), rng.integers(0, 4, n), rng.standard_normal(n).astype(np.float32),
                    (ids + 1).astype(np.float32), np.zeros(n, np.float32)]
            service.Client(addr).append(cols, np.ones(n, np.float32))
            for c, col in enumerate(po.nstep_fold(cols, step, gamma)[0]):
                want[c].append(col[:n - step])
        eng.drain()
        assert eng.cnt == 4 * (n - step) and len(eng._nstep_streams) == 2
        for c in range(5):
            assert np.array_equal(eng.shard.ring_read(c, 0, eng.cnt), np.concatenate(want[c])), c
    finally:
        proc.terminate()
