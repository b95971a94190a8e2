"""The Uniform and FIFO sampler plugins on the device against the golden
vectors of the reference's own classes and against the oracle on larger tapes, incl. the gather
of the rows they select and their use through the asynchronous pipe."""
import numpy as np
import pytest

from oracle import per_oracle as po

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import __graft_entry__ as entry
    entry.build()
    from reth_b200 import ReplayShard, sampler
    return ReplayShard, sampler


def test_plugins_golden(mods, golden):
    _, sampler = mods
    g = golden("sampler_plugins")
    us, fs = sampler.UniformSampler(300, max_batch=64), sampler.FIFOSampler(50, max_batch=64)
    for k in range(5):
        us.update(g[f"msg{k}"], g[f"w{k}"])
        fs.update(g[f"msg{k}"], g[f"w{k}"])
        assert us.tail == int(g[f"u_tail{k}"])
        ui, uw = us.sample(16, draws=g[f"u_draws{k}"])
        assert np.array_equal(ui, g[f"u_idx{k}"]) and np.array_equal(uw, g[f"u_w{k}"]) and uw.dtype == np.int64
        if fs.ready_sample(8):
            fi, fw = fs.sample(8)
            assert np.array_equal(fi, g[f"f_idx{k}"]) and np.array_equal(fw, g[f"f_w{k}"])
        assert len(fs) == int(g[f"f_len{k}"])
    us.update(np.arange(5), np.ones(5))
    assert us.tail == int(g["u_tail_final"])
    from reth_b200 import ReplayError
    with pytest.raises(ReplayError):
        sampler.FIFOSampler(10, max_batch=8).sample(4)             # nothing queued: ready_sample is False
    us.clear()
    assert us.tail == 0
    # a message that crosses the capacity fills the log and raises like the reference (uniform_sampler.py:20-22)
    us.update(np.arange(290), np.ones(290))
    with pytest.raises(IndexError):
        us.update(np.arange(20), np.ones(20))
    assert us.tail == 300
    us.update(np.arange(5), np.ones(5))                            # a full log ignores further messages (:18-19)
    assert us.tail == 300


def test_plugins_with_rows_and_pipe(mods):
    """Appended rows notified to both samplers by slot range; draws gathered by the common gather kernel;
    FIFO through the asynchronous pipe hands every row out exactly once, oldest first."""
    ReplayShard, _ = mods
    cap, B, G = 4096, 64, 2
    struct = [(np.uint8, (4, 84, 84)), (np.int64, ())]
    rng = np.random.default_rng(8)
    sh = ReplayShard(cap, struct, max_batch=B, ring_depth=3 * G)
    ring = po.OracleRing(cap, struct)
    ous, ofs = po.OracleUniformSampler(cap), po.OracleFIFOSampler(cap)
    for n in (1000, 2000, 1500):                                   # the ring wraps; the uniform log fills up
        cols = [rng.integers(0, 256, (n, 4, 84, 84), dtype=np.uint8), rng.integers(0, 1 << 40, n)]
        w = rng.random(n).astype(np.float32)
        first = sh.append(cols, w)
        slots = ring.append_batch(cols)
        sh.uniform_note(None, first_slot=first, n=n)
        sh.fifo_push(None, w, first_slot=first)
        ous.update(slots)
        ofs.update(slots, w)
    st = sh.sampler_state()
    assert st["uniform_tail"] == ous.tail == cap and st["fifo_len"] == len(ofs.buffer) == cap
    draws = rng.integers(0, ous.tail, B * G)
    sh.sample_uniform(B, 0, G, draws=draws)
    sh.gather_multi(0, G)
    for g in range(G):
        data, idx, w = sh.slot_read(g, with_rows=True)
        oi, _ = ous.sample(draws[g * B:(g + 1) * B])
        assert np.array_equal(idx, oi) and np.all(w == 1.0)
        for got, want in zip(data, ring.select(oi)):
            assert np.array_equal(got, want)
    sh.sample_uniform(B, 0, 1, seed=5)                             # device generator: positions inside the log
    _, idx, _ = sh.slot_read(0)
    assert np.all(np.isin(idx, ous.indices[:ous.tail]))
    # FIFO through the pipe
    sh.pipe_sampler("fifo")
    pipe = sh.open_pipe(B, G)
    pipe.issue(0.4)
    for it in range(5):
        pipe.issue(0.4)
        g = pipe.wait()
        oi, ow = ofs.sample(B * G)
        assert np.array_equal(pipe.idx_host[g].reshape(-1), oi)
        assert np.array_equal(pipe.w_host[g].reshape(-1), ow)
        data, _, _, _ = sh.slot_tensors(g * G)
        assert np.array_equal(data[1].cpu().numpy(), ring.select(oi[:B])[1])
    pipe.close()
    sh.close()
