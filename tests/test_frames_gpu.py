"""Frame-de-duplicated storage on the GPU: single frames live once in a frame
ring, transitions store k frame slots for s0 / s1, the gather rebuilds the k-stacks.  The checker is
the layout the reference actually stores -- full stacks per transition (FrameStack,
reth/reth/env/util.py:179-209, appended as NumpyBuffer rows): what comes out of the gather must be
byte-identical to gathering those rows."""
from collections import deque

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

K, FRAME = 4, (84, 84)
STRUCT = [(np.uint8, (K,) + FRAME), (np.int64, ()), (np.float32, ()), (np.uint8, (K,) + FRAME), (np.float32, ())]


@pytest.fixture(scope="module")
def mods():
    import __graft_entry__ as entry
    entry.build()
    from reth_b200 import ReplayShard
    from reth_b200.frames import FrameRefWriter
    return ReplayShard, FrameRefWriter


def _episodes(rng, writer, n_steps, n_env=3, p_done=0.03):
    """Drive `writer` like n_env Atari actors and return the full-stack rows FrameStack would store."""
    full = []
    stacks = {}
    for env in range(n_env):
        f = rng.integers(0, 256, FRAME, dtype=np.uint8)
        stacks[env] = deque([f] * K, maxlen=K)
        writer.reset(env, f)
    for t in range(n_steps):
        env = t % n_env
        f = rng.integers(0, 256, FRAME, dtype=np.uint8)
        a, r, done = int(rng.integers(0, 18)), float(rng.standard_normal()), float(rng.random() < p_done)
        s0 = np.stack(stacks[env])
        stacks[env].append(f)
        full.append((s0, a, np.float32(r), np.stack(stacks[env]), np.float32(done)))
        writer.step(env, f, [a, r, done], prio=float(rng.random() + 0.1))
        if done:
            f = rng.integers(0, 256, FRAME, dtype=np.uint8)
            stacks[env] = deque([f] * K, maxlen=K)
            writer.reset(env, f)
        if t % 64 == 63:
            writer.flush()
    writer.flush()
    return full


@pytest.mark.parametrize("cast", [None, "f32", "f32/255"])
def test_gather_rebuilds_frame_stacks(mods, cast):
    ReplayShard, FrameRefWriter = mods
    rng = np.random.default_rng(17)
    cap, B, n = 2048, 128, 1500
    sh = ReplayShard(cap, STRUCT, max_batch=B, ring_depth=2,
                     frames={"shape": FRAME, "capacity": 4096, "columns": [0, 3]},
                     out_cast=[cast, None, None, cast, None] if cast else None)
    assert sh.struct[0] == (np.dtype(np.int32), (K,)) and sh.out_struct[0][1] == (K,) + FRAME
    full = _episodes(rng, FrameRefWriter(sh, k=K, frame_cols=(0, 3)), n)
    assert sh.size == n
    for it in range(4):
        sh.sample(B, uniforms=rng.random(B), slot=it % 2)
        sh.gather(it % 2)
        data, idx, w = sh.slot_read(it % 2, with_rows=True)
        want_s0 = np.stack([full[i][0] for i in idx])
        want_s1 = np.stack([full[i][3] for i in idx])
        if cast:
            scale = np.float32(255.0) if cast == "f32/255" else np.float32(1.0)
            want_s0 = want_s0.astype(np.float32) / scale
            want_s1 = want_s1.astype(np.float32) / scale
            assert data[0].dtype == np.float32
        assert np.array_equal(data[0], want_s0) and np.array_equal(data[3], want_s1)
        assert np.array_equal(data[1], [full[i][1] for i in idx])
        assert np.array_equal(data[2], np.array([full[i][2] for i in idx], np.float32))
        assert np.array_equal(data[4], np.array([full[i][4] for i in idx], np.float32))
    sh.close()


def test_frame_ring_wraps_and_group_gather(mods):
    """The frame ring wraps (slots reused after frame_capacity frames) and a prefetch group gathers
    several batches in one launch."""
    ReplayShard, FrameRefWriter = mods
    rng = np.random.default_rng(18)
    cap, B, G = 512, 64, 4
    sh = ReplayShard(cap, STRUCT, max_batch=B, ring_depth=G, frames={"shape": FRAME, "capacity": 700, "columns": [0, 3]})
    full = _episodes(rng, FrameRefWriter(sh, k=K, frame_cols=(0, 3)), 1200, p_done=0.0)     # wraps both rings
    # transitions are in slots (i mod cap); only the newest `cap` are alive and their frames are within the
    # last 700 frames (1 frame per step + K at reset)
    alive = {i % cap: i for i in range(len(full))}
    import torch
    u = torch.rand(G * B, device="cuda", dtype=torch.float64)
    sh.sample_multi(B, 0, G, uniforms_dev=u)
    sh.gather_multi(0, G)
    for g in range(G):
        data, idx, w = sh.slot_read(g, with_rows=True)
        rows = [full[alive[int(s)]] for s in idx]
        assert np.array_equal(data[0], np.stack([r[0] for r in rows]))
        assert np.array_equal(data[3], np.stack([r[3] for r in rows]))
        assert np.array_equal(data[1], [r[1] for r in rows])
    sh.close()


def test_frame_ring_wraps_before_row_ring_is_reported(mods):
    """the frame ring advances one frame per step plus one per reset, so with frame_capacity close to
    capacity old but still sampleable transitions can outlive their frames.  The gather must say so (RETH_ERR_STATE)
    instead of rebuilding stacks from overwritten slots; a big enough ring over the same steps stays clean."""
    ReplayShard, FrameRefWriter = mods
    from reth_b200 import ReplayError
    cap, B, n = 512, 64, 500
    for fcap, stale in ((300, True), (600, False)):
        rng = np.random.default_rng(21)
        sh = ReplayShard(cap, STRUCT, max_batch=B, ring_depth=2, frames={"shape": FRAME, "capacity": fcap, "columns": [0, 3]})
        full = _episodes(rng, FrameRefWriter(sh, k=K, frame_cols=(0, 3)), n, p_done=0.0)
        assert sh.size == n                                           # the ROW ring has not wrapped: all 500 rows are alive
        # stratified draw over all rows: the oldest ones refer to frames 500 + 3 - 300 steps behind the frame cursor
        sh.sample(B, uniforms=rng.random(B), slot=0)
        sh.gather(0)
        if stale:
            with pytest.raises(ReplayError) as ei:
                sh.slot_read(0, with_rows=True)
            assert ei.value.code == -5 and "overwritten" in str(ei.value)
            # nearly all priority on the newest rows: the same shard keeps serving the rows whose frames are alive
            sh.update(np.arange(n - 150), np.full(n - 150, 1e-12), transform=False)
            sh.sample(B, uniforms=rng.random(B), slot=1)
            sh.gather(1)
            data, idx, _ = sh.slot_read(1, with_rows=True)
            assert idx.min() >= n - 150
        else:
            data, idx, _ = sh.slot_read(0, with_rows=True)
        assert np.array_equal(data[0], np.stack([full[i][0] for i in idx]))
        assert np.array_equal(data[3], np.stack([full[i][3] for i in idx]))
        sh.close()


def test_frame_column_errors(mods):
    ReplayShard, _ = mods
    with pytest.raises(ValueError):
        ReplayShard(64, [(np.float32, (K,) + FRAME)], max_batch=8, frames={"shape": FRAME, "capacity": 64, "columns": [0]})
    sh = ReplayShard(64, [(np.uint8, (K,) + FRAME), (np.float32, ())], max_batch=8,
                     frames={"shape": FRAME, "capacity": 16, "columns": [0]})
    with pytest.raises(Exception):
        sh.append_frames(np.zeros((17,) + FRAME, np.uint8))          # more than the ring holds
    first = sh.append_frames(np.arange(3 * 84 * 84, dtype=np.uint8).reshape((3,) + FRAME))
    assert first == 0 and np.array_equal(sh.frame_read(1, 1)[0].ravel()[:4], (np.arange(4) + 84 * 84).astype(np.uint8))
    # a reference outside the frame ring is reported, not followed
    sh.append([np.full((8, K), 99, np.int32), np.zeros(8, np.float32)], np.ones(8, np.float32))
    sh.sample(8, uniforms=np.random.default_rng(0).random(8))
    sh.gather(0)
    with pytest.raises(Exception):
        sh.slot_read(0, with_rows=True)
    sh.close()


@pytest.mark.parametrize("layout", ["inproc", "process"])
def test_service_append_steps(layout):
    """The frame-de-duplicated layout behind the service surface: actors send one new frame per step
    (Client.append_steps), loaders hand out the same k-stacks FrameStack would have produced."""
    import __graft_entry__ as entry
    entry.build()
    from reth_b200 import service, service_mp
    mod = service if layout == "inproc" else service_mp
    B, k = 32, K
    frames_cfg = {"k": k, "shape": FRAME, "capacity": 2048, "columns": (0, 3)}
    proc, addr = mod.start_per(1024, B, alpha=0.6, beta=0.4, sample_start=B, frames=frames_cfg)
    try:
        client = mod.Client(addr)
        rng = np.random.default_rng(23)
        full = []                                           # transitions in arrival order, full stacks
        stacks = {}
        for env in range(2):
            f0 = rng.integers(0, 256, FRAME, dtype=np.uint8)
            stacks[env] = deque([f0] * k, maxlen=k)
            client.append_steps(env, np.zeros((0,) + FRAME, np.uint8), [np.zeros(0, np.int64), np.zeros(0, np.float32),
                                np.zeros(0, np.float32)], np.zeros(0, np.float32), reset=f0)
        for rnd in range(6):
            for env in range(2):
                n = 40
                fr = rng.integers(0, 256, (n,) + FRAME, dtype=np.uint8)
                a = rng.integers(0, 18, n)
                r = rng.standard_normal(n).astype(np.float32)
                d = np.zeros(n, np.float32)
                for i in range(n):
                    s0 = np.stack(stacks[env])
                    stacks[env].append(fr[i])
                    full.append((s0, int(a[i]), r[i], np.stack(stacks[env])))
                client.append_steps(env, fr, [a, r, d], rng.random(n).astype(np.float32) + 0.1)
        loader = mod.TorchCudaLoader(addr, buffer_size=2)
        for _ in range(6):
            data, indices, weights = loader.sample()
            assert tuple(data[0].shape) == (B, k) + FRAME
            s0, s1 = data[0].cpu().numpy(), data[3].cpu().numpy()
            for j, i in enumerate(indices):
                assert np.array_equal(s0[j], full[i][0]) and np.array_equal(s1[j], full[i][3])
            assert np.array_equal(data[1].cpu().numpy(), [full[i][1] for i in indices])
            client.update_priorities(indices, np.ones(B, np.float32))
        loader.close()
    finally:
        proc.terminate()
        proc.join()
