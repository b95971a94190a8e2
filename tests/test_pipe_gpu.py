"""GPU tests of the asynchronous loader pipeline (reth_replay_pipe_*) and of the service layer
on top of it: ordering (updates and draws are totally ordered by call order), last-writer-wins
over a sequence of fire-and-forget write-backs, the borrow rule of lent batch slots, and
row/weight consistency of pre-sampled groups.  Reference behaviour restated:
  client/torch_cuda_loader.py:152-163 (lend until the next sample),
  server/sampler_loop.py:29-39 (drain updates, then emit a pre-sampled batch),
  sampler/per_sampler.py:24-28 (IS weights from the tree state at draw time)."""
import numpy as np
import pytest

from oracle import per_oracle as po

pytestmark = pytest.mark.gpu

ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]


@pytest.fixture(scope="module")
def mods():
    import __graft_entry__ as entry
    entry.build()
    from reth_b200 import ReplayShard, service
    return ReplayShard, service


def _fill(rng, n):
    s0 = rng.integers(0, 256, (n, 4, 84, 84), dtype=np.uint8)
    return [s0, rng.integers(0, 18, n), rng.standard_normal(n).astype(np.float32), s0[::-1].copy(),
            np.zeros(n, np.float32)]


@pytest.mark.parametrize("G,ahead", [(1, 0), (1, 2), (4, 1), (2, 3)])
def test_pipe_shard_level(mods, G, ahead):
    """C-ABI level: groups delivered in issue order, host index rows == device slots, weights match the
    tree state the draw saw, async write-backs land in order (final tree == oracle replay)."""
    ReplayShard, _ = mods
    cap, B, n = 4000, 32, 3000
    rng = np.random.default_rng(5)
    sh = ReplayShard(cap, ATARI, max_batch=B, ring_depth=(ahead + 1) * G + G)
    cols = _fill(rng, n)
    stream = sh.pipe_stream
    sh.append(cols, rng.random(n).astype(np.float32) + 0.05, alpha=0.6, stream=stream)
    pipe = sh.open_pipe(B, G, seed=7)
    assert pipe.n_groups == ahead + 2
    log = []
    for _ in range(ahead + 1):
        pipe.issue(0.4)
    for it in range(12):
        g = pipe.wait()
        assert g == it % pipe.n_groups
        idx = np.array(pipe.idx_host[g])                      # (G, B)
        w = np.array(pipe.w_host[g])
        for k in range(G):
            data, didx, dw, _ = sh.slot_tensors(g * G + k)
            assert np.array_equal(didx.cpu().numpy(), idx[k])
            assert np.array_equal(dw.cpu().numpy(), w[k])
            assert np.array_equal(data[0].cpu().numpy(), cols[0][idx[k]])
            assert np.array_equal(data[1].cpu().numpy(), cols[1][idx[k]])
            assert np.array_equal(data[3].cpu().numpy(), cols[3][idx[k]])
        assert w.max() <= 1.0 + 1e-12 and w.min() > 0
        # write-backs: one staging entry per group, committed before the next draw is issued
        e = pipe.update_begin()
        newp = np.abs(rng.standard_normal((G, B))).astype(np.float32)
        pipe.upd_idx[e, :G * B] = idx.reshape(-1)
        pipe.upd_p32[e, :G * B] = newp.reshape(-1)
        pipe.update_commit(e, G * B, True, 0.6)
        log.append((idx.reshape(-1).copy(), newp.reshape(-1).copy()))
        pipe.issue(0.4)
    pipe.drain()
    s, m, v = sh.tree_arrays(stream=stream)
    # last writer wins over the whole sequence, transform within 1e-5 of numpy's float32 pow
    expect = {}
    for idx, p in log:
        for i, x in zip(idx, p):
            expect[int(i)] = x
    keys = np.fromiter(expect.keys(), np.int64)
    want = po.normalize_weights(np.fromiter(expect.values(), np.float32), 0.6)
    assert np.allclose(v[keys], want.astype(np.float64), rtol=1e-5, atol=0)
    ref = po.OracleSumTree(cap)
    ref.update(np.arange(n), v[:n])
    assert np.array_equal(s, ref._sum) and np.array_equal(m, ref._min)
    pipe.close()
    sh.close()


def test_pipe_strict_weights_match_tree(mods):
    """ahead = 0: a draw sees every earlier update; its weights are (val/min)^-beta of that tree."""
    ReplayShard, _ = mods
    cap, B = 1000, 64
    rng = np.random.default_rng(9)
    sh = ReplayShard(cap, [], max_batch=B, ring_depth=2, store_rows=False)
    stream = sh.pipe_stream
    sh.update(np.arange(cap), rng.random(cap) + 0.01, transform=False, stream=stream)
    pipe = sh.open_pipe(B, 1, seed=1)
    for it in range(6):
        pipe.issue(0.7)
        g = pipe.wait()
        idx, w = np.array(pipe.idx_host[g, 0]), np.array(pipe.w_host[g, 0])
        s, m, v = sh.tree_arrays(stream=stream)
        mn = m[0] if s[0] != 0 else 1.0
        assert np.allclose(w, (v[idx] / mn) ** -0.7, rtol=1e-5, atol=0)
        pipe.update_async(idx, rng.random(B) + 0.01, alpha=0.6)
    with pytest.raises(Exception):
        pipe.wait()                                            # nothing issued
    pipe.close()
    sh.close()


def test_pipe_errors(mods):
    ReplayShard, _ = mods
    from reth_b200 import ReplayError
    sh = ReplayShard(100, [], max_batch=8, ring_depth=4, store_rows=False)
    sh.update(np.arange(100), np.ones(100), transform=False, stream=sh.pipe_stream)
    with pytest.raises(ReplayError):
        sh.open_pipe(8, 4)                                     # fewer than 2 groups fit the ring
    pipe = sh.open_pipe(8, 2)
    with pytest.raises(ReplayError):
        sh.open_pipe(8, 1)                                     # already open
    pipe.issue(0.4)
    pipe.issue(0.4)
    with pytest.raises(ReplayError):
        pipe.issue(0.4)                                        # both group slots are issued
    pipe.wait()
    with pytest.raises(ReplayError):
        pipe.issue(0.4)                                        # one lent + one issued
    pipe.wait()                                                # releases the first group
    pipe.issue(0.4)
    e = pipe.update_begin()
    with pytest.raises(ReplayError):
        pipe.update_commit(e, 17, True, 0.6)                   # more than the entry holds
    with pytest.raises(ReplayError):
        pipe.update_commit((e + 1) % pipe.upd_entries, 8, True, 0.6)
    pipe.close()
    sh.close()


@pytest.mark.parametrize("prefetch,ahead", [(1, 0), (4, 1), (2, 2)])
def test_service_pipeline_final_tree(mods, prefetch, ahead):
    """Service level: TorchCudaLoader + Client.update_priorities, learner-style loop; the final tree
    equals an oracle replay of the write-back sequence whatever the prefetch depth."""
    _, service = mods
    cap, B, n = 5000, 32, 2000
    rng = np.random.default_rng(2)
    proc, addr = service.start_per(cap, B, alpha=0.6, beta=0.4, sample_start=B)
    try:
        client = service.Client(addr)
        cols = _fill(rng, n)
        for st in range(0, n, 500):
            client.append([c[st:st + 500] for c in cols], rng.random(500).astype(np.float32) + 0.05)
        loader = service.TorchCudaLoader(addr, buffer_size=(ahead + 1) * prefetch, prefetch=prefetch, ahead=ahead)
        eng = service._ENGINES[addr]
        log = []
        held = None
        for it in range(25):
            data, indices, weights = loader.sample()
            assert indices.dtype == np.int64 and indices.shape == (B,)
            assert np.array_equal(data[0].cpu().numpy(), cols[0][indices])
            if held is not None and prefetch > 1 and it % prefetch:
                # the previous batch of the same group is still intact (lent until the group is released)
                assert np.array_equal(held[0].cpu().numpy(), cols[0][held[1]])
            held = (data[0], indices)
            loss = np.abs(rng.standard_normal(B)).astype(np.float32)
            client.update_priorities(indices, loss)
            log.append((indices.copy(), loss))
        eng.drain()
        s, m, v = eng.shard.tree_arrays(stream=eng.stream)
        expect = {}
        for idx, p in log:
            for i, x in zip(idx, p):
                expect[int(i)] = x
        keys = np.fromiter(expect.keys(), np.int64)
        want = po.normalize_weights(np.fromiter(expect.values(), np.float32), 0.6)
        assert np.allclose(v[keys], want.astype(np.float64), rtol=1e-5, atol=0)
        ref = po.OracleSumTree(cap)
        ref.update(np.arange(n), v[:n])
        assert np.array_equal(s, ref._sum) and np.array_equal(m, ref._min)
        loader.close()
        assert eng.pipe is None
        # a second loader with another geometry can attach after the first one closed
        l2 = service.NumpyLoader(addr)
        d2, i2, w2 = l2.sample()
        assert np.array_equal(d2[0], cols[0][i2]) and w2.dtype == np.float64
        l2.close()
    finally:
        proc.terminate()


def test_appends_never_tear_rows_of_groups_in_flight(mods):
    """gathers alternate between two streams and several groups are in flight; an append that recycles
    sampled slots must wait for ALL of them (reth_replay_append waits for the newest gather of each gather stream), so
    a gathered row is entirely old or entirely new -- across its bytes and across its columns -- never half and half
    (the reference's LMDB rows are replaced atomically, server/main_loop.py:55-57).  Every row carries its version
    in every byte of the frame column and in the scalar column."""
    ReplayShard, _ = mods
    import torch
    cap, B, G, ahead = 2048, 256, 2, 2
    struct = [(np.uint8, (4, 84, 84)), (np.int64, ())]
    sh = ReplayShard(cap, struct, max_batch=B, ring_depth=(ahead + 2) * G)
    stream = sh.pipe_stream

    def rows(n, first_slot, version):
        frames = torch.full((n, 4, 84, 84), version % 251, dtype=torch.uint8, device="cuda")
        tag = (torch.arange(first_slot, first_slot + n, device="cuda") % cap) + version * cap
        torch.cuda.current_stream().synchronize()      # the append runs on the shard's own stream, not on torch's
        return [frames, tag]
    sh.append(rows(cap, 0, 0), torch.ones(cap, device="cuda"), alpha=1.0, transform=False, stream=stream)
    pipe = sh.open_pipe(B, G, seed=3)
    for _ in range(ahead + 1):
        pipe.issue(0.4)
    version, tail = 0, 0
    for it in range(40):
        g = pipe.wait()
        for k in range(G):
            data, idx, _, _ = sh.slot_tensors(g * G + k)
            frames, tag = data[0].reshape(B, -1), data[1]
            v = tag // cap
            assert torch.equal(tag % cap, idx), it                                 # the scalar column is the sampled slot's
            assert bool((frames == (v % 251).to(torch.uint8)[:, None]).all()), it   # and the frames are the same version's
        # recycle 768 slots (more than a third of the ring: in-flight groups have sampled many of them)
        version += 1
        n = 768
        sh.append(rows(n, tail, version), torch.ones(n, device="cuda"), alpha=1.0, transform=False, stream=stream)
        tail = (tail + n) % cap
        pipe.issue(0.4)
    pipe.drain()
    pipe.close()
    sh.close()
