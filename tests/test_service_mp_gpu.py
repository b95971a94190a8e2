"""The drop-in layout of the reference on the GPU: server in its own process, clients and
loaders built from the address string in other processes (test/apex-dqn/trainer.py:49-66,
reth_buffer/test/test_rb.py:12-35,63-85), ZeroMQ + the reference wire format for append/update,
CUDA IPC for the batch ring."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def reth_buffer():
    import __graft_entry__ as entry
    entry.build()
    sys.path.insert(0, os.path.join(ROOT, "reth_b200", "compat"))
    os.environ.pop("RETH_B200_SERVER", None)
    import reth_buffer
    yield reth_buffer
    sys.path.remove(os.path.join(ROOT, "reth_b200", "compat"))


@pytest.mark.parametrize("compress", [False, True])
def test_rb_basic_server_process(reth_buffer, compress):
    """test_rb.py:12-35 against a spawned server; append in the reference's ROW wire format, plain
    and with the lz4-framed body the Ape-X workers send (worker.py:60 compress=True)."""
    from reth_b200 import service_mp
    CAPACITY, BATCH_SIZE = 100000, 32
    rb_proc, addr = reth_buffer.start_server(CAPACITY, BATCH_SIZE)
    try:
        assert addr.startswith("tcp://") and rb_proc.is_alive()
        client = service_mp.Client(addr, wire="rows")
        loader = reth_buffer.NumpyLoader(addr)
        data = [np.random.rand(1000, 4, 84), np.random.rand(1000), np.arange(1000)]
        buf_weights = np.random.rand(1000) + 1
        for start in range(0, 1000, 100):
            batch = [x[start:start + 100] for x in data]
            client.append(batch, buf_weights[start:start + 100], compress=compress)
        for _ in range(10):
            batch, indices, weights = loader.sample()
            assert indices.dtype == np.int64 and weights.dtype == np.float64
            for i, idx in enumerate(indices):
                assert batch[2][i] == idx
                assert batch[1][i] == data[1][idx]
                assert np.array_equal(batch[0][i], data[0][idx])
        buf_weights = np.random.rand(1000) + 10
        client.update_priorities(np.arange(1000), buf_weights)
        assert loader.drain()["cnt"] == 2000                  # sampler_loop.py:36: appended + updated indices
        loader.close()
    finally:
        rb_proc.terminate()
        rb_proc.join()
    assert not rb_proc.is_alive()


@pytest.mark.parametrize("wire", ["shm", "columns"])
def test_torch_cuda_loader_ipc(reth_buffer, wire):
    """trainer.py:19-36 shape: TorchCudaLoader in a process other than the server's; the lent CUDA
    tensors are views of the server's batch ring (CUDA IPC) and hold the rows of `indices`."""
    import torch
    B = 64
    proc, addr = reth_buffer.start_per(5000, B, alpha=0.5, beta="0.4,1,2000", sample_start=200)
    try:
        from reth_b200 import service_mp
        client = service_mp.Client(addr, wire=wire, shm_mb=8) # 8 MB ring: 64 Atari rows x 2 frames wrap it often
        assert reth_buffer.Client(addr).wire == "shm"         # same host: the default picks the shared-memory ring
        rng = np.random.default_rng(0)
        s0 = rng.integers(0, 256, (600, 4, 84, 84), dtype=np.uint8)
        cols = [s0, rng.integers(0, 18, 600), rng.standard_normal(600).astype(np.float32), s0[::-1].copy(),
                np.zeros(600, np.float32)]
        for st in range(0, 600, 64):
            client.append([c[st:st + 64] for c in cols], np.ones(len(cols[0][st:st + 64]), np.float32))
        loader = reth_buffer.TorchCudaLoader(addr, buffer_size=4, num_procs=2)     # trainer.py:23
        for it in range(12):
            data, indices, weights = loader.sample()
            assert isinstance(indices, np.ndarray) and indices.dtype == np.int64 and indices.shape == (B,)
            assert weights.is_cuda and weights.dtype == torch.float64 and tuple(weights.shape) == (B,)
            assert data[0].is_cuda and data[0].dtype == torch.uint8 and tuple(data[0].shape) == (B, 4, 84, 84)
            assert np.array_equal(data[0].cpu().numpy(), s0[indices])
            assert np.array_equal(data[1].cpu().numpy(), cols[1][indices])
            assert np.array_equal(data[3].cpu().numpy(), cols[3][indices])
            assert float(weights.max()) <= 1.0 + 1e-12
            loss = np.abs(rng.standard_normal(B)).astype(np.float32)
            client.update_priorities(np.asarray(indices), np.asarray(loss))         # trainer.py:36
        assert loader.group == 4                              # the whole buffer_size is one group across processes
        st = loader.drain()
        assert st["cnt"] == 600 + 12 * B
        with pytest.raises(AssertionError):
            client.append([c[:10] for c in cols], np.ones(9))
        loader.close()
    finally:
        proc.terminate()
        proc.join()


@pytest.mark.parametrize("wire", ["columns", "rows"])
def test_draw_sees_earlier_writes_of_the_same_process(reth_buffer, wire):
    """Appends, write-backs and loader requests travel on three sockets; a strict-alternation loader
    (prefetch 1, nothing ahead) must still see every write its own process made before sample():
    the request names the writes and the server takes them in first."""
    from reth_b200 import service_mp
    B, n = 32, 512
    proc, addr = reth_buffer.start_per(4096, B, alpha=1.0, beta=0.4, sample_start=B)
    try:
        client = service_mp.Client(addr, wire=wire, coalesce=8 if wire == "columns" else 1)
        loader = service_mp.NumpyLoader(addr)                 # prefetch 1 / ahead 0
        ids = np.arange(n, dtype=np.int64)
        client.append([ids, ids.astype(np.float32)], np.full(n, 1e-3, np.float32))
        rng = np.random.default_rng(0)
        for it in range(30):
            hot = int(rng.integers(0, n))
            client.update_priorities(np.array([hot]), np.array([1e9], np.float32))   # held back for coalescing
            data, indices, weights = loader.sample()
            if wire == "columns":                             # numbered frames: exact ordering
                assert np.all(indices == hot), (it, hot, indices[:8])
                assert np.array_equal(data[0], indices)
            client.update_priorities(np.array([hot]), np.array([1e-3], np.float32))
        assert loader.drain()["cnt"] == n + 60
        loader.close()
        client.close()
    finally:
        proc.terminate()
        proc.join()


def test_second_topic_across_processes(reth_buffer):
    """Two sampler topics on a spawned server: the loader of the second topic gets its batches in the
    topic's reserved slot of the shared batch ring (CUDA IPC), rows consistent with its indices."""
    from reth_buffer.sampler import UniformSampler
    B, n = 32, 512
    samplers = [{"sampler_cls": reth_buffer.PERSampler, "sample_start": B},
                {"sampler_cls": UniformSampler, "topic": "uni", "sample_start": B}]
    proc, addr = reth_buffer.start_server(2048, B, samplers=samplers)
    try:
        client = reth_buffer.Client(addr)
        ids = np.arange(n)
        feat = np.repeat(ids[:, None], 8, 1).astype(np.float32)
        client.append([ids, feat], np.ones(n, np.float32))
        per = reth_buffer.TorchCudaLoader(addr, buffer_size=2)
        uni = reth_buffer.TorchCudaLoader(addr, topic="uni")
        uni_np = reth_buffer.NumpyLoader(addr, topic="uni")
        for _ in range(5):
            for ld in (per, uni):
                data, indices, weights = ld.sample()
                assert np.array_equal(data[0].cpu().numpy(), indices)
                assert np.array_equal(data[1].cpu().numpy(), feat[indices])
            batch, indices, weights = uni_np.sample()
            assert np.array_equal(batch[0], indices) and np.all(weights == 1.0)
            client.update_priorities(indices, np.ones(B, np.float32))
        assert per.drain()["cnt"] == n + 5 * B
        for ld in (per, uni, uni_np):
            ld.close()
    finally:
        proc.terminate()
        proc.join()


def _producer(addr, n_batches, seed):
    sys.path.insert(0, ROOT)
    from reth_b200 import service
    client = service.Client(addr)
    rng = np.random.default_rng(seed)
    for _ in range(n_batches):
        ids = rng.integers(0, 1 << 30, 32)
        client.append([ids, np.repeat(ids[:, None], 16, 1).astype(np.float32)], np.ones(32))
    client.close()


def test_mp_producers(reth_buffer):
    """test_rb.py:63-85: several producer processes feed one server; rows stay self-consistent."""
    import multiprocessing as mp
    B = 32
    proc, addr = reth_buffer.start_per(20000, B, sample_start=B)
    try:
        ctx = mp.get_context("spawn")
        ps = [ctx.Process(target=_producer, args=(addr, 20, k)) for k in range(3)]
        [p.start() for p in ps]
        loader = reth_buffer.NumpyLoader(addr)
        for _ in range(20):
            data, indices, weights = loader.sample()
            assert len(data) == 2
            assert np.array_equal(data[1], np.repeat(data[0][:, None], 16, 1).astype(np.float32))
        [p.join(60) for p in ps]
        assert all(p.exitcode == 0 for p in ps)
        assert loader.drain()["cnt"] == 3 * 20 * 32
        loader.close()
    finally:
        proc.terminate()
        proc.join()


@pytest.mark.timeout(240)
def test_256_actor_connections_row_format(reth_buffer):
    """BASELINE.json config 4 as worded -- 256 actor append streams: 256 concurrent PUSH connections
    (worker.py:22,57-61) send 64-row messages in the reference's row wire format, half of them lz4-framed,
    through the server's ingest threads.  Every row must land exactly once and self-consistent, the messages of
    one connection in the order they were sent (slot = arrival order, fifo_policy.py:11-18); a malformed
    message in between is dropped without stopping the server."""
    import time
    import zmq
    from reth_b200 import pack, service_mp
    n_conn, per_conn, rows, B = 256, 3, 64, 64
    total = n_conn * per_conn * rows
    proc, addr = reth_buffer.start_per(total, B, sample_start=B)
    ctx = zmq.Context(2)
    socks = []
    try:
        meta = service_mp.get_meta(addr)
        for k in range(n_conn):
            s = ctx.socket(zmq.PUSH)
            s.connect(meta["append_addr"])
            socks.append(s)
        for m in range(per_conn):
            for k, s in enumerate(socks):
                ids = (k * per_conn + m) * rows + np.arange(rows)
                cols = [ids, np.repeat(ids[:, None], 24, 1).astype(np.float32), np.full(rows, m, np.int32)]
                s.send(pack.encode_append_rows(cols, np.ones(rows, np.float32), compress=bool(k & 1)), copy=False)
            if m == 1:
                socks[0].send(b"\x00\x00\x00\x02garbage")     # not a message: dropped, the server keeps going
        loader = reth_buffer.NumpyLoader(addr)
        t0 = time.time()
        while True:
            st = loader.drain()
            if st["cnt"] >= total or time.time() - t0 > 90:
                break
            time.sleep(0.01)
        assert st["cnt"] == total and st["dropped"]["append"] == 1
        ids = loader.ring_read(0, 0, total)
        feat = loader.ring_read(1, 0, total)
        mcol = loader.ring_read(2, 0, total)
        assert np.array_equal(np.sort(ids), np.arange(total))                     # every row exactly once
        assert np.array_equal(feat, np.repeat(ids[:, None], 24, 1).astype(np.float32))
        msg = ids // rows
        assert np.array_equal(ids.reshape(-1, rows), (msg.reshape(-1, rows)[:, :1] * rows + np.arange(rows)))  # messages stay whole
        assert np.array_equal(mcol, msg % per_conn)
        conn, m = msg[::rows] // per_conn, msg[::rows] % per_conn
        for k in range(n_conn):                                                    # per connection: arrival order = send order
            assert np.array_equal(m[conn == k], np.arange(per_conn)), k
        data, indices, weights = loader.sample()
        assert np.array_equal(data[0], ids[indices])
        loader.close()
    finally:
        for s in socks:
            s.close(0)
        ctx.term()
        proc.terminate()
        proc.join()
