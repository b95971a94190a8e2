"""Host-side logic that needs no GPU: the loader geometry derived from the reference's staleness budget, the numpy
checkers of bench.py's Mode G parity leg pinned against the oracle (they are what `mode_g.parity_ok` rests on), and the
recipe that places the reference's own modules under oracle/_ref."""
import os
import sys

import numpy as np
import pytest

from oracle import per_oracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_loader_geometry_follows_the_reference_budget():
    from reth_b200.service import loader_geometry
    # TorchCudaLoader(addr, buffer_size=4, num_procs=2) (test/apex-dqn/trainer.py:22-24): 4 + 2 + 2 = 8 batches in flight
    assert loader_geometry(4, 2, 16) == (4, 1)
    # the reference's defaults (8, 6): 16 batches in flight (client/torch_cuda_loader.py:70)
    assert loader_geometry(8, 6, 32) == (8, 1)
    assert loader_geometry(8, 6, 16) == (8, 1)
    assert loader_geometry(8, 6, 8) == (4, 1)            # a batch ring of 8 slots bounds the budget
    assert loader_geometry(1, 0, 16) == (1, 1)
    for b, p, r in ((4, 2, 16), (8, 6, 32), (2, 1, 4), (16, 8, 64)):
        g, a = loader_geometry(b, p, r)
        assert g * (a + 1) <= min(b + p + 2, r)          # never more batches between draw and consumption than the reference


@pytest.mark.parametrize("cap", [1, 2, 7, 1000, 4096, 65537])
def test_mode_g_numpy_checkers_match_the_oracle(cap):
    sys.path.insert(0, ROOT)
    import bench_modeg
    rng = np.random.default_rng(cap)
    tree = po.OracleSumTree(cap)
    n = max(1, (3 * cap) // 4)
    idx = rng.permutation(cap)[:n]
    tree.update(idx, rng.random(n) * (1 + 50 * (rng.random(n) < 0.03)) + 1e-3)
    tree.update(idx[: n // 3], rng.random(n // 3) + 1e-3)          # rewrite some leaves
    s, m = bench_modeg.np_tree_recompute(tree._val.copy())
    assert np.array_equal(s, tree._sum) and np.array_equal(m, tree._min)   # incl. never-maintained nodes (sum = min = 0)
    total = tree.sum()
    targets = (np.arange(257) + rng.random(257)) * (total / 257)
    want = np.array([tree.find_index(t) for t in targets])
    assert np.array_equal(bench_modeg.np_find_index(tree._sum, tree._val, targets), want)


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference checkout is only present in the build container")
def test_make_ref_places_the_reference_modules_unchanged():
    from oracle import make_ref
    path = make_ref.build()
    assert path and os.path.isdir(os.path.join(path, "reth_ref"))
    for src, dst in make_ref.FILES.items():
        a = open(os.path.join("/root/reference", src), "rb").read()
        b = open(os.path.join(path, "reth_ref", dst), "rb").read()
        assert a == b, dst                                           # byte for byte: the reference itself is what gets timed
    # oracle/_ref never enters the history
    assert "oracle/_ref/" in open(os.path.join(ROOT, ".gitignore")).read()


def test_shared_memory_ingest_ring_is_mapped_without_the_resource_tracker():
    """The server side of the same-host ingest ring (service_mp._ShmSegment) maps the client's segment directly: the
    producer's bump allocator, the consumed counter the server writes back, and a segment that is gone."""
    from reth_b200 import service_mp as m
    ring = m._ShmRing(1 << 16)
    a = ring.reserve(1000)
    b = ring.reserve(3000)
    assert a == 64 and b == 64 + 1024 and ring.seq == 2                   # 64-byte header, 64-byte granules
    ring.buf[a:a + 1000] = 7
    name = ring.shm.name
    seg = m._ShmSegment(name, ring.size, 0)
    assert not seg.pinned or seg.pinned is True
    assert len(seg.buf) >= ring.size and int(seg.buf[a:a + 1000].sum()) == 7000
    seg.consumed[0] = 1                                                    # the server has copied message 1 out
    ring._retire()
    assert [s for s, _, _ in ring.inflight] == [2]
    assert ring.reserve(1 << 17) is None                                   # larger than the ring: the caller falls back
    with pytest.raises(ValueError):
        m._ShmSegment(name, ring.size * 2, 0)                              # the descriptor disagrees with the mapping
    seg.consumed[0] = 2
    seg.close()
    ring.close(timeout=1.0)
    with pytest.raises(FileNotFoundError):
        m._ShmSegment(name, 64, 0)
