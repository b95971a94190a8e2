"""reth_replay_tree_read_noalloc (no counterpart in the reference): the
tree readback with its device scratch in a static device array instead of cudaMalloc.  Called through the C ABI
and checked against the existing entry point reth_replay_tree_read bit for bit (integer/heap-order copies of the
same float64 values: the tolerance of the other parity tests for tree arrays is exact equality), on two calls in a
row -- with different tree contents, and with a smaller shard second so that stale scratch contents lie beyond it.
Collected last (file name) so that a failure here hides no other test under `pytest -x`."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _read(sh, fn_name):
    lib = sh._lib
    out = [np.full(sh.capacity, -7.0, np.float64) for _ in range(3)]
    rc = getattr(lib, fn_name)(sh._h, out[0].ctypes.data, out[1].ctypes.data, out[2].ctypes.data, C.c_void_p(0))
    sh._ck(rc)
    return out


def test_tree_read_noalloc_matches_tree_read_two_calls_in_a_row():
    import __graft_entry__ as entry
    entry.build()
    from reth_b200 import ReplayError, ReplayShard
    rng = np.random.default_rng(123)
    big = ReplayShard(50000, [], max_batch=64, store_rows=False)
    small = ReplayShard(3001, [], max_batch=64, store_rows=False)
    try:
        big.append([], (rng.random(50000) + 0.01).astype(np.float32), alpha=0.6)
        small.append([], rng.random(2000) + 0.01, alpha=0.5)
        # call 1: the large tree
        want = _read(big, "reth_replay_tree_read")
        got = _read(big, "reth_replay_tree_read_noalloc")
        for g, w in zip(got, want):
            assert np.array_equal(g, w)
        assert got[0][0] > 0 and not np.any(got[2] == -7.0)
        # call 2, straight after: other contents (a write-back in between) ...
        idx = rng.integers(0, 50000, 4000)
        big.update(idx, rng.random(4000) + 0.5, alpha=0.6)
        want = _read(big, "reth_replay_tree_read")
        got = _read(big, "reth_replay_tree_read_noalloc")
        for g, w in zip(got, want):
            assert np.array_equal(g, w)
        # ... and a smaller shard, whose read must not see what the larger one left in the shared scratch
        want = _read(small, "reth_replay_tree_read")
        got = _read(small, "reth_replay_tree_read_noalloc")
        for g, w in zip(got, want):
            assert np.array_equal(g, w)
        assert not got[2][2000:].any()
        # only some of the arrays requested
        v = np.empty(small.capacity)
        small._ck(small._lib.reth_replay_tree_read_noalloc(small._h, None, None, v.ctypes.data, C.c_void_p(0)))
        assert np.array_equal(v, want[2])
        # beyond the static scratch: rejected with a clear error, nothing launched
        huge = ReplayShard(65537, [], max_batch=8, store_rows=False)
        try:
            with pytest.raises(ReplayError) as ei:
                _read(huge, "reth_replay_tree_read_noalloc")
            assert ei.value.code == -1 and "exceeds the static scratch" in str(ei.value)
            _read(huge, "reth_replay_tree_read")                   # the allocating entry point still serves it
        finally:
            huge.close()
    finally:
        big.close()
        small.close()
