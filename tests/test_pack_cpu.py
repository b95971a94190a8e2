"""Wire codec (reth_b200/pack.py) against bytes produced by the reference's own
utils/pack.py + client/client.py (tests/golden/pack_wire.npz, made by make_golden.py with a
fixed clock): decode equality, BYTE-exact encoding, and the columnar fast path."""
import numpy as np
import pytest

from reth_b200 import pack


def test_decode_reference_append_message(golden):
    g = golden("pack_wire")
    cols, w = pack.decode_append(g["msg"].tobytes())
    assert len(cols) == 3
    for got, want in zip(cols, (g["c0"], g["c1"], g["c2"])):
        assert got.dtype == want.dtype and np.array_equal(got, want) and got.flags.c_contiguous
    assert w.dtype == np.float32 and np.array_equal(w, g["weights"])
    idx, w2, step = pack.decode(g["upd"].tobytes())
    assert np.array_equal(idx, [5, 1, 5]) and np.array_equal(w2, g["weights"]) and step is True
    row0 = pack.decode(g["row0"].tobytes())
    assert np.array_equal(row0[0], g["c0"][0]) and row0[1] == g["c1"][0]


def test_encode_is_byte_exact(golden):
    g = golden("pack_wire")
    cols = [g["c0"], g["c1"], g["c2"]]
    assert bytes(pack.encode([c[0, ...] for c in cols], now=1234.5)) == g["row0"].tobytes()
    assert bytes(pack.encode_append_rows(cols, g["weights"], now=1234.5)) == g["msg"].tobytes()
    assert bytes(pack.encode([np.array([5, 1, 5], np.int64), g["weights"], True], now=1234.5)) == g["upd"].tobytes()


@pytest.mark.parametrize("n", [1, 7, 64])
def test_roundtrip_formats(n):
    rng = np.random.default_rng(n)
    cols = [rng.integers(0, 256, (n, 4, 84, 84), dtype=np.uint8), rng.integers(0, 18, n),
            rng.standard_normal(n).astype(np.float32), rng.random((n, 3)), (rng.random(n) < 0.5).astype(np.float32)]
    w = rng.random(n).astype(np.float32)
    for msg in (pack.encode_append_rows(cols, w), pack.encode_append_columns(cols, w)):
        got, gw = pack.decode_append(msg)
        assert len(got) == len(cols)
        for a, b in zip(got, cols):
            assert a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a, b)
        assert np.array_equal(gw, w)
    assert pack.is_columnar(pack.encode_append_columns(cols, w)) and not pack.is_columnar(pack.encode_append_rows(cols, w))


def test_general_path_and_errors():
    # rows of different struct sizes force the per-row path
    rows = [pack.encode([np.arange(3), np.float32(1.0)]), pack.encode([np.arange(3), np.float32(2.0)])]
    msg = pack.encode([rows, np.ones(2)])
    cols, w = pack.decode_append(msg)
    assert np.array_equal(cols[0], [[0, 1, 2], [0, 1, 2]]) and np.array_equal(cols[1], [1.0, 2.0])
    nested = {"a": [np.eye(2), b"xyz"], "b": 3}
    back = pack.decode(pack.encode(nested))
    assert np.array_equal(back["a"][0], np.eye(2)) and bytes(back["a"][1]) == b"xyz" and back["b"] == 3
    f = np.asfortranarray(np.arange(6).reshape(2, 3))
    assert np.array_equal(pack.decode(pack.encode([f]))[0], f)
    bad = bytearray(pack.encode([np.arange(4)]))
    with pytest.raises(ValueError):
        pack.decode(bad[:-1])


def test_row_scatter_in_the_library_and_staging_alloc(golden):
    """The ingest threads' path: reth_pack_scatter_rows (host code of the C-ABI library) transposes the (n, L) row
    matrix of a reference-format append into columns written straight into a caller-provided buffer."""
    import ctypes as C
    from reth_b200 import _cabi
    g = golden("pack_wire")
    slabs = []

    def alloc(nbytes):
        slabs.append(np.full(nbytes + 64, 0xEE, np.uint8))      # guard bytes behind the slab
        return slabs[-1][:nbytes]
    kind, cols, w, src = pack.open_append(g["msg"].tobytes(), alloc)
    assert kind == "columns" and src is None and len(slabs) == 1
    for got, want in zip(cols, (g["c0"], g["c1"], g["c2"])):
        assert got.dtype == want.dtype and np.array_equal(got, want) and got.flags.c_contiguous
        assert np.shares_memory(got, slabs[0])                   # the columns live in the staging slab
        assert got.ctypes.data % 256 == slabs[0].ctypes.data % 256   # each column 256-byte aligned inside it
    assert np.all(slabs[0][-64:] == 0xEE)
    assert np.array_equal(w, g["weights"])
    # lz4-framed and columnar messages go through the same allocator
    rng = np.random.default_rng(4)
    big = [rng.integers(0, 256, (16, 4, 84, 84), dtype=np.uint8) // 32, rng.integers(0, 18, 16), rng.random(16).astype(np.float32)]
    for msg in (pack.encode_append_rows(big, np.ones(16, np.float32), compress=True), pack.encode_append_columns(big, np.ones(16, np.float32))):
        n0 = len(slabs)
        _, cols, _, _ = pack.open_append(msg, alloc)
        assert len(slabs) == n0 + 1
        for a, b in zip(cols, big):
            assert np.array_equal(a, b) and np.shares_memory(a, slabs[-1])
    # the entry point itself: argument checks (a column outside its row, null pointers) return -1, nothing is written
    L = _cabi.lib()
    mat = np.arange(40, dtype=np.uint8).reshape(4, 10)
    out = np.zeros(12, np.uint8)
    dst = (C.c_void_p * 1)(out.ctypes.data)
    lo, ln = np.array([2], np.int64), np.array([3], np.int64)
    assert L.reth_pack_scatter_rows(mat.ctypes.data, 4, 10, 1, lo.ctypes.data, ln.ctypes.data, dst) == 12
    assert np.array_equal(out.reshape(4, 3), mat[:, 2:5])
    bad_len = np.array([9], np.int64)
    assert L.reth_pack_scatter_rows(mat.ctypes.data, 4, 10, 1, lo.ctypes.data, bad_len.ctypes.data, dst) == -1
    assert L.reth_pack_scatter_rows(None, 4, 10, 1, lo.ctypes.data, ln.ctypes.data, dst) == -1
    # a truncated message is refused
    with pytest.raises(ValueError):
        pack.open_append(g["msg"].tobytes()[:-5], alloc)
