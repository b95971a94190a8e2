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
