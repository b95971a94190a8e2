"""lz4 frame codec of the ingest wire format (reth_lz4f_*, reth_b200/csrc/lz4f.h): host code, so these
run without a GPU.  Pins: frames written by liblz4 (tests/golden/lz4_frames.npz), the published
XXH32 known answers, the frame header bytes of the lz4 CLI, and -- when pyarrow is importable --
liblz4 decoding what this build's compressor writes."""
import os
import struct

import numpy as np
import pytest

from reth_b200 import _cabi, pack

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "lz4_frames.npz"))
NAMES = sorted(k[6:] for k in GOLD.files if k.startswith("plain_"))


def raw_decompress(buf, cap):
    L = _cabi.lib()
    src = np.frombuffer(bytes(buf), np.uint8)
    out = np.empty(max(cap, 1), np.uint8)
    n = L.reth_lz4f_decompress(src.ctypes.data, src.size, out.ctypes.data, cap)
    return n, out[:max(n, 0)].tobytes()


@pytest.mark.parametrize("name", NAMES)
def test_decodes_liblz4_frames(name):
    plain, frame = GOLD["plain_" + name], GOLD["frame_" + name]
    assert np.array_equal(pack.lz4_decompress(frame), plain)
    n, out = raw_decompress(frame, plain.size)               # exact-size output buffer
    assert n == plain.size and out == plain.tobytes()
    if plain.size:
        assert raw_decompress(frame, plain.size - 1)[0] == -4


@pytest.mark.parametrize("name", NAMES)
def test_round_trip_own_frames(name):
    plain = GOLD["plain_" + name]
    frame = pack.lz4_compress(plain)
    L = _cabi.lib()
    assert L.reth_lz4f_content_size(frame.ctypes.data, frame.size) == plain.size
    assert frame.size <= L.reth_lz4f_compress_bound(plain.size)
    assert np.array_equal(pack.lz4_decompress(frame), plain)
    if name in ("frame", "period5"):
        assert frame.size < plain.size // 2                  # it does compress


@pytest.mark.parametrize("name", NAMES)
def test_liblz4_decodes_own_frames(name):
    pa = pytest.importorskip("pyarrow")
    plain = GOLD["plain_" + name]
    if plain.size == 0:
        pytest.skip("pyarrow rejects a zero-size output buffer")
    frame = pack.lz4_compress(plain)
    back = pa.Codec("lz4").decompress(frame.tobytes(), decompressed_size=plain.size)
    assert np.array_equal(np.frombuffer(back, np.uint8), plain)


def test_known_answers():
    # lz4 CLI header: FLG 0x64 (independent blocks, content checksum), BD 0x40 -> HC 0xA7;
    # BD 0x70 -> HC 0xB9.  XXH32("") = 0x02CC5D05; XXH32("Nobody inspects the spammish repetition") = 0xE2293B2F.
    empty = bytes.fromhex("04224d186440a7") + bytes(4) + struct.pack("<I", 0x02CC5D05)
    assert raw_decompress(empty, 0)[0] == 0
    assert raw_decompress(bytes.fromhex("04224d186470b9") + bytes(4) + struct.pack("<I", 0x02CC5D05), 0)[0] == 0
    assert raw_decompress(bytes.fromhex("04224d186440a8") + bytes(8), 0)[0] == -3          # header checksum
    msg = b"Nobody inspects the spammish repetition"
    f = bytes.fromhex("04224d186440a7") + struct.pack("<I", len(msg) | 0x80000000) + msg + bytes(4) + struct.pack("<I", 0xE2293B2F)
    assert raw_decompress(f, 64) == (len(msg), msg)
    assert raw_decompress(f[:-1] + b"\x00", 64)[0] == -3                                    # content checksum
    # block checksum (FLG 0x74 = 0x64 | B.Checksum): computed by the frame we build, verified on decode
    L = _cabi.lib()
    assert raw_decompress(f[:25], 64)[0] == -1                                              # truncated
    assert raw_decompress(f, 10)[0] == -4
    assert raw_decompress(bytes(16), 64)[0] == -2                                           # bad magic
    skip = struct.pack("<II", 0x184D2A53, 5) + b"abcde"
    assert raw_decompress(skip + f + f, 128) == (2 * len(msg), msg * 2)                     # skippable + concatenated
    assert L.reth_lz4f_content_size(np.frombuffer(f, np.uint8).ctypes.data, len(f)) == -1


def test_hand_built_sequences():
    """Block format corner cases, built by hand: overlapping matches (offset < length), a long
    literal run and a long match with 255-continued length bytes."""
    def frame(block):
        return bytes.fromhex("04224d18604082") + struct.pack("<I", len(block)) + block + bytes(4)
    # "ab" then match offset 2 length 10 -> "abababababab", then 5 literals
    blk = bytes([0x26]) + b"ab" + struct.pack("<H", 2) + bytes([0x50]) + b"vwxyz"
    assert raw_decompress(frame(blk), 64) == (17, b"ab" * 6 + b"vwxyz")
    # offset 1 run of 300 'q': literal 'q', match len 300 = 4 + 15 + 255 + 26
    blk = bytes([0x1F]) + b"q" + struct.pack("<H", 1) + bytes([255, 26]) + bytes([0x50]) + b"12345"
    assert raw_decompress(frame(blk), 400) == (306, b"q" * 301 + b"12345")
    # 300 literals (15 + 255 + 30) then offset 3 (pattern of period 3) length 20
    lit = bytes(range(256)) + bytes(range(44))
    blk = bytes([0xFF]) + bytes([255, 30]) + lit + struct.pack("<H", 3) + bytes([1]) + bytes([0x50]) + b"12345"
    n, out = raw_decompress(frame(blk), 400)
    exp = bytearray(lit)
    for _ in range(20):
        exp.append(exp[-3])
    assert (n, out) == (325, bytes(exp) + b"12345")
    # offset beyond the start of the data is rejected
    assert raw_decompress(frame(bytes([0x10]) + b"a" + struct.pack("<H", 2) + bytes([0x50]) + b"12345"), 64)[0] == -2


def test_pack_compress_round_trip():
    rng = np.random.default_rng(0)
    obj = [np.repeat(rng.integers(0, 4, (8, 441), dtype=np.uint8), 16, axis=1), {"k": np.arange(5.0), "b": b"xyz"}, 3]
    buf = pack.encode(obj, compress=True)
    assert len(buf) < obj[0].nbytes // 2
    back = pack.decode(buf)
    assert np.array_equal(back[0], obj[0]) and np.array_equal(back[1]["k"], obj[1]["k"]) and bytes(back[1]["b"]) == b"xyz"
    assert back[2] == 3


@pytest.mark.parametrize("wire", ["rows", "columns"])
def test_compressed_append_message(wire):
    """worker.py:60 sends compress=True appends; the server side decodes them to the same columns."""
    rng = np.random.default_rng(1)
    n = 16
    cols = [np.repeat(rng.integers(0, 3, (n, 4, 21, 84), dtype=np.uint8), 4, axis=2), rng.integers(0, 18, n),
            rng.standard_normal(n).astype(np.float32)]
    w = rng.random(n).astype(np.float32)
    enc = pack.encode_append_rows if wire == "rows" else pack.encode_append_columns
    plain, comp = enc(cols, w), enc(cols, w, compress=True)
    assert len(comp) < len(plain) // 2
    got, gw = pack.decode_append(comp)
    assert np.array_equal(gw, w)
    for a, b in zip(got, cols):
        assert a.dtype == b.dtype and np.array_equal(a, b)


def test_sanitizer_fuzz(tmp_path):
    """The codec under AddressSanitizer + UBSan: round trips and mutated frames (tests/native/fuzz_lz4f.cpp)."""
    import shutil
    import subprocess
    if shutil.which("g++") is None:
        pytest.skip("g++ not available")
    src = os.path.join(os.path.dirname(__file__), "native", "fuzz_lz4f.cpp")
    exe = str(tmp_path / "fuzz_lz4f")
    build = subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", src, "-o", exe],
                           capture_output=True, text=True)
    if build.returncode != 0 and "sanitize" in build.stderr:
        pytest.skip("sanitizer runtime not available")
    assert build.returncode == 0, build.stderr
    run = subprocess.run([exe, "600"], capture_output=True, text=True, timeout=300)
    assert run.returncode == 0 and run.stdout.startswith("ok"), run.stdout + run.stderr
