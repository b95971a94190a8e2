"""Wire codec of the reth_buffer service (host side, no kernel).

The byte format is the reference's (reth_buffer/reth_buffer/utils/pack.py:60-98,148-164), kept so
that actors built against the reference can feed this server unchanged:

    [4-byte big-endian header length][msgpack header][raw body]
    header = {"meta": <tree>, "body_len": int, "time": float}
    <tree> mirrors the python object; an ndarray becomes
        {KEY_NUMPY: True, "header": {"shape","fortran_order","descr"}, "raw_size", "offset", "length"}
    and a bytes-like becomes {KEY_BYTES: True, "raw_size", "offset", "length"}; offset/length
    address the body.  An lz4-framed body is flagged by KEY_COMPRESS (utils/pack.py:67-84); the
    reference's actors use it (test/apex-dqn/worker.py:60).  The frame codec is the `lz4` package
    when it is installed, else the C implementation in this build's library
    (reth_lz4f_*, include/reth_replay.h; reth_b200/csrc/lz4f.h).

Client.append sends  encode([[encode(row_0), ..., encode(row_n-1)], weights])  with one inner
message per transition (client/client.py:26-35); `decode_append` turns that back into columns
with one strided copy per column instead of the reference's per-row Python loop
(client/numpy_loader.py:35-50).  This build's own clients send the columns directly
(`encode([columns, weights])` with COLUMNAR set in the header), which skips the row split.
"""
import threading
import time

import msgpack
import numpy as np

KEY_COMPRESS = "reth_compress_953531b51ab8"
KEY_NUMPY = "reth_numpy_356b759ef2b3"
KEY_BYTES = "reth_bytes_a32b00c47e72"
KEY_COLUMNAR = "reth_b200_columnar"
KEY_SHM = "reth_b200_shm"              # [descriptor of columns in a shared-memory ring, weights]
KEY_STEPS = "reth_b200_steps"          # [head, frames, other columns, weights]: frame-de-duplicated ingest
KEY_SRC = "reth_b200_src"              # [client id, message number]: lets the server order a loader's
                                       # request after the writes the same process sent on other sockets


def _native():
    from . import _cabi
    return _cabi.lib()


def lz4_compress(buf):
    """lz4.frame.compress(buf) (utils/pack.py:69) -> uint8 array holding one LZ4 frame."""
    global _LZ4_PKG
    if _LZ4_PKG is None:
        try:
            import lz4.frame
            _LZ4_PKG = lz4.frame
        except ImportError:
            _LZ4_PKG = False
    if _LZ4_PKG:
        return np.frombuffer(_LZ4_PKG.compress(buf, return_bytearray=True), np.uint8)
    src = np.ascontiguousarray(np.frombuffer(buf, np.uint8))
    L = _native()
    out = np.empty(L.reth_lz4f_compress_bound(src.size), np.uint8)
    n = L.reth_lz4f_compress(src.ctypes.data, src.size, out.ctypes.data, out.size)
    if n < 0:
        raise ValueError(f"lz4 frame compression failed ({n})")
    return out[:n]


_LZ4F_ERR = {-1: "truncated frame", -2: "malformed frame", -3: "checksum mismatch", -4: "output too small",
             -5: "frames with a dictionary id are not supported"}


_LZ4_PKG = None          # the `lz4` package: None = not probed yet, False = absent (a failed import is not cached by Python)
_SCRATCH = threading.local()


def _scratch(nbytes):
    """A per-thread buffer that is reused from call to call (warm pages: a fresh 3.6 MB array per message costs
    more in page faults than the decode itself)."""
    buf = getattr(_SCRATCH, "buf", None)
    if buf is None or buf.size < nbytes:
        buf = _SCRATCH.buf = np.empty(max(nbytes, 1 << 20), np.uint8)
    return buf[:nbytes]


def lz4_decompress(buf, reuse=False):
    """lz4.frame.decompress(buf) (utils/pack.py:155) -> uint8 array.  reuse: the result may live in this thread's
    scratch buffer, i.e. it is only valid until the thread's next call (callers that copy everything out at once)."""
    global _LZ4_PKG
    if _LZ4_PKG is None:
        try:
            import lz4.frame
            _LZ4_PKG = lz4.frame
        except ImportError:
            _LZ4_PKG = False
    if _LZ4_PKG:
        return np.frombuffer(_LZ4_PKG.decompress(buf, return_bytearray=True), np.uint8)
    src = np.ascontiguousarray(np.frombuffer(buf, np.uint8))
    L = _native()
    size = L.reth_lz4f_content_size(src.ctypes.data, src.size)
    if size < -1:
        raise ValueError(f"lz4 frame: {_LZ4F_ERR.get(size, size)}")
    if size > 256 * src.size + (1 << 16):                     # LZ4 cannot expand more than ~255:1
        raise ValueError("lz4 frame: the stored content size is not credible")
    cap = size if size >= 0 else max(4 * src.size, 1 << 16)
    while True:
        out = _scratch(cap) if reuse else np.empty(cap, np.uint8)
        n = L.reth_lz4f_decompress(src.ctypes.data, src.size, out.ctypes.data, cap)
        if n == -4 and cap < (1 << 40):                       # size not stored / concatenated frames: grow
            cap = max(4 * cap, 1 << 16)
            continue
        if n < 0:
            raise ValueError(f"lz4 frame: {_LZ4F_ERR.get(n, n)}")
        return out[:n]


class _Body:
    """Body segments in emission order; offsets are assigned as segments are added."""

    def __init__(self):
        self.size = 0
        self.parts = []

    def add(self, buf):
        mv = memoryview(buf).cast("B") if not isinstance(buf, np.ndarray) else memoryview(
            np.ascontiguousarray(buf).reshape(-1).view(np.uint8))
        at = self.size
        self.parts.append(mv)
        self.size += mv.nbytes
        return at, mv.nbytes


def _describe(obj, body):
    if isinstance(obj, (list, tuple)):
        return [_describe(x, body) for x in obj]
    if isinstance(obj, dict):
        return {k: _describe(v, body) for k, v in obj.items()}
    if isinstance(obj, np.ndarray):
        node = {KEY_NUMPY: True, "header": np.lib.format.header_data_from_array_1_0(obj), "raw_size": obj.nbytes}
        if obj.flags.f_contiguous and not obj.flags.c_contiguous:
            at, n = body.add(obj.T)                  # fortran data is emitted in memory order
        else:
            at, n = body.add(obj)
        node["offset"], node["length"] = at, n
        return node
    if isinstance(obj, (bytes, bytearray, memoryview)):
        node = {KEY_BYTES: True, "raw_size": memoryview(obj).nbytes}
        node["offset"], node["length"] = body.add(obj)
        return node
    if isinstance(obj, np.generic):
        return obj.item()
    return obj


def encode(obj, compress=False, columnar=False, now=None, extra=None):
    """-> bytes-like (a uint8 numpy buffer) in the reference wire format.  The body segments are
    copied ONCE, straight into the output buffer.  extra: additional header fields (the reference's
    deserialize ignores header keys it does not know)."""
    body = _Body()
    meta = _describe(obj, body)
    header = {}
    raw = None
    if compress:
        flat = np.empty(body.size, np.uint8)
        at = 0
        for mv in body.parts:
            flat[at:at + mv.nbytes] = np.frombuffer(mv, np.uint8)
            at += mv.nbytes
        raw = lz4_compress(flat)
        header[KEY_COMPRESS] = True
    header["meta"] = meta
    header["body_len"] = len(raw) if raw is not None else body.size
    header["time"] = time.time() if now is None else now
    if columnar:
        header[KEY_COLUMNAR] = True
    if extra:
        header.update(extra)
    hb = msgpack.packb(header)
    out = np.empty(4 + len(hb) + header["body_len"], np.uint8)
    out[:4] = np.frombuffer(len(hb).to_bytes(4, "big"), np.uint8)
    out[4:4 + len(hb)] = np.frombuffer(hb, np.uint8)
    at = 4 + len(hb)
    if raw is not None:
        out[at:] = np.frombuffer(raw, np.uint8)
    else:
        for mv in body.parts:
            if mv.nbytes:
                out[at:at + mv.nbytes] = np.frombuffer(mv, np.uint8)
            at += mv.nbytes
    return out.data


def _split(buf, reuse=None):
    """reuse(header) -> bool: the decompressed body may live in the thread's scratch buffer (see lz4_decompress)."""
    mv = memoryview(buf).cast("B")
    hl = int.from_bytes(mv[:4], "big")
    header = msgpack.unpackb(mv[4:4 + hl])
    body = mv[4 + hl:]
    if body.nbytes != header["body_len"]:
        raise ValueError("truncated message: body_len does not match")
    if header.get(KEY_COMPRESS):
        body = memoryview(lz4_decompress(body, reuse=bool(reuse and reuse(header))))
    return header, body, 4 + hl


def _array(node, body):
    h = node["header"]
    dt = np.lib.format.descr_to_dtype(h["descr"])
    a = np.frombuffer(body[node["offset"]:node["offset"] + node["length"]], dtype=dt)
    shape = tuple(h["shape"])
    a = a.reshape(shape[::-1]).T if h["fortran_order"] else a.reshape(shape)
    if a.nbytes != node["raw_size"]:
        raise ValueError("ndarray size mismatch")
    return a


def _rebuild(node, body):
    if isinstance(node, (list, tuple)):
        return [_rebuild(x, body) for x in node]
    if isinstance(node, dict):
        if KEY_NUMPY in node:
            return _array(node, body)
        if KEY_BYTES in node:
            return body[node["offset"]:node["offset"] + node["length"]]
        return {k: _rebuild(v, body) for k, v in node.items()}
    return node


def decode(buf):
    """Inverse of encode (utils/pack.py:148-164); arrays are views into `buf` (or into the
    decompressed body)."""
    header, body, _ = _split(buf)
    return _rebuild(header["meta"], body)


def decode_src(buf):
    """decode() plus the sender's [client id, message number] header field (None from a reference client)."""
    header, body, _ = _split(buf)
    return _rebuild(header["meta"], body), header.get(KEY_SRC)


def is_columnar(buf):
    mv = memoryview(buf).cast("B")
    hl = int.from_bytes(mv[:4], "big")
    return bool(msgpack.unpackb(mv[4:4 + hl]).get(KEY_COLUMNAR))


def is_shm(buf):
    """True for an append whose columns sit in a shared-memory ring (service_mp._ShmRing)."""
    mv = memoryview(buf).cast("B")
    hl = int.from_bytes(mv[:4], "big")
    meta = msgpack.unpackb(mv[4:4 + hl])["meta"]
    return isinstance(meta, list) and len(meta) == 2 and isinstance(meta[0], dict) and KEY_SHM in meta[0]


def open_append(buf, alloc=None):
    """An append message, header parsed once -> (kind, payload, weights, src):
    kind "shm": payload = descriptor of columns in a shared-memory ring (service_mp._ShmRing);
    kind "columns": payload = list of C-contiguous (n, ...) arrays (row or columnar wire format);
    src = the sender's [client id, message number] or None (a reference client).
    alloc(nbytes) -> writable uint8 buffer: where the columns of a "columns" message are put (pinned staging
    memory on the server's ingest threads); default: fresh arrays."""
    def plain_columns(header):
        # row / columnar appends with a destination allocator copy every byte out of the body before they return
        # (scatter / _place / np.array(weights)): the decompressed body can live in the thread's scratch buffer
        m = header["meta"]
        special = isinstance(m, list) and m and isinstance(m[0], dict) and (KEY_SHM in m[0] or KEY_STEPS in m[0])
        return alloc is not None and not special
    header, body, _ = _split(buf, plain_columns)
    meta = header["meta"]
    src = header.get(KEY_SRC)
    if isinstance(meta, list) and len(meta) == 2 and isinstance(meta[0], dict) and KEY_SHM in meta[0]:
        head, weights = _rebuild(meta, body)
        return "shm", head, np.asarray(weights), src
    if isinstance(meta, list) and len(meta) == 4 and isinstance(meta[0], dict) and KEY_STEPS in meta[0]:
        head, frames, others, weights = _rebuild(meta, body)
        return "steps", (head, frames, others), np.asarray(weights), src
    cols, weights = _decode_append(header, body, alloc)
    return "columns", cols, weights, src


def decode_append(buf):
    """An append message -> (columns, weights): columns[c] is a C-contiguous (n, ...) array.

    Accepts the reference's row format ([[row messages], weights], client/client.py:26-35) and this
    build's columnar format.  Row format, fast path: when all row messages have the same length and
    the same inner header length (same struct: always, for a fixed-size transition), the rows form
    an (n, L) byte matrix and each column is ONE strided copy out of it."""
    header, body, _ = _split(buf)
    return _decode_append(header, body)


def _place(arrays, alloc):
    """Copy arrays into one buffer obtained from alloc (columns back to back, 256-byte aligned)."""
    sizes = [(a.nbytes + 255) & ~255 for a in arrays]
    slab = np.frombuffer(alloc(sum(sizes)), np.uint8)
    out, at = [], 0
    for a, sz in zip(arrays, sizes):
        dst = slab[at:at + a.nbytes].view(a.dtype).reshape(a.shape)
        np.copyto(dst, a)
        out.append(dst)
        at += sz
    return out


def _decode_append(header, body, alloc=None):
    meta = header["meta"]
    if header.get(KEY_COLUMNAR):
        cols, weights = _rebuild(meta, body)
        cols = [np.ascontiguousarray(c) for c in cols]
        return (_place(cols, alloc) if alloc else cols), np.array(weights)
    rows, wnode = meta
    weights = _rebuild(wnode, body)
    n = len(rows)
    if n == 0:
        return [], np.asarray(weights)
    L, first = rows[0]["length"], rows[0]["offset"]
    regular = all(r["length"] == L and r["offset"] == first + i * L for i, r in enumerate(rows))
    if regular:
        mat = np.frombuffer(body[first:first + n * L], np.uint8).reshape(n, L)
        h0, _, skip = _split(mat[0])
        hl = mat[:, :4]
        if np.all(hl == hl[0]) and all(isinstance(x, dict) and KEY_NUMPY in x for x in h0["meta"]) and not any(
                x["header"]["fortran_order"] and len(x["header"]["shape"]) > 1 for x in h0["meta"]):
            # every column of every row in ONE call of the library (reth_pack_scatter_rows, no interpreter in the
            # loop, GIL released), straight into the destination buffer (pinned staging on the ingest threads)
            nodes = h0["meta"]
            ncol = len(nodes)
            lens = np.array([x["length"] for x in nodes], np.int64)
            los = np.array([skip + x["offset"] for x in nodes], np.int64)
            sizes = [(int(n * ln) + 255) & ~255 for ln in lens]
            total = sum(sizes)
            slab = np.frombuffer(alloc(total), np.uint8) if alloc else np.empty(total, np.uint8)
            import ctypes as C
            dst = (C.c_void_p * max(1, ncol))()
            cols, at = [], 0
            for c, node in enumerate(nodes):
                hd = node["header"]
                dt = np.lib.format.descr_to_dtype(hd["descr"])
                view = slab[at:at + n * int(lens[c])]
                dst[c] = view.ctypes.data
                cols.append(view.view(dt).reshape((n,) + tuple(hd["shape"])))
                at += sizes[c]
            got = _native().reth_pack_scatter_rows(mat.ctypes.data, n, L, ncol, los.ctypes.data, lens.ctypes.data, dst)
            if got != int((lens * n).sum()):
                raise ValueError("row message: a column lies outside its row")
            return cols, np.array(weights)
    # general path: decode row by row
    per_row = [decode(body[r["offset"]:r["offset"] + r["length"]]) for r in rows]
    ncol = len(per_row[0])
    cols = [np.stack([np.asarray(per_row[i][c]) for i in range(n)]) for c in range(ncol)]
    return (_place(cols, alloc) if alloc else cols), np.array(weights)


def encode_append_rows(columns, weights, compress=False, now=None, extra=None):
    """What the reference's Client.append puts on the wire (client/client.py:26-35)."""
    n = len(weights)
    rows = [encode([col[i, ...] for col in columns], now=now) for i in range(n)]
    return encode([rows, weights], compress=compress, now=now, extra=extra)


def encode_append_columns(columns, weights, compress=False, now=None, extra=None):
    return encode([list(columns), weights], compress=compress, columnar=True, now=now, extra=extra)
