// LZ4 frame codec (host code) for the ingest wire format.
//
// The reference's actors send lz4-framed message bodies: test/apex-dqn/worker.py:60 calls
// Client.append(..., compress=True), which wraps the body with lz4.frame.compress
// (reth_buffer/reth_buffer/utils/pack.py:63-84) and the server undoes it with lz4.frame.decompress
// (utils/pack.py:152-156).  The python `lz4` package (pinned 3.1.3 in the reference's poetry.lock)
// is a third-party dependency that is absent from this image, so the published LZ4 frame format
// (v1.6.x) and block format are restated here:
//
//   frame  = magic 0x184D2204 | FLG | BD | [content size u64] | [dict id u32] | HC
//            | { block size u32 (bit 31: stored uncompressed) | data | [xxh32 of data] }*
//            | end mark u32 0 | [xxh32 of the content]
//   FLG    = version 01 <<6 | B.Indep <<5 | B.Checksum <<4 | C.Size <<3 | C.Checksum <<2 | DictID
//   HC     = (xxh32(FLG .. before HC, seed 0) >> 8) & 0xff
//   block  = sequences of  token (literal len <<4 | match len - 4) [len bytes 255...] literals
//            offset u16 LE [match len bytes 255...];  the last sequence stops after its literals.
//   skippable frames (magic 0x184D2A50..5F | size u32 | bytes) are stepped over; frames may be
//   concatenated.
//
// Linked blocks (python-lz4's default, block_linked=True) may reference up to 64 KB of the output
// that precedes them; decoding into one contiguous buffer gives that for free.
// The compressor is a greedy single-probe hash matcher writing independent 64 KB blocks with the
// content size stored, which every conforming decoder accepts; it exists so that this build's own
// clients can send compress=True and for round-trip tests.
#pragma once
#include <cstdint>
#include <cstring>

namespace lz4f {

constexpr uint32_t kMagic = 0x184D2204u;
constexpr uint32_t kSkipMagicLo = 0x184D2A50u;

enum : int64_t {
    kErrTruncated = -1,      // input ends inside a header, block or sequence
    kErrFormat = -2,         // bad magic, version, reserved bits, block size code or offset
    kErrChecksum = -3,       // header, block or content checksum mismatch
    kErrCapacity = -4,       // output does not fit
    kErrUnsupported = -5,    // dictionary id present
};

inline uint32_t rd32(const uint8_t *p) { return (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24; }
inline uint64_t rd64(const uint8_t *p) { return (uint64_t)rd32(p) | (uint64_t)rd32(p + 4) << 32; }
inline void wr32(uint8_t *p, uint32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24); }
inline uint32_t rotl(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

// XXH32, the checksum of the frame format
inline uint32_t xxh32(const void *data, size_t len, uint32_t seed)
{
    constexpr uint32_t P1 = 2654435761u, P2 = 2246822519u, P3 = 3266489917u, P4 = 668265263u, P5 = 374761393u;
    const uint8_t *p = (const uint8_t *)data, *end = p + len;
    uint32_t h;
    if (len >= 16) {
        uint32_t v1 = seed + P1 + P2, v2 = seed + P2, v3 = seed, v4 = seed - P1;
        const uint8_t *lim = end - 16;
        do {
            v1 = rotl(v1 + rd32(p) * P2, 13) * P1;
            v2 = rotl(v2 + rd32(p + 4) * P2, 13) * P1;
            v3 = rotl(v3 + rd32(p + 8) * P2, 13) * P1;
            v4 = rotl(v4 + rd32(p + 12) * P2, 13) * P1;
            p += 16;
        } while (p <= lim);
        h = rotl(v1, 1) + rotl(v2, 7) + rotl(v3, 12) + rotl(v4, 18);
    } else {
        h = seed + P5;
    }
    h += (uint32_t)len;
    for (; p + 4 <= end; p += 4) h = rotl(h + rd32(p) * P3, 17) * P4;
    for (; p < end; ++p) h = rotl(h + (uint32_t)*p * P5, 11) * P1;
    h ^= h >> 15; h *= P2; h ^= h >> 13; h *= P3; h ^= h >> 16;
    return h;
}

struct FrameInfo {
    int64_t header_bytes = 0;
    int64_t content_size = -1;      // -1: not stored
    int64_t block_max = 0;
    bool block_indep = false, block_checksum = false, content_checksum = false;
};

// -> 0 or an error; `src` points at the magic number
inline int64_t parse_header(const uint8_t *src, int64_t n, FrameInfo *fi)
{
    if (n < 7) return kErrTruncated;
    if (rd32(src) != kMagic) return kErrFormat;
    const uint8_t flg = src[4], bd = src[5];
    if ((flg >> 6) != 1 || (flg & 0x02) || (bd & 0x8f)) return kErrFormat;
    if (flg & 0x01) return kErrUnsupported;
    const int code = (bd >> 4) & 7;
    if (code < 4) return kErrFormat;
    fi->block_max = (int64_t)1 << (8 + 2 * code);           // 4: 64 KB, 5: 256 KB, 6: 1 MB, 7: 4 MB
    fi->block_indep = flg & 0x20; fi->block_checksum = flg & 0x10; fi->content_checksum = flg & 0x04;
    int64_t at = 6;
    if (flg & 0x08) {
        if (n < at + 8 + 1) return kErrTruncated;
        const uint64_t cs = rd64(src + at);
        if (cs >> 62) return kErrFormat;
        fi->content_size = (int64_t)cs;
        at += 8;
    }
    if (n < at + 1) return kErrTruncated;
    if (((xxh32(src + 4, (size_t)(at - 4), 0) >> 8) & 0xff) != src[at]) return kErrChecksum;
    fi->header_bytes = at + 1;
    return 0;
}

// One block of sequences: src[0, n) -> dst[at, ...); dst[0, at) is the history matches may reach
// into.  -> new output position or an error.
inline int64_t decode_block(const uint8_t *src, int64_t n, uint8_t *dst, int64_t at, int64_t cap)
{
    const uint8_t *ip = src, *iend = src + n;
    while (ip < iend) {
        const unsigned token = *ip++;
        int64_t lit = token >> 4;
        // short sequence (the common case): fixed-size copies, no length bytes.  Needs 16 literal
        // bytes + the offset readable and 16 + 18 bytes of output room; the last sequence of a block
        // (literals only, fewer than 18 input bytes left) never qualifies.
        if (lit < 15 && iend - ip >= 18 && cap - at >= 34) {
            memcpy(dst + at, ip, 16);
            const int64_t off = (int64_t)ip[lit] | (int64_t)ip[lit + 1] << 8;
            const int64_t ml = token & 15;
            if (ml < 15 && off >= 8 && off <= at + lit) {
                at += lit; ip += lit + 2;
                uint8_t *d = dst + at;
                const uint8_t *m = d - off;
                memcpy(d, m, 8); memcpy(d + 8, m + 8, 8); memcpy(d + 16, m + 16, 2);
                at += ml + 4;
                continue;
            }
        }
        if (lit == 15) {
            unsigned b;
            do {
                if (ip >= iend) return kErrTruncated;
                b = *ip++;
                lit += b;
            } while (b == 255);
        }
        if (lit > iend - ip) return kErrTruncated;
        if (lit > cap - at) return kErrCapacity;
        memcpy(dst + at, ip, (size_t)lit);
        at += lit; ip += lit;
        if (ip == iend) break;                               // last sequence: literals only
        if (iend - ip < 2) return kErrTruncated;
        const int64_t off = (int64_t)ip[0] | (int64_t)ip[1] << 8;
        ip += 2;
        if (off == 0 || off > at) return kErrFormat;
        int64_t ml = token & 15;
        if (ml == 15) {
            unsigned b;
            do {
                if (ip >= iend) return kErrTruncated;
                b = *ip++;
                ml += b;
            } while (b == 255);
        }
        ml += 4;
        if (ml > cap - at) return kErrCapacity;
        uint8_t *d = dst + at;
        const uint8_t *m = d - off;
        if (off >= ml) {
            memcpy(d, m, (size_t)ml);
        } else if (off == 1) {
            memset(d, *m, (size_t)ml);
        } else {                                             // overlapping match: the pattern repeats
            int64_t done = 0;
            while (done < ml) {                              // copy in chunks that double the period
                const int64_t span = (off + done) < (ml - done) ? (off + done) : (ml - done);
                // source [m, m + span) lies entirely before d + done when span <= off + done
                memmove(d + done, m, (size_t)span);
                done += span;
            }
        }
        at += ml;
    }
    return at;
}

// Total content size when every frame stores it, else -1 (unknown); <-1: error.
inline int64_t content_size(const uint8_t *src, int64_t n)
{
    // only the first frame is inspected: python-lz4 writes one frame per message
    if (n >= 8 && (rd32(src) & 0xfffffff0u) == kSkipMagicLo) return -1;
    FrameInfo fi;
    const int64_t rc = parse_header(src, n, &fi);
    if (rc) return rc < -1 ? rc : -2;
    return fi.content_size;
}

// Decodes every frame in src[0, n) into dst[0, cap).  -> bytes written or an error.
// verify: check block / content checksums when the frame carries them.
inline int64_t decompress(const uint8_t *src, int64_t n, uint8_t *dst, int64_t cap, bool verify = true)
{
    int64_t ip = 0, out = 0;
    while (ip < n) {
        if (n - ip < 4) return kErrTruncated;
        const uint32_t magic = rd32(src + ip);
        if ((magic & 0xfffffff0u) == kSkipMagicLo) {
            if (n - ip < 8) return kErrTruncated;
            const int64_t sz = rd32(src + ip + 4);
            if (sz > n - ip - 8) return kErrTruncated;
            ip += 8 + sz;
            continue;
        }
        FrameInfo fi;
        int64_t rc = parse_header(src + ip, n - ip, &fi);
        if (rc) return rc;
        ip += fi.header_bytes;
        const int64_t frame_start = out;
        for (;;) {
            if (n - ip < 4) return kErrTruncated;
            const uint32_t bs = rd32(src + ip);
            ip += 4;
            if (bs == 0) break;
            const int64_t len = bs & 0x7fffffffu;
            if (len > fi.block_max) return kErrFormat;
            if (len > n - ip) return kErrTruncated;
            if (bs & 0x80000000u) {
                if (len > cap - out) return kErrCapacity;
                memcpy(dst + out, src + ip, (size_t)len);
                out += len;
            } else {
                // history: the whole frame so far when blocks are linked, nothing otherwise
                const int64_t base = fi.block_indep ? out : frame_start;
                int64_t lim = cap - base;
                if (out - base + fi.block_max < lim) lim = out - base + fi.block_max;
                rc = decode_block(src + ip, len, dst + base, out - base, lim);
                if (rc < 0) return rc;
                out = base + rc;
            }
            if (fi.block_checksum) {
                if (n - ip - len < 4) return kErrTruncated;
                if (verify && xxh32(src + ip, (size_t)len, 0) != rd32(src + ip + len)) return kErrChecksum;
                ip += 4;
            }
            ip += len;
        }
        if (fi.content_checksum) {
            if (n - ip < 4) return kErrTruncated;
            if (verify && xxh32(dst + frame_start, (size_t)(out - frame_start), 0) != rd32(src + ip)) return kErrChecksum;
            ip += 4;
        }
        if (fi.content_size >= 0 && out - frame_start != fi.content_size) return kErrFormat;
    }
    return out;
}

constexpr int64_t kBlock = 65536;

inline int64_t compress_bound(int64_t n)
{
    const int64_t blocks = (n + kBlock - 1) / kBlock;
    return 4 + 2 + 8 + 1 + blocks * 4 + n + n / 255 + 16 * (blocks + 1) + 4;
}

// One independent block.  -> compressed size, or 0 when it does not beat storing the block raw.
inline int64_t encode_block(const uint8_t *src, int64_t n, uint8_t *dst, int64_t cap)
{
    constexpr int kHashBits = 13;
    constexpr int64_t kMfLimit = 12, kLastLiterals = 5;
    uint16_t table[1 << kHashBits];
    memset(table, 0, sizeof table);
    uint8_t *op = dst, *const oend = dst + cap;
    int64_t anchor = 0, i = 0;
    auto emit = [&](int64_t lit, int64_t ml, int64_t off) -> bool {   // ml = 0: final literals
        if (op + 1 + lit / 255 + 1 + lit + 2 + ml / 255 + 1 > oend) return false;
        uint8_t *tok = op++;
        if (lit >= 15) {
            *tok = 15 << 4;
            int64_t r = lit - 15;
            for (; r >= 255; r -= 255) *op++ = 255;
            *op++ = (uint8_t)r;
        } else {
            *tok = (uint8_t)(lit << 4);
        }
        memcpy(op, src + anchor, (size_t)lit);
        op += lit;
        if (ml) {
            *op++ = (uint8_t)off; *op++ = (uint8_t)(off >> 8);
            int64_t r = ml - 4;
            if (r >= 15) {
                *tok |= 15;
                r -= 15;
                for (; r >= 255; r -= 255) *op++ = 255;
                *op++ = (uint8_t)r;
            } else {
                *tok |= (uint8_t)r;
            }
        }
        return true;
    };
    if (n >= kMfLimit + 1) {
        const int64_t match_limit = n - kMfLimit;            // a match may not start after this
        while (i < match_limit) {
            const uint32_t seq = rd32(src + i);
            const uint32_t hsh = (seq * 2654435761u) >> (32 - kHashBits);
            const int64_t cand = table[hsh];
            table[hsh] = (uint16_t)i;
            if (cand < i && rd32(src + cand) == seq && (cand > 0 || i > 0) && i - cand <= 65535 && i != cand) {
                int64_t ml = 4;
                const int64_t max_ml = n - kLastLiterals - i;
                while (ml < max_ml && src[cand + ml] == src[i + ml]) ++ml;
                if (ml >= 4 && ml <= max_ml) {
                    if (!emit(i - anchor, ml, i - cand)) return 0;
                    i += ml;
                    anchor = i;
                    continue;
                }
            }
            ++i;
        }
    }
    if (!emit(n - anchor, 0, 0)) return 0;
    const int64_t sz = op - dst;
    return sz < n ? sz : 0;
}

// -> frame bytes written or kErrCapacity
inline int64_t compress(const uint8_t *src, int64_t n, uint8_t *dst, int64_t cap)
{
    if (cap < compress_bound(n)) return kErrCapacity;
    uint8_t *op = dst;
    wr32(op, kMagic); op += 4;
    op[0] = 0x68;                                            // version 01, independent blocks, content size
    op[1] = 0x40;                                            // 64 KB blocks
    wr32(op + 2, (uint32_t)((uint64_t)n & 0xffffffffu)); wr32(op + 6, (uint32_t)((uint64_t)n >> 32));
    op[10] = (uint8_t)((xxh32(op, 10, 0) >> 8) & 0xff);
    op += 11;
    for (int64_t at = 0; at < n; at += kBlock) {
        const int64_t len = n - at < kBlock ? n - at : kBlock;
        const int64_t c = encode_block(src + at, len, op + 4, len);
        if (c > 0) {
            wr32(op, (uint32_t)c);
            op += 4 + c;
        } else {
            wr32(op, (uint32_t)len | 0x80000000u);
            memcpy(op + 4, src + at, (size_t)len);
            op += 4 + len;
        }
    }
    wr32(op, 0); op += 4;
    return op - dst;
}

}  // namespace lz4f
