// copy.cuh -- HBM ring storage movers (sm_100a): batched append, slot scatter,
// and the sampled-row gather into a contiguous learner batch.
//
// Storage is columnar like the reference's (one array per column:
// reth/reth/buffer/buffer.py:19-23; the LMDB rows of server/main_loop.py:55-57
// become ring[c][slot]).  An Atari frame-stack row is 28 224 B = 1764 x 16 B,
// 32 B-sector aligned, so all bulk traffic is 16-byte vectors; small scalar
// columns (a, r, done) take a word path.  These kernels are pure copies: the
// roofline is HBM bandwidth, algorithmic bytes = 2 x bytes moved.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace reth {

constexpr int kMaxCols = 16;
constexpr int kCopyThreads = 256;

__device__ __forceinline__ uint4 ld_stream(const uint4 *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// 4 packed uint8 -> 4 float32 (exact); unit: IEEE division by 255 like numpy's float32 x / 255.0
__device__ __forceinline__ uint4 u8x4_to_f32(uint32_t w, bool unit)
{
    float f0 = (float)(w & 0xffu), f1 = (float)((w >> 8) & 0xffu), f2 = (float)((w >> 16) & 0xffu), f3 = (float)(w >> 24);
    if (unit) {
        f0 = __fdiv_rn(f0, 255.0f); f1 = __fdiv_rn(f1, 255.0f); f2 = __fdiv_rn(f2, 255.0f); f3 = __fdiv_rn(f3, 255.0f);
    }
    return make_uint4(__float_as_uint(f0), __float_as_uint(f1), __float_as_uint(f2), __float_as_uint(f3));
}
__device__ __forceinline__ uint32_t ld_stream32(const uint32_t *p)
{
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream(uint4 *p, const uint4 &v)
{
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// Cooperative copy of `bytes` by `nthr` threads (this thread = `t`), widest
// vector both pointers and the length allow.  4 independent 16 B loads in flight.
__device__ __forceinline__ void span_copy(uint8_t *dst, const uint8_t *src, int64_t bytes,
                                          int64_t t, int64_t nthr)
{
    const uint64_t al = (uint64_t)dst | (uint64_t)src | (uint64_t)bytes;
    if ((al & 15ull) == 0) {
        const uint4 *s = (const uint4 *)src;
        uint4 *d = (uint4 *)dst;
        const int64_t n = bytes >> 4;
        int64_t j = t;
        for (; j + 3 * nthr < n; j += 4 * nthr) {
            const uint4 v0 = ld_stream(s + j), v1 = ld_stream(s + j + nthr);
            const uint4 v2 = ld_stream(s + j + 2 * nthr), v3 = ld_stream(s + j + 3 * nthr);
            st_stream(d + j, v0); st_stream(d + j + nthr, v1);
            st_stream(d + j + 2 * nthr, v2); st_stream(d + j + 3 * nthr, v3);
        }
        for (; j < n; j += nthr) st_stream(d + j, ld_stream(s + j));
    } else if ((al & 3ull) == 0) {
        const uint32_t *s = (const uint32_t *)src;
        uint32_t *d = (uint32_t *)dst;
        for (int64_t j = t; j < (bytes >> 2); j += nthr) d[j] = s[j];
    } else {
        for (int64_t j = t; j < bytes; j += nthr) dst[j] = src[j];
    }
}

// ------------------------------------------------------------ append (FIFO) --
// Slots start, start+1, ... mod capacity (fifo_policy.py:11-18): at most two
// contiguous spans per column (buffer.py:68-74).
struct AppendArgs {
    const uint8_t *src[kMaxCols];
    uint8_t *ring[kMaxCols];
    int64_t rb[kMaxCols];
    int32_t ncols;
    int64_t start, n, capacity;
};

__global__ void __launch_bounds__(kCopyThreads) ring_append_kernel(const AppendArgs a)
{
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
    const int64_t len1 = min(a.n, a.capacity - a.start), len2 = a.n - len1;
    for (int c = 0; c < a.ncols; ++c) {
        const int64_t rb = a.rb[c];
        span_copy(a.ring[c] + a.start * rb, a.src[c], len1 * rb, t, nthr);
        if (len2 > 0) span_copy(a.ring[c], a.src[c] + len1 * rb, len2 * rb, t, nthr);
    }
}

// Tiled variant for 16-byte-aligned spans: one CTA per 16 KB tile (256 threads x 4 vectors in flight),
// so each CTA streams one contiguous piece like the gather does -- the grid-strided kernel above spreads
// a thread's four loads 4.8 MB apart and reaches 0.85 of the copy peak, this one is on par with the
// driver's device-to-device copy (profiles/r01_append_ab.json).
constexpr int kAppendTileUnits = 2048;
struct AppendSeg {
    uint8_t *dst;
    const uint8_t *src;
    int64_t units;        // 16 B units
    int64_t tile_base;    // first tile of this segment
};
struct AppendTiledArgs {
    AppendSeg seg[2 * kMaxCols];
    int32_t nseg;
    int64_t tiles;
};

__global__ void __launch_bounds__(kCopyThreads) ring_append_tiled_kernel(const AppendTiledArgs a)
{
    for (int64_t tile = blockIdx.x; tile < a.tiles; tile += gridDim.x) {
        int sidx = 0;
#pragma unroll 1
        for (int k = 1; k < a.nseg; ++k)
            if (tile >= a.seg[k].tile_base) sidx = k;
        const AppendSeg &sg = a.seg[sidx];
        const int64_t u0 = (tile - sg.tile_base) * kAppendTileUnits;
        const int n = (int)min((int64_t)kAppendTileUnits, sg.units - u0);
        const uint4 *s = (const uint4 *)sg.src + u0;
        uint4 *d = (uint4 *)sg.dst + u0;
        const int j = threadIdx.x;
        constexpr int kPer = kAppendTileUnits / kCopyThreads;
        uint4 v[kPer];
#pragma unroll
        for (int q = 0; q < kPer; ++q)
            if (j + q * kCopyThreads < n) v[q] = ld_stream(s + j + q * kCopyThreads);
#pragma unroll
        for (int q = 0; q < kPer; ++q)
            if (j + q * kCopyThreads < n) st_stream(d + j + q * kCopyThreads, v[q]);
    }
}

// -------------------------------------------------- row gather / row scatter --
struct RowCol {
    uint8_t *ring;        // ring column base
    uint8_t *batch;       // contiguous batch column base (gather dst / scatter src)
    int64_t rb;           // row bytes
    int32_t chunks;       // work items per row (wide columns), 0 = narrow column
    int32_t units;        // 16 B units per chunk
    int32_t cast;         // gather only: 0 raw, 1 u8 -> f32, 2 u8 -> f32 / 255 (batch rows are 4*rb bytes)
    int64_t item_base;    // first work item of this column
    int32_t chunk_base;   // chunks of the wide columns before this one (row-major item order, kDense)
};

struct RowArgs {
    RowCol col[kMaxCols];
    int32_t ncols;
    const int64_t *idx;        // slot per batch row
    int32_t batch;
    const int64_t *meta;       // optional {first,count}: rows = meta[1]
    int32_t dst_row_from_meta; // batch row of local row i is meta[0] + i (cross-shard assembly into a peer's batch)
    int64_t wide_items;        // items over all wide columns (batch-relative layout)
    int32_t narrow_groups;     // ceil(batch / kCopyThreads) when narrow columns exist
    int32_t chunks_per_row;    // chunks of one row over all wide columns
    int32_t share_div;         // host-side launch hint: about batch / share_div rows are valid (Mode G: n_shards), 0 = all
    // scatter only: last-writer-wins arbitration (same stamps as the tree update)
    const uint32_t *stamp;
    int64_t capacity;
    int32_t *err;
};

// kScatter=false: batch[i] <- ring[idx[i]]   (numpy_loader.py:35-50, buffer.py:92-93)
// kScatter=true : ring[idx[i]] <- batch[i]   (append to policy-chosen slots)
// kWiden (gather only): some column is emitted as float32 (RowCol::cast); a separate instantiation so
// that the plain byte mover keeps its register budget.
// kDense (gather of a Mode G slot): the launch covers `batch` rows but only the first meta[1] belong to this
// shard (1/n_shards of a global batch on average).  Items are then numbered row-major over the VALID rows only,
// so surplus CTAs leave after one comparison -- with the column-major numbering every skipped item costs its
// CTA a column search and a 64-bit division in all eight warps, which at 8 shards is more issue time than
// the copy itself.
// (Tried and rejected, profiles/r02_lean_gather.jsonl: a <= 32-register build capped at 6 CTAs per SM through a
// dynamic shared-memory request, to keep 16 K registers of every SM free for the tree kernels.  The request alone
// costs the gather a quarter of its rate -- 41 vs 54 M rows/s gather-only -- and uncapped the 32-register build
// is 5 % slower than this one.)
template <bool kScatter, bool kWiden = false, bool kDense = false>
__global__ void __launch_bounds__(kCopyThreads, kWiden ? 4 : 6) row_move_kernel(const RowArgs a)
{
    const int rows = a.meta ? (int)a.meta[1] : a.batch;
    const int64_t row0 = (a.meta && a.dst_row_from_meta) ? a.meta[0] : 0;
    const int64_t wide_items = kDense ? (int64_t)rows * a.chunks_per_row : a.wide_items;
    const int64_t total = wide_items + (kDense ? (a.narrow_groups ? (rows + kCopyThreads - 1) / kCopyThreads : 0)
                                               : a.narrow_groups);
    for (int64_t item = blockIdx.x; item < total; item += gridDim.x) {
        if (item < wide_items) {
            int c = 0, row, chunk;
            if (kDense) {
                row = (int)(item / a.chunks_per_row);
                const int k = (int)(item - (int64_t)row * a.chunks_per_row);
#pragma unroll 1
                for (int q = 0; q < a.ncols; ++q)
                    if (a.col[q].chunks > 0 && k >= a.col[q].chunk_base) c = q;
                chunk = k - a.col[c].chunk_base;
            } else {
#pragma unroll 1
                for (int k = 0; k < a.ncols; ++k)
                    if (a.col[k].chunks > 0 && item >= a.col[k].item_base) c = k;
                const int64_t rel = item - a.col[c].item_base;
                row = (int)(rel / a.col[c].chunks);
                chunk = (int)(rel - (int64_t)row * a.col[c].chunks);
            }
            const RowCol &col = a.col[c];
            if (row >= rows) continue;
            const int64_t slot = a.idx[row];
            if (slot < 0 || slot >= a.capacity) { if (threadIdx.x == 0) atomicOr(a.err, 1); continue; }
            if (kScatter && a.stamp && a.stamp[slot] != 1u + (uint32_t)row) continue;
            const int64_t upr = col.rb >> 4;
            const int64_t u0 = (int64_t)chunk * col.units;
            const int64_t n = min((int64_t)col.units, upr - u0);
            const uint4 *rp = (const uint4 *)(col.ring + slot * col.rb) + u0;
            const uint4 *bp = (const uint4 *)(col.batch + ((int64_t)row + row0) * col.rb) + u0;
            if (kWiden && !kScatter && col.cast) {
                // widening gather: lane = one 32-bit word (4 uint8) in, one 16 B vector (4 float32) out,
                // so a warp reads 128 contiguous bytes and writes 512 contiguous bytes per instruction
                uint4 *d4 = (uint4 *)(col.batch + ((int64_t)row + row0) * col.rb * 4) + u0 * 4;
                const uint32_t *w32 = (const uint32_t *)rp;
                const bool unit = col.cast == 2;
                const int64_t n4 = n * 4;
                int64_t j = threadIdx.x;
                for (; j + 3 * kCopyThreads < n4; j += 4 * kCopyThreads) {
                    const uint32_t a0 = ld_stream32(w32 + j), a1 = ld_stream32(w32 + j + kCopyThreads);
                    const uint32_t a2 = ld_stream32(w32 + j + 2 * kCopyThreads), a3 = ld_stream32(w32 + j + 3 * kCopyThreads);
                    st_stream(d4 + j, u8x4_to_f32(a0, unit));
                    st_stream(d4 + j + kCopyThreads, u8x4_to_f32(a1, unit));
                    st_stream(d4 + j + 2 * kCopyThreads, u8x4_to_f32(a2, unit));
                    st_stream(d4 + j + 3 * kCopyThreads, u8x4_to_f32(a3, unit));
                }
                for (; j < n4; j += kCopyThreads) st_stream(d4 + j, u8x4_to_f32(ld_stream32(w32 + j), unit));
                continue;
            }
            const uint4 *s = kScatter ? bp : rp;
            uint4 *d = (uint4 *)(kScatter ? rp : bp);
            int64_t j = threadIdx.x;
            for (; j + 3 * kCopyThreads < n; j += 4 * kCopyThreads) {
                const uint4 v0 = ld_stream(s + j), v1 = ld_stream(s + j + kCopyThreads);
                const uint4 v2 = ld_stream(s + j + 2 * kCopyThreads);
                const uint4 v3 = ld_stream(s + j + 3 * kCopyThreads);
                st_stream(d + j, v0); st_stream(d + j + kCopyThreads, v1);
                st_stream(d + j + 2 * kCopyThreads, v2); st_stream(d + j + 3 * kCopyThreads, v3);
            }
            // tail: up to 3 predicated loads issued together
            uint4 v[3];
            int cnt = 0;
#pragma unroll
            for (int q = 0; q < 3; ++q)
                if (j + q * kCopyThreads < n) { v[q] = ld_stream(s + j + q * kCopyThreads); cnt = q + 1; }
#pragma unroll
            for (int q = 0; q < 3; ++q)
                if (q < cnt) st_stream(d + j + q * kCopyThreads, v[q]);
        } else {
            // narrow / unaligned columns: one thread per row, all such columns
            const int row = (int)(item - wide_items) * kCopyThreads + threadIdx.x;
            if (row >= rows) continue;
            const int64_t slot = a.idx[row];
            if (slot < 0 || slot >= a.capacity) { atomicOr(a.err, 1); continue; }
            if (kScatter && a.stamp && a.stamp[slot] != 1u + (uint32_t)row) continue;
            for (int c = 0; c < a.ncols; ++c) {
                const RowCol &col = a.col[c];
                if (col.chunks != 0) continue;
                uint8_t *rp = col.ring + slot * col.rb;
                if (kWiden && !kScatter && col.cast) {
                    float *fp = (float *)(col.batch + ((int64_t)row + row0) * col.rb * 4);
                    for (int64_t q = 0; q < col.rb; ++q)
                        fp[q] = col.cast == 2 ? __fdiv_rn((float)rp[q], 255.0f) : (float)rp[q];
                    continue;
                }
                uint8_t *bp = col.batch + ((int64_t)row + row0) * col.rb;
                span_copy(kScatter ? rp : bp, kScatter ? bp : rp, col.rb, 0, 1);
            }
        }
    }
}

// ------------------------------------------- gather of frame-reference columns --
// Frame de-duplication (the reference's FrameStack re-stores every frame k times,
// reth/reth/env/util.py:179-209, its LazyFrames remedy is commented out): single frames live ONCE in
// a frame ring; a frame-reference column of the row struct stores K int32 frame slots per
// transition and the gather expands them into the K-frame stack the learner expects:
//   batch[row][k] = frames[ ring_refs[idx[row]][k] ]      (optionally widened to float32)
// One work item = (column, row, k, chunk of the frame), 256 threads moving 16-byte vectors.
struct FrameCol {
    const uint8_t *refs;  // ring column of the references: int32[K] per slot (direct: unused)
    const uint8_t *src;   // the frame ring; direct: the column's own ring, "frame" = the stored row
    uint8_t *batch;       // batch column base: K * fb (x4 when cast) bytes per row
    int64_t rb;           // bytes per slot of the reference column = 4 * K
    int64_t fb;           // bytes per frame (multiple of 16)
    int64_t fcap;         // frames in the source ring
    int32_t direct;       // 1: no indirection, frame slot = row slot, K = 1 (a stored wide column being widened)
    int32_t K;
    int32_t cast;         // 0 raw, 1 u8 -> f32, 2 u8 -> f32 / 255
    int32_t split;        // pieces per frame: K * split pieces are dealt to the warps of the CTA
};

struct FrameArgs {
    FrameCol col[kMaxCols];
    int32_t ncols;
    const int64_t *idx;
    int32_t batch;
    const int64_t *meta;
    int32_t dst_row_from_meta;
    int64_t capacity;
    int32_t *err;
    // liveness of frame references: row_oldest[slot] = sequence number of the oldest frame the stored row refers
    // to (frame_mark_kernel); a frame is overwritten once frame_capacity later frames arrived
    const int64_t *row_oldest;
    int64_t frames_appended, frame_capacity;
};
constexpr int kErrStaleFrame = 0x200;

// After rows were stored at `n` slots (FIFO: (first + i) % capacity, or an explicit list): record for each the
// sequence number of the OLDEST frame its frame-reference columns point at.  Frames are numbered in arrival order
// (frame `seq` lives in ring slot seq % frame_capacity), so a reference r made when `frames_appended` frames had
// arrived names frame  frames_appended - 1 - ((frame_tail - 1 - r) mod frame_capacity).  The gather compares this
// against the frames appended since: a transition that outlived one of its frames is reported, never handed out.
struct MarkArgs {
    const uint8_t *refs[kMaxCols];   // ring columns holding int32[K] references
    int64_t rb[kMaxCols];
    int32_t K[kMaxCols];
    int32_t ncols;
    const int64_t *slots;            // explicit slots, or nullptr: (first + i) % capacity
    int64_t first, n, capacity;
    int64_t frame_tail, frames_appended, frame_capacity;
    int64_t *row_oldest;
    int32_t *err;
};

__global__ void frame_mark_kernel(const MarkArgs a)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const int64_t slot = a.slots ? a.slots[i] : (a.first + i) % a.capacity;
    if (slot < 0 || slot >= a.capacity) return;          // reported by the row mover
    int64_t max_age = 0;
    for (int c = 0; c < a.ncols; ++c) {
        const int32_t *r = (const int32_t *)(a.refs[c] + slot * a.rb[c]);
        for (int k = 0; k < a.K[c]; ++k) {
            const int64_t ref = r[k];
            if (ref < 0 || ref >= a.frame_capacity) { atomicOr(a.err, 1); continue; }
            int64_t age = (a.frame_tail - 1 - ref) % a.frame_capacity;
            if (age < 0) age += a.frame_capacity;
            max_age = age > max_age ? age : max_age;
        }
    }
    a.row_oldest[slot] = a.frames_appended - 1 - max_age;
}

// Also the widening gather of stored wide columns (direct = 1): the same warp-per-part streaming, which
// keeps 6 CTAs per SM resident where the general row mover's widening path manages 4.
// One work item = the K-frame stack of (batch row, frame column); items of the same row are adjacent in
// the grid, so the s0 and s1 stacks of a transition (which share K - 1 frames) are read at about the same
// time and the second read of a shared frame hits L2.  Inside an item every warp owns (frame k, part of
// it): the reference is one uniform load per warp, then 32 lanes stream 16-byte vectors, 4 in flight each.
__global__ void __launch_bounds__(kCopyThreads, 6) frame_gather_kernel(const FrameArgs a)
{
    constexpr int kWarps = kCopyThreads / 32;
    const int rows = a.meta ? (int)a.meta[1] : a.batch;
    const int64_t row0 = (a.meta && a.dst_row_from_meta) ? a.meta[0] : 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t items = (int64_t)a.batch * a.ncols;
    for (int64_t item = blockIdx.x; item < items; item += gridDim.x) {
        const int row = (int)(item / a.ncols);
        const FrameCol &col = a.col[(int)(item - (int64_t)row * a.ncols)];
        if (row >= rows) continue;
        const int64_t slot = a.idx[row];
        if (slot < 0 || slot >= a.capacity) { if (threadIdx.x == 0) atomicOr(a.err, 1); continue; }
        const int32_t *refs = (const int32_t *)(col.refs + slot * col.rb);
        const int64_t out_row = (int64_t)row + row0;
        if (!col.direct && a.row_oldest && a.frames_appended - a.row_oldest[slot] > a.frame_capacity) {
            if (threadIdx.x == 0) atomicOr(a.err, kErrStaleFrame);      // a referenced frame has been overwritten
            continue;
        }
        const int pieces = col.K * col.split;
        const int upf = (int)(col.fb >> 4);
        for (int p = warp; p < pieces; p += kWarps) {
            const int k = p / col.split, part = p - k * col.split;
            const int64_t ref = col.direct ? slot : (int64_t)__ldg(refs + k);
            if (ref < 0 || ref >= col.fcap) { if (lane == 0) atomicOr(a.err, 1); continue; }
            const int u0 = (int)((int64_t)upf * part / col.split), u1 = (int)((int64_t)upf * (part + 1) / col.split);
            if (col.cast) {
                // lane = one 32-bit word (4 uint8) in, one 16 B vector (4 float32) out
                const uint32_t *w32 = (const uint32_t *)(col.src + ref * col.fb) + u0 * 4;
                uint4 *d4 = (uint4 *)(col.batch + (out_row * col.K + k) * col.fb * 4) + u0 * 4;
                const bool unit = col.cast == 2;
                const int n4 = (u1 - u0) * 4;
                int j = lane;
                for (; j + 96 < n4; j += 128) {
                    const uint32_t a0 = ld_stream32(w32 + j), a1 = ld_stream32(w32 + j + 32);
                    const uint32_t a2 = ld_stream32(w32 + j + 64), a3 = ld_stream32(w32 + j + 96);
                    st_stream(d4 + j, u8x4_to_f32(a0, unit));
                    st_stream(d4 + j + 32, u8x4_to_f32(a1, unit));
                    st_stream(d4 + j + 64, u8x4_to_f32(a2, unit));
                    st_stream(d4 + j + 96, u8x4_to_f32(a3, unit));
                }
                for (; j < n4; j += 32) st_stream(d4 + j, u8x4_to_f32(ld_stream32(w32 + j), unit));
                continue;
            }
            const uint4 *sp = (const uint4 *)(col.src + ref * col.fb) + u0;
            uint4 *d = (uint4 *)(col.batch + (out_row * col.K + k) * col.fb) + u0;
            const int n = u1 - u0;
            int j = lane;
            for (; j + 96 < n; j += 128) {
                const uint4 v0 = ld_stream(sp + j), v1 = ld_stream(sp + j + 32);
                const uint4 v2 = ld_stream(sp + j + 64), v3 = ld_stream(sp + j + 96);
                st_stream(d + j, v0); st_stream(d + j + 32, v1);
                st_stream(d + j + 64, v2); st_stream(d + j + 96, v3);
            }
            uint4 v[3];
            int cnt = 0;
#pragma unroll
            for (int q = 0; q < 3; ++q)
                if (j + q * 32 < n) { v[q] = ld_stream(sp + j + q * 32); cnt = q + 1; }
#pragma unroll
            for (int q = 0; q < 3; ++q)
                if (q < cnt) st_stream(d + j + q * 32, v[q]);
        }
    }
}

// indices / weights of a shard's slice of a global batch -> the same rows of a peer's batch slot
__global__ void slice_copy_kernel(const int64_t *idx, const double *w, const int64_t *meta,
                                  int64_t *dst_idx, double *dst_w, int64_t *dst_meta = nullptr)
{
    const int64_t first = meta[0], count = meta[1];
    if (dst_meta && blockIdx.x == 0 && threadIdx.x == 0) { dst_meta[0] = first; dst_meta[1] = count; }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        dst_idx[first + i] = idx[i];
        dst_w[first + i] = w[i];
    }
}


// ------------------------------------------------ gather, TMA bulk-copy variant --
// The sampled-row gather as pure DMA: a persistent grid (one CTA per SM), each CTA
// owning a ring of row-sized shared-memory stages.  One thread per CTA issues
//   cp.async.bulk  global -> shared   (completion on an mbarrier, complete_tx bytes)
//   cp.async.bulk  shared -> global   (completion through bulk async-groups)
// so up to kStages x chunk bytes are in flight per SM with ~no instruction issue and
// no register staging; the other warps move the narrow scalar columns meanwhile.
// Items are (row, wide column, chunk) triples, grid-strided; chunk = the whole column
// row when it fits a stage (28 224 B for an Atari frame stack), else equal 16 B-multiple
// pieces.  Replaces the per-row Python loop of client/numpy_loader.py:35-50 and
// client/torch_cuda_loader.py:49-61.
constexpr int kBulkThreads = 128;
constexpr int kBulkSlotCache = 512;      // items whose ring slots are staged in shared memory per round

struct BulkArgs {
    RowArgs r;                  // same decomposition as row_move_kernel (chunks/units per column)
    int32_t stage_bytes;        // bytes of one shared-memory stage (>= the largest chunk)
    int32_t stages;             // ring depth
    int32_t evict_first_loads;  // L2 policy for the ring reads (rows are touched once)
    int32_t evict_first_stores;
};

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}"
        :: "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_first()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void *src, uint32_t bytes, uint32_t bar,
                                         uint64_t policy, bool hint)
{
    if (hint)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
                     "[%0], [%1], %2, [%3], %4;"
                     :: "r"(dst_smem), "l"(src), "r"(bytes), "r"(bar), "l"(policy) : "memory");
    else
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(dst_smem), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst, uint32_t src_smem, uint32_t bytes, uint64_t policy, bool hint)
{
    if (hint)
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
                     :: "l"(dst), "r"(src_smem), "r"(bytes), "l"(policy) : "memory");
    else
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                     :: "l"(dst), "r"(src_smem), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

struct BulkItem {
    const uint8_t *src;
    uint8_t *dst;
    uint32_t bytes;             // 0 = skip (row beyond meta count, bad slot)
};

// item -> (source in the ring, destination in the batch, bytes); slot comes from the shared cache
__device__ __forceinline__ BulkItem bulk_item(const RowArgs &a, int64_t item, int64_t slot, int rows)
{
    int c = 0;
#pragma unroll 1
    for (int k = 0; k < a.ncols; ++k)
        if (a.col[k].chunks > 0 && item >= a.col[k].item_base) c = k;
    const RowCol &col = a.col[c];
    const int64_t rel = item - col.item_base;
    const int row = (int)(rel / col.chunks), chunk = (int)(rel - (int64_t)row * col.chunks);
    BulkItem it;
    it.bytes = 0;
    if (row >= rows || slot < 0 || slot >= a.capacity) return it;
    const int64_t off = (int64_t)chunk * col.units * 16;
    const int64_t left = col.rb - off;
    it.bytes = (uint32_t)(left < (int64_t)col.units * 16 ? left : (int64_t)col.units * 16);
    it.src = col.ring + slot * col.rb + off;
    it.dst = col.batch + (int64_t)row * col.rb + off;
    return it;
}

__device__ __forceinline__ int bulk_item_row(const RowArgs &a, int64_t item)
{
    int c = 0;
#pragma unroll 1
    for (int k = 0; k < a.ncols; ++k)
        if (a.col[k].chunks > 0 && item >= a.col[k].item_base) c = k;
    return (int)((item - a.col[c].item_base) / a.col[c].chunks);
}

__global__ void __launch_bounds__(kBulkThreads, 1) row_gather_bulk_kernel(const BulkArgs b)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const RowArgs &a = b.r;
    const int S = b.stages;
    uint8_t *stage0 = smem;
    uint64_t *bars = (uint64_t *)(smem + (size_t)S * b.stage_bytes);
    int64_t *slot_cache = (int64_t *)(bars + 16);
    const int rows = a.meta ? (int)a.meta[1] : a.batch;
    const int64_t total = a.wide_items;
    // this CTA's items: blockIdx.x, blockIdx.x + gridDim.x, ...
    const int64_t n_local = total > blockIdx.x ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) mbar_init(smem_addr(&bars[s]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    uint64_t pol = 0;
    if (b.evict_first_loads || b.evict_first_stores) pol = l2_policy_evict_first();

    int64_t issued = 0, stored = 0;      // running item counters of the issuing thread (persist across rounds)
    for (int64_t base = 0; base < n_local; base += kBulkSlotCache) {
        const int64_t cnt = (n_local - base < kBulkSlotCache) ? n_local - base : kBulkSlotCache;
        __syncthreads();                 // previous round's cache fully consumed
        for (int64_t j = threadIdx.x; j < cnt; j += kBulkThreads) {
            const int64_t item = blockIdx.x + (base + j) * gridDim.x;
            const int row = bulk_item_row(a, item);
            int64_t slot = -1;
            if (row < rows) {
                slot = a.idx[row];
                if (slot < 0 || slot >= a.capacity) atomicOr(a.err, 1);
            }
            slot_cache[j] = slot;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            // software pipeline over this round: S-1 loads in flight while one stage drains; stage = item % S
            const int64_t end = base + cnt;
            for (;;) {
                // refill: issue loads while a stage is free (a stage is free once the store that read it
                // has finished READING shared memory: all but the newest bulk group)
                while (issued < end && issued - stored < S - 1) {
                    if (issued >= S) {
                        // stage of item `issued` was last used by item issued-S, whose store was committed
                        // at position issued-S; allow only groups newer than that to be pending
                        const int64_t newer = stored - 1 - (issued - S);   // stores committed after it
                        if (newer <= 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                        else if (newer == 1) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                        else if (newer == 2) asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
                        else if (newer == 3) asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory");
                        // newer >= 4: with S <= 8 the store is old enough only if we also wait; be safe
                        else asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory");
                    }
                    const int s = (int)(issued % S);
                    const BulkItem it = bulk_item(a, blockIdx.x + issued * gridDim.x, slot_cache[issued - base], rows);
                    const uint32_t bar = smem_addr(&bars[s]);
                    mbar_expect_tx(bar, it.bytes);           // bytes == 0: the phase completes at once
                    if (it.bytes)
                        bulk_g2s(smem_addr(stage0 + (size_t)s * b.stage_bytes), it.src, it.bytes, bar, pol,
                                 b.evict_first_loads != 0);
                    ++issued;
                }
                if (stored >= end) break;
                // drain one: wait for the oldest load, push it out
                const int s = (int)(stored % S);
                mbar_wait(smem_addr(&bars[s]), (uint32_t)((stored / S) & 1));
                const BulkItem it = bulk_item(a, blockIdx.x + stored * gridDim.x, slot_cache[stored - base], rows);
                if (it.bytes)
                    bulk_s2g(it.dst, smem_addr(stage0 + (size_t)s * b.stage_bytes), it.bytes, pol,
                             b.evict_first_stores != 0);
                else
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");   // keep group count == item count
                ++stored;
            }
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");

    // narrow / unaligned columns by warps 1..3 (one thread per row), overlapping the DMA above
    if (a.narrow_groups > 0 && threadIdx.x >= 32) {
        const int per = kBulkThreads - 32;
        for (int row = blockIdx.x * per + (threadIdx.x - 32); row < rows; row += gridDim.x * per) {
            const int64_t slot = a.idx[row];
            if (slot < 0 || slot >= a.capacity) { atomicOr(a.err, 1); continue; }
            for (int c = 0; c < a.ncols; ++c) {
                const RowCol &col = a.col[c];
                if (col.chunks != 0) continue;
                span_copy(col.batch + (int64_t)row * col.rb, col.ring + slot * col.rb, col.rb, 0, 1);
            }
        }
    }
}

}  // namespace reth
