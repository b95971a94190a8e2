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

// -------------------------------------------------- row gather / row scatter --
struct RowCol {
    uint8_t *ring;        // ring column base
    uint8_t *batch;       // contiguous batch column base (gather dst / scatter src)
    int64_t rb;           // row bytes
    int32_t chunks;       // work items per row (wide columns), 0 = narrow column
    int32_t units;        // 16 B units per chunk
    int64_t item_base;    // first work item of this column
};

struct RowArgs {
    RowCol col[kMaxCols];
    int32_t ncols;
    const int64_t *idx;        // slot per batch row
    int32_t batch;
    const int64_t *meta;       // optional {first,count}: rows = meta[1]
    int64_t wide_items;        // items over all wide columns (batch-relative layout)
    int32_t narrow_groups;     // ceil(batch / kCopyThreads) when narrow columns exist
    // scatter only: last-writer-wins arbitration (same stamps as the tree update)
    const uint32_t *stamp;
    int64_t capacity;
    int32_t *err;
};

// kScatter=false: batch[i] <- ring[idx[i]]   (numpy_loader.py:35-50, buffer.py:92-93)
// kScatter=true : ring[idx[i]] <- batch[i]   (append to policy-chosen slots)
template <bool kScatter>
__global__ void __launch_bounds__(kCopyThreads, 6) row_move_kernel(const RowArgs a)
{
    const int rows = a.meta ? (int)a.meta[1] : a.batch;
    const int64_t total = a.wide_items + a.narrow_groups;
    for (int64_t item = blockIdx.x; item < total; item += gridDim.x) {
        if (item < a.wide_items) {
            int c = 0;
#pragma unroll 1
            for (int k = 0; k < a.ncols; ++k)
                if (a.col[k].chunks > 0 && item >= a.col[k].item_base) c = k;
            const RowCol &col = a.col[c];
            const int64_t rel = item - col.item_base;
            const int row = (int)(rel / col.chunks), chunk = (int)(rel - (int64_t)row * col.chunks);
            if (row >= rows) continue;
            const int64_t slot = a.idx[row];
            if (slot < 0 || slot >= a.capacity) { if (threadIdx.x == 0) atomicOr(a.err, 1); continue; }
            if (kScatter && a.stamp && a.stamp[slot] != 1u + (uint32_t)row) continue;
            const int64_t upr = col.rb >> 4;
            const int64_t u0 = (int64_t)chunk * col.units;
            const int64_t n = min((int64_t)col.units, upr - u0);
            const uint4 *rp = (const uint4 *)(col.ring + slot * col.rb) + u0;
            const uint4 *bp = (const uint4 *)(col.batch + (int64_t)row * col.rb) + u0;
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
            const int row = (int)(item - a.wide_items) * kCopyThreads + threadIdx.x;
            if (row >= rows) continue;
            const int64_t slot = a.idx[row];
            if (slot < 0 || slot >= a.capacity) { atomicOr(a.err, 1); continue; }
            if (kScatter && a.stamp && a.stamp[slot] != 1u + (uint32_t)row) continue;
            for (int c = 0; c < a.ncols; ++c) {
                const RowCol &col = a.col[c];
                if (col.chunks != 0) continue;
                uint8_t *rp = col.ring + slot * col.rb;
                uint8_t *bp = col.batch + (int64_t)row * col.rb;
                span_copy(kScatter ? rp : bp, kScatter ? bp : rp, col.rb, 0, 1);
            }
        }
    }
}

}  // namespace reth
