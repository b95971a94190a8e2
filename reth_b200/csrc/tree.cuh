// tree.cuh -- warp-tiled float64 sum/min tree for prioritized replay (sm_100a).
//
// Semantics are those of the reference's implicit-heap tree
// (reth_buffer/reth_buffer/utils/sumtree.py): node i owns val[i] and has
// children 2i+1, 2i+2; sum[i] = (val[i] + sum[l]) + sum[r]; min[i] per
// _numba_maintain_node (:5-21).  The LAYOUT is new: the heap is re-blocked into
// tiles of 5 binary levels (31 nodes, 32 slots) so one warp reads a tile with a
// single coalesced load and walks its 5 levels with register shuffles:
//
//   desc[tile*32 + local] = (sumL, val)   sumL = sum of the LEFT child subtree
//   agg [tile*64 + local] = (sum, min)    local in 1..31: the tile's own nodes
//   agg [tile*64 + 32 + c] = (sum, min)   copy of the root of child tile c (0..31)
//
// `local` is the 1-based heap position inside the tile; tile-level T holds heap
// depths 5T..5T+4; tiles of one level are contiguous, level T starts at
// (32^T-1)/31.  A descent therefore needs ceil(depth/5) dependent memory round
// trips (5 for capacity 2^24) instead of `depth`, and performs exactly the
// reference's comparisons and subtractions in the reference's order, so sampled
// indices are bit-identical for identical leaves and uniforms.  Slots of nodes
// >= capacity stay zero forever, which makes the reference's `left < capacity`
// guards arithmetic no-ops (x + 0.0, min skipped when 0).
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace reth {
namespace cg = cooperative_groups;

constexpr int kMaxTileLevels = 7;   // capacity <= 2^31 -> heap depth <= 31 -> tile levels 0..6
constexpr int kUpdateThreads = 512;

struct TreeGeom {
    int64_t capacity;
    int64_t n_tiles;
    int32_t n_tile_levels;
};

__host__ __device__ __forceinline__ int64_t tile_base(int T)
{
    return (int64_t)(((1ull << (5 * T)) - 1ull) / 31ull);
}

inline TreeGeom make_geom(int64_t capacity)
{
    TreeGeom g;
    g.capacity = capacity;
    int d = 0;
    while ((2ull << d) <= (uint64_t)capacity) ++d;  // d = floor(log2(capacity)) = depth of node k=capacity
    const int T = d / 5;
    const uint64_t lo = 1ull << (5 * T), hi = (2ull << (5 * T)) - 1ull;
    const uint64_t last = (uint64_t)capacity < hi ? (uint64_t)capacity : hi;
    g.n_tile_levels = T + 1;
    g.n_tiles = tile_base(T) + (int64_t)(last - lo + 1);
    return g;
}

struct NodeLoc {
    int64_t tile;
    int32_t local, T, r;
};

__device__ __forceinline__ NodeLoc locate(int64_t idx)
{
    const uint64_t k = (uint64_t)idx + 1ull;
    const int d = 63 - __clzll((long long)k);
    NodeLoc n;
    n.T = d / 5;
    n.r = d - 5 * n.T;
    n.tile = tile_base(n.T) + (int64_t)((k >> n.r) - (1ull << (5 * n.T)));
    n.local = (int)((1u << n.r) | (uint32_t)(k & ((1ull << n.r) - 1ull)));
    return n;
}

__device__ __forceinline__ double2 ldcg2(const double2 *p) { return __ldcg(p); }

// counter-based uniform in [0,1) with 53 random bits (splitmix64 finaliser);
// the reference draws from numba's private generator (sumtree.py:75), so only
// the distribution, not the stream, is part of the contract.
__device__ __forceinline__ double uniform53(uint64_t seed, uint64_t ctr)
{
    uint64_t z = seed + (ctr + 1ull) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}

// ---------------------------------------------------------------- sampling --
struct SampleArgs {
    TreeGeom g;
    const double2 *desc;
    const double2 *agg;
    int32_t batch;             // stratified segments (global batch in global mode)
    const double *uniforms;    // device double[batch] or nullptr
    uint64_t seed, draw_base;
    double beta;
    int32_t min_mode;
    int64_t *idx_out;
    double *w_out, *prio_out;
    int64_t *meta;             // {first, count}
    const double *all_roots;   // global mode: double[2*n_shards]
    int32_t n_shards, shard;
    int32_t groups;            // independent stratified draws of `batch` in one launch (>= 1)
    int32_t out_stride;        // elements between the outputs of consecutive draws
};

__device__ __forceinline__ double draw_u(const SampleArgs &a, int64_t i)
{
    return a.uniforms ? a.uniforms[i] : uniform53(a.seed, a.draw_base + (uint64_t)i);
}

// smallest i in [0,batch] with ((i + u_i) * segment) >= bound (targets are
// monotone in i); evaluated with the exact target formula over a 32-wide window.
__device__ __forceinline__ int64_t lower_bound_target(const SampleArgs &a, double segment,
                                                      double bound, int lane)
{
    if (!(bound > 0.0)) return 0;
    double q = bound / segment;
    if (!(q < (double)a.batch + 8.0)) q = (double)a.batch + 8.0;
    const int64_t i0 = (int64_t)q - 15;
    const int64_t j = i0 + lane;
    bool ge = true;
    if (j < 0) ge = false;
    else if (j < a.batch) ge = (((double)j + draw_u(a, j)) * segment >= bound);
    const unsigned b = __ballot_sync(0xffffffffu, ge);
    int64_t res = b ? i0 + (__ffs(b) - 1) : i0 + 32;
    if (res < 0) res = 0;
    if (res > a.batch) res = a.batch;
    return res;
}

// One warp per stratified sample (utils/sumtree.py:70-79 + :34-58), IS weight
// fused (sampler/per_sampler.py:26).
__global__ void __launch_bounds__(256) per_sample_kernel(const SampleArgs a)
{
    const int lane = threadIdx.x & 31;
    const int64_t W = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (W >= (int64_t)a.batch * a.groups) return;
    const int64_t grp = W / a.batch, i = W - grp * a.batch;   // draw `grp`, stratum i
    const int64_t cap = a.g.capacity;

    double total, pmin, prefix = 0.0;
    int64_t first = 0, end = a.batch;
    if (a.n_shards > 1) {
        // global CDF over shard totals in rank order (Mode G)
        total = 0.0;
        double mn = 0.0, mine = 0.0;
        for (int s = 0; s < a.n_shards; ++s) {
            const double t = a.all_roots[2 * s], m = a.all_roots[2 * s + 1];
            if (s == a.shard) { prefix = total; mine = t; }
            total += t;
            if (t != 0.0 && m != 0.0) mn = (mn == 0.0 || m < mn) ? m : mn;
        }
        pmin = (mn != 0.0) ? mn : 1.0;
        const double seg = total / (double)a.batch;
        first = lower_bound_target(a, seg, prefix, lane);
        end = (a.shard == a.n_shards - 1) ? (int64_t)a.batch
                                          : lower_bound_target(a, seg, prefix + mine, lane);
        if (end < first) end = first;
        if (i == 0 && lane == 0) { a.meta[0] = first; a.meta[1] = end - first; }
        if (i < first || i >= end) return;
    } else {
        const double2 root = ldcg2(&a.agg[1]);
        total = root.x;
        if (a.min_mode == 0) pmin = (root.x != 0.0) ? root.y : 1.0;   // utils/sumtree.py:109-110
        else pmin = (root.y != 0.0) ? root.y : 1.0;                   // reth/buffer/sumtree.py:112-113
        if (i == 0 && lane == 0) { a.meta[2 * grp] = 0; a.meta[2 * grp + 1] = a.batch; }
    }
    const double segment = total / (double)a.batch;                  // sumtree.py:73
    double w = ((double)i + draw_u(a, W)) * segment;                 // sumtree.py:74
    w -= prefix;                                                      // 0.0 for a single shard

    int64_t idx = 0, tile = 0, tile_rel = 0;
    int T = 0;
    double val = 0.0;
    for (;;) {
        const double2 d = a.desc[tile * 32 + lane];
        int c = 1;
        bool found = false;
#pragma unroll
        for (int l = 0; l < 5; ++l) {
            const double sl = __shfl_sync(0xffffffffu, d.x, c);
            const double v = __shfl_sync(0xffffffffu, d.y, c);
            const int64_t left = 2 * idx + 1;
            if (left < cap) {                       // sumtree.py:42-47
                if (w < sl) { idx = left; c = 2 * c; continue; }
                w -= sl;
            }
            if (w < v + 1e-5) { val = v; found = true; break; }   // :49-50
            w -= v;                                               // :52
            if (left + 1 >= cap) { val = v; found = true; break; }  // :54-55
            idx = left + 1;
            c = 2 * c + 1;
        }
        if (found) break;
        tile_rel = tile_rel * 32 + (c - 32);
        ++T;
        tile = tile_base(T) + tile_rel;
    }
    if (lane == 0) {
        const int64_t o = grp * a.out_stride + i - first;
        a.idx_out[o] = idx;
        a.prio_out[o] = val;
        a.w_out[o] = pow(val / pmin, -a.beta);      // per_sampler.py:26
    }
}

// ------------------------------------------------------------------ update --
struct UpdateArgs {
    TreeGeom g;
    double2 *desc;
    double2 *agg;
    uint32_t *stamp;      // [capacity]  last-writer-wins arbitration
    uint32_t *mask;       // [n_tiles]   touched-node bits of the running update
    int32_t *worklist;    // [kMaxTileLevels][max_n] owned tiles per tile level
    int32_t *counts;      // [kMaxTileLevels] this call
    int32_t *counts_next; // [kMaxTileLevels] zeroed for the next call
    int32_t *err;
    int64_t n, max_n;
    const void *idx;
    int32_t idx_kind;     // 0 int64[], 1 int32[], 2 iota: (iota_base + i) % capacity
    int64_t iota_base;
    const void *prio;
    int32_t prio_kind;    // 0 float32[], 1 float64[], 2 constant 1.0 (raw)
    double alpha;
    int32_t transform;
    int32_t keep_stamps;  // leave the winners' stamps in place (policy-slot scatter reads them)
    int32_t dedupe;       // 0 none (distinct slots), 1 stamps (arbitrary order), 2 adjacent runs
    long long *dbg;       // optional: SM-clock stamps of stage boundaries (CTA 0, thread 0)
};

#define RETH_STAMP(k) do { if (a.dbg && brank == 0 && threadIdx.x == 0) a.dbg[k] = clock64(); } while (0)

__device__ __forceinline__ int64_t upd_index(const UpdateArgs &a, int64_t i)
{
    if (a.idx_kind == 0) return ((const int64_t *)a.idx)[i];
    if (a.idx_kind == 1) return (int64_t)((const int32_t *)a.idx)[i];
    return (a.iota_base + i) % a.g.capacity;
}

// (w + 1e-6) ** alpha in the dtype of the incoming array (per_sampler.py:16-17):
// float32 in -> float32 add and a correctly rounded float32 power.
__device__ __forceinline__ double upd_leaf(const UpdateArgs &a, int64_t i)
{
    if (a.prio_kind == 2) return 1.0;                    // prioritized_buffer.py:36-37
    const bool f32 = a.prio_kind == 0;
    double base, expo = a.alpha;
    if (f32) {
        const float w = ((const float *)a.prio)[i];
        if (!a.transform) return (double)w;
        base = (double)(w + 1e-6f);
        expo = (double)(float)a.alpha;
    } else {
        const double w = ((const double *)a.prio)[i];
        if (!a.transform) return w;
        base = w + 1e-6;
    }
    const double r = pow(base, expo);                    // one call site keeps the kernel small
    return f32 ? (double)(float)r : r;
}

__device__ __forceinline__ double2 maintain_node(double v, const double2 &Lc, const double2 &Rc)
{
    // _numba_maintain_node, utils/sumtree.py:5-21 ((val + sum_l) + sum_r; min skips zeros)
    double s = v + Lc.x;
    s = s + Rc.x;
    double mn = (v != 0.0) ? v : 1.0;
    if (Lc.y != 0.0) mn = (Lc.y < mn) ? Lc.y : mn;
    if (Rc.y != 0.0) mn = (Rc.y < mn) ? Rc.y : mn;
    return make_double2(s, mn);
}

// Warp-cooperative repair of one tile with an arbitrary touched mask (several paths):
// lane = node, 5 levels bottom-up through register shuffles.
__device__ __forceinline__ void repair_tile_warp(const UpdateArgs &a, int64_t tile, int T, int lane)
{
    double2 ag = ldcg2(&a.agg[tile * 64 + lane]);
    const double2 ex = ldcg2(&a.agg[tile * 64 + 32 + lane]);
    double2 de = ldcg2(&a.desc[tile * 32 + lane]);
    const uint32_t m = __ldcg(&a.mask[tile]);
    const bool touched = (m >> lane) & 1u;
#pragma unroll
    for (int l = 4; l >= 0; --l) {
        const int src = ((l == 4) ? 2 * (lane - 16) : 2 * lane) & 31;
        const double sx = (l == 4) ? ex.x : ag.x, sy = (l == 4) ? ex.y : ag.y;
        double2 Lc, Rc;
        Lc.x = __shfl_sync(0xffffffffu, sx, src);
        Lc.y = __shfl_sync(0xffffffffu, sy, src);
        Rc.x = __shfl_sync(0xffffffffu, sx, src | 1);
        Rc.y = __shfl_sync(0xffffffffu, sy, src | 1);
        if (touched && (lane >> l) == 1) {
            ag = maintain_node(de.y, Lc, Rc);
            de.x = Lc.x;
        }
    }
    if (touched) {
        __stcg(&a.agg[tile * 64 + lane], ag);
        __stcg(&a.desc[tile * 32 + lane].x, de.x);
    }
    if (lane == 1 && T > 0) {
        const int64_t rel = tile - tile_base(T);
        __stcg(&a.agg[(tile_base(T - 1) + (rel >> 5)) * 64 + 32 + (rel & 31)], ag);
    }
    if (lane == 0) __stcg(&a.mask[tile], 0u);
}

// One THREAD repairs one tile whose touched set is a single root-ward chain (the
// common case below the top levels): every load address follows from the deepest
// touched node, so all (<= 11) loads are issued together and the <= 5 node
// recomputations run in registers.  32 tiles per warp instead of one.
__device__ __forceinline__ void repair_tile_chain(const UpdateArgs &a, int64_t tile, uint32_t m, int T)
{
    const int j = 31 - __clz(m);      // deepest touched node (1-based position in the tile)
    const int r = 31 - __clz(j);      // its depth inside the tile, 0..4
    const double2 *ag = a.agg + tile * 64;
    double2 *dsc = a.desc + tile * 32;
    double v[5];
    double2 sib[5];
#pragma unroll
    for (int l = 0; l < 5; ++l) {
        if (l > r) continue;
        const int p = j >> (r - l);
        v[l] = __ldcg(&dsc[p].y);
        if (l > 0) sib[l] = ldcg2(&ag[p ^ 1]);
    }
    const int c = (r == 4) ? 32 + 2 * (j - 16) : 2 * j;   // children of the deepest node
    const double2 c0 = ldcg2(&ag[c]), c1 = ldcg2(&ag[c + 1]);
    double2 cur = make_double2(0.0, 0.0);
#pragma unroll
    for (int l = 4; l >= 0; --l) {
        if (l > r) continue;
        const int p = j >> (r - l);
        double2 Lc, Rc;
        if (l == r) { Lc = c0; Rc = c1; }
        else {
            const bool left = (((j >> (r - l - 1)) & 1) == 0);   // is the path child the left child?
            Lc = left ? cur : sib[l + 1];
            Rc = left ? sib[l + 1] : cur;
        }
        cur = maintain_node(v[l], Lc, Rc);
        __stcg(&a.agg[tile * 64 + p], cur);
        __stcg(&dsc[p].x, Lc.x);
    }
    if (T > 0) {
        const int64_t rel = tile - tile_base(T);
        __stcg(&a.agg[(tile_base(T - 1) + (rel >> 5)) * 64 + 32 + (rel & 31)], cur);
    }
    __stcg(&a.mask[tile], 0u);
}

// Tile levels 0 and 1 (heap depths 0..9, <= 1023 nodes) swept by one CTA in shared
// memory, one __syncthreads per tree level.  A node is recomputed when its own mask
// bit is set (it was written by this batch) or a child has ever been maintained
// (child min != 0, which _numba_maintain_node guarantees for every node it visited):
// its ancestors were maintained then too, so recomputing them is idempotent.  Deep
// paths therefore need no marking atomics on the two top tile levels.
__device__ __forceinline__ void repair_top(const UpdateArgs &a)
{
    __shared__ double sS[2048], sM[2048], sV[1024];
    __shared__ uint8_t sK[1024];
    const int NT = a.g.n_tile_levels;
    const int D = (NT >= 2) ? 10 : 5;                    // tree levels in the sweep
    const int nn = 1 << D;                                // nodes are heap k = 1 .. nn-1
    const int64_t n1 = (NT >= 2) ? ((a.g.n_tiles - 1 < 32) ? a.g.n_tiles - 1 : 32) : 0;
    // 2*nn <= 2048 entries over 512 threads: 4 per thread, all loads issued before any use
    double2 g[4];
    double v[4];
    uint32_t mk[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int k = threadIdx.x + q * kUpdateThreads;
        g[q] = make_double2(0.0, 0.0);
        v[q] = 0.0;
        mk[q] = 0u;
        if (k >= 2 * nn) continue;
        if (k >= 1 && k < nn) {
            const int d = 31 - __clz(k);
            const int64_t tile = (d < 5) ? 0 : 1 + ((k >> (d - 5)) - 32);
            const int local = (d < 5) ? k : ((1 << (d - 5)) | (k & ((1 << (d - 5)) - 1)));
            if (tile <= n1) {
                g[q] = ldcg2(&a.agg[tile * 64 + local]);
                v[q] = __ldcg(&a.desc[tile * 32 + local].y);
                mk[q] = (__ldcg(&a.mask[tile]) >> local) & 1u;
            }
        } else if (k >= nn) {                             // children of the last swept level
            const int64_t tile = (NT >= 2) ? 1 + ((k >> 5) - 32) : 0;
            if (tile <= n1) g[q] = ldcg2(&a.agg[tile * 64 + 32 + (k & 31)]);
        }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int k = threadIdx.x + q * kUpdateThreads;
        if (k >= 2 * nn) continue;
        sS[k] = g[q].x;
        sM[k] = g[q].y;
        if (k < nn) { sV[k] = v[q]; sK[k] = (uint8_t)mk[q]; }
    }
    if (a.dbg && threadIdx.x == 0) a.dbg[23] = clock64();
    __syncthreads();
    for (int l = D - 1; l >= 0; --l) {
        for (int k = (1 << l) + threadIdx.x; k < (2 << l); k += kUpdateThreads) {
            const double2 Lc = make_double2(sS[2 * k], sM[2 * k]), Rc = make_double2(sS[2 * k + 1], sM[2 * k + 1]);
            if (sK[k] || Lc.y != 0.0 || Rc.y != 0.0) {
                const double2 res = maintain_node(sV[k], Lc, Rc);
                sS[k] = res.x;
                sM[k] = res.y;
                sK[k] = 2;
            }
        }
        __syncthreads();
    }
    if (a.dbg && threadIdx.x == 0) a.dbg[24] = clock64();
    for (int k = 1 + threadIdx.x; k < nn; k += kUpdateThreads) {
        if (sK[k] != 2) continue;
        const int d = 31 - __clz(k);
        const int64_t tile = (d < 5) ? 0 : 1 + ((k >> (d - 5)) - 32);
        const int local = (d < 5) ? k : ((1 << (d - 5)) | (k & ((1 << (d - 5)) - 1)));
        const double2 res = make_double2(sS[k], sM[k]);
        __stcg(&a.agg[tile * 64 + local], res);
        __stcg(&a.desc[tile * 32 + local].x, sS[2 * k]);
        if (d == 5) __stcg(&a.agg[32 + (k - 32)], res);  // root of a level-1 tile: copy in tile 0
    }
    if (threadIdx.x <= n1) __stcg(&a.mask[threadIdx.x], 0u);
}

// Batched priority write-back with tree-path repair (utils/sumtree.py:61-67,
// 24-31).  One thread-block cluster; stages separated by cluster barriers:
//   P0 stamp  : atomicMax(stamp[idx], 1+i) -> the last i of a duplicate index wins
//   P1 leaves : winners write val.  For tile levels >= 2 they OR their path bits into
//               the per-tile masks (all atomics of a path issued together); the first
//               marker of a tile owns it: worklist[T][i] = tile (else -1).  A leaf that
//               itself lives in tile level 0/1 only sets its own mask bit.
//   R  repair : tile levels deepest -> 2 (skipped when nobody owns a tile there); a
//               lane takes one owned tile: single-chain tiles are repaired by that
//               thread alone, multi-path tiles by the whole warp; cluster barrier
//   TOP       : tile levels 1 and 0 by CTA 0 in shared memory (repair_top)
// The result equals the reference's sequential loop because sum/min are pure
// functions of the final leaves on the union of touched paths.
__global__ void __launch_bounds__(kUpdateThreads) per_update_kernel(const UpdateArgs a)
{
    cg::cluster_group cluster = cg::this_cluster();
    const int nblk = (int)cluster.num_blocks(), brank = (int)cluster.block_rank();
    const int64_t nthr = (int64_t)nblk * kUpdateThreads;
    // consecutive work items are spread over the CTAs of the cluster
    const int64_t L = (int64_t)threadIdx.x * nblk + brank;
    const int64_t cap = a.g.capacity;
    const int lane = threadIdx.x & 31;
    const int NT = a.g.n_tile_levels;

    RETH_STAMP(0);
    if (brank == 0 && threadIdx.x < kMaxTileLevels) a.counts_next[threadIdx.x] = 0;
    if (a.dedupe == 1) {
        for (int64_t i = L; i < a.n; i += nthr) {
            const int64_t idx = upd_index(a, i);
            if (idx < 0 || idx >= cap) { atomicOr(a.err, 1); continue; }
            atomicMax(&a.stamp[idx], 1u + (uint32_t)i);
        }
        RETH_STAMP(1);
        cluster.sync();
    }
    RETH_STAMP(2);
    const int64_t iters = (a.n + nthr - 1) / nthr;      // warp-uniform trip count (ballots inside)
    for (int64_t it = 0; it < iters; ++it) {
        const int64_t i = it * nthr + L;
        const bool inrange = i < a.n;
        bool active = inrange;
        int64_t idx = 0;
        if (active) {
            idx = upd_index(a, i);
            if (idx < 0 || idx >= cap) { atomicOr(a.err, 1); active = false; }
        }
        uint32_t st = 0;
        if (active && a.dedupe == 1) st = __ldcg(&a.stamp[idx]);
        // dedupe 2: indices drawn by per_sample_kernel.  Stratified targets are monotone and
        // _numba_find_index is a monotone step function of its target, so equal indices are
        // always adjacent: the last element of a run is the sequential loop's last writer.
        if (active && a.dedupe == 2 && i + 1 < a.n && upd_index(a, i + 1) == idx) active = false;
        if (it == 0) RETH_STAMP(11);
        const double leaf = active ? upd_leaf(a, i) : 0.0;       // pow overlaps the stamp load
        if (it == 0) RETH_STAMP(12);
        if (active && a.dedupe == 1) {
            if (st != 1u + (uint32_t)i) active = false;
            else if (!a.keep_stamps) __stcg(&a.stamp[idx], 0u);   // stamps are zero between launches
        }
        int64_t tiles[kMaxTileLevels];
        uint32_t old[kMaxTileLevels];
        int Tn = -1;
        if (active) {
            const NodeLoc n = locate(idx);
            Tn = n.T;
            __stcg(&a.desc[n.tile * 32 + n.local].y, leaf);
            if (Tn <= 1) {
                atomicOr(&a.mask[n.tile], 1u << n.local);         // own bit; ancestors are derived
            } else {
                uint32_t bits[kMaxTileLevels];
                int64_t tile = n.tile;
                int local = n.local, r = n.r;
                uint64_t kroot = ((uint64_t)idx + 1ull) >> n.r;   // heap k of the tile root
#pragma unroll
                for (int T = kMaxTileLevels - 1; T >= 2; --T) {
                    if (T > Tn) continue;
                    uint32_t b = 0;
                    for (int j = 0; j <= r; ++j) b |= 1u << (local >> (r - j));
                    tiles[T] = tile;
                    bits[T] = b;
                    // parent of the tile root is a bottom-row node of the parent tile
                    const uint64_t kp = kroot >> 1;
                    local = 16 | (int)(kp & 15ull);
                    r = 4;
                    kroot = kp >> 4;
                    tile = tile_base(T - 1) + (int64_t)(kroot - (1ull << (5 * (T - 1))));
                }
#pragma unroll
                for (int T = kMaxTileLevels - 1; T >= 2; --T)
                    if (T <= Tn) old[T] = atomicOr(&a.mask[tiles[T]], bits[T]);
            }
        }
        if (it == 0) RETH_STAMP(13);
#pragma unroll
        for (int T = kMaxTileLevels - 1; T >= 2; --T) {
            if (T >= NT) continue;                                // uniform
            const bool own = active && T <= Tn && old[T] == 0u;
            if (inrange) a.worklist[(int64_t)T * a.max_n + i] = own ? (int32_t)tiles[T] : -1;
            const unsigned bal = __ballot_sync(0xffffffffu, own);
            if (bal != 0u && lane == __ffs(bal) - 1) atomicOr(&a.counts[0], 1 << T);   // level has work
        }
    }
    RETH_STAMP(3);
    cluster.sync();
    RETH_STAMP(4);

    const int64_t warp = (int64_t)(threadIdx.x >> 5) * nblk + brank, nwarps = nthr >> 5;
    // entry e of a level's worklist goes to lane (e / nwarps) of warp (e % nwarps): the owned
    // tiles are spread thinly so that multi-path tiles rarely queue up inside one warp
    const int64_t e0 = (int64_t)lane * nwarps + warp;
    const int flags = __ldcg(&a.counts[0]);                       // bit T: somebody owns a tile at level T
#pragma unroll 1
    for (int T = NT - 1; T >= 2; --T) {
        if (!((flags >> T) & 1)) continue;                        // uniform over the cluster
        for (int64_t base = 0; base < a.n; base += 32 * nwarps) {
            const int64_t e = base + e0;
            int32_t t = -1;
            if (e < a.n) t = __ldcg(&a.worklist[(int64_t)T * a.max_n + e]);
            uint32_t m = 0;
            if (t >= 0) m = __ldcg(&a.mask[t]);
            if (m) RETH_STAMP(14 + (T & 1));
            const bool chain = t >= 0 && __popc(m) == (31 - __clz(31 - __clz(m | 1u))) + 1 && m != 0u;
            if (chain) repair_tile_chain(a, (int64_t)t, m, T);
            RETH_STAMP(16 + (T & 1));
            unsigned rest = __ballot_sync(0xffffffffu, t >= 0 && !chain);
            while (rest) {
                const int src = __ffs(rest) - 1;
                rest &= rest - 1;
                repair_tile_warp(a, (int64_t)__shfl_sync(0xffffffffu, t, src), T, lane);
            }
        }
        RETH_STAMP(5 + 2 * (T - 2));
        cluster.sync();
        RETH_STAMP(6 + 2 * (T - 2));
    }
    if (brank != 0) return;
    RETH_STAMP(20);
    repair_top(a);
    RETH_STAMP(22);
}

// Zero the arbitration stamps of `n` slots (after a policy-slot scatter that needed them).
__global__ void stamp_clear_kernel(uint32_t *stamp, const int64_t *slots, int64_t n, int64_t cap)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && slots[i] >= 0 && slots[i] < cap) stamp[slots[i]] = 0u;
}

// ---------------------------------------------------------------- readback --
// Tiled -> heap order (the three arrays of NumbaSumTree, utils/sumtree.py:85-88).
__global__ void tree_read_kernel(TreeGeom g, const double2 *desc, const double2 *agg,
                                 double *sum, double *mn, double *val)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.capacity) return;
    const NodeLoc n = locate(i);
    const double2 ag = agg[n.tile * 64 + n.local];
    if (sum) sum[i] = ag.x;
    if (mn) mn[i] = ag.y;
    if (val) val[i] = desc[n.tile * 32 + n.local].y;
}

}  // namespace reth
