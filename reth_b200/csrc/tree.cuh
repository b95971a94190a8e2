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

constexpr int kMaxTileLevels = 8;
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
    const int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= a.batch) return;
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
        if (i == 0 && lane == 0) { a.meta[0] = 0; a.meta[1] = a.batch; }
    }
    const double segment = total / (double)a.batch;                  // sumtree.py:73
    double w = ((double)i + draw_u(a, i)) * segment;                 // sumtree.py:74
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
        const int64_t o = i - first;
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
    uint32_t stamp_base;
    int32_t dedupe;
};

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
    if (a.prio_kind == 0) {
        const float w = ((const float *)a.prio)[i];
        if (!a.transform) return (double)w;
        const float x = w + 1e-6f;
        return (double)(float)pow((double)x, (double)(float)a.alpha);
    }
    const double w = ((const double *)a.prio)[i];
    return a.transform ? pow(w + 1e-6, a.alpha) : w;
}

// Recompute the touched nodes of one tile bottom-up (5 levels in registers).
__device__ __forceinline__ void repair_tile(const UpdateArgs &a, int64_t tile, int T, int lane)
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
        const double Lx = __shfl_sync(0xffffffffu, sx, src);
        const double Ly = __shfl_sync(0xffffffffu, sy, src);
        const double Rx = __shfl_sync(0xffffffffu, sx, src | 1);
        const double Ry = __shfl_sync(0xffffffffu, sy, src | 1);
        if (touched && (lane >> l) == 1) {
            // _numba_maintain_node, utils/sumtree.py:5-21
            double s = de.y + Lx;
            s = s + Rx;
            double mn = (de.y != 0.0) ? de.y : 1.0;
            if (Ly != 0.0) mn = (Ly < mn) ? Ly : mn;
            if (Ry != 0.0) mn = (Ry < mn) ? Ry : mn;
            ag.x = s;
            ag.y = mn;
            de.x = Lx;
        }
    }
    if (touched) {
        __stcg(&a.agg[tile * 64 + lane], ag);
        __stcg(&a.desc[tile * 32 + lane].x, de.x);
    }
    if (lane == 1 && T > 0) {
        const int64_t rel = tile - tile_base(T);
        const int64_t parent = tile_base(T - 1) + (rel >> 5);
        __stcg(&a.agg[parent * 64 + 32 + (rel & 31)], ag);
    }
    if (lane == 0) __stcg(&a.mask[tile], 0u);
}

// Batched priority write-back with tree-path repair (utils/sumtree.py:61-67,
// 24-31).  One thread-block cluster; phases are separated by cluster barriers:
//   P0 stamp   : atomicMax(stamp[idx], base+1+i)  -> the last i of a duplicate wins
//   P1 leaves  : winners write val, mark their path bits per tile; the first marker
//                of a tile owns it and queues it on that tile level's worklist
//   R  repair  : tile levels deepest -> root, one warp per owned tile
// The result equals the reference's sequential loop because sum/min are pure
// functions of the final leaves on the union of touched paths.
__global__ void __launch_bounds__(kUpdateThreads) per_update_kernel(const UpdateArgs a)
{
    cg::cluster_group cluster = cg::this_cluster();
    const int64_t tid = cluster.thread_rank();
    const int64_t nthr = cluster.num_threads();
    const int64_t cap = a.g.capacity;

    if (tid < kMaxTileLevels) a.counts_next[tid] = 0;
    if (a.dedupe) {
        for (int64_t i = tid; i < a.n; i += nthr) {
            const int64_t idx = upd_index(a, i);
            if (idx < 0 || idx >= cap) { atomicOr(a.err, 1); continue; }
            atomicMax(&a.stamp[idx], a.stamp_base + 1u + (uint32_t)i);
        }
        cluster.sync();
    }
    for (int64_t i = tid; i < a.n; i += nthr) {
        const int64_t idx = upd_index(a, i);
        if (idx < 0 || idx >= cap) { atomicOr(a.err, 1); continue; }
        if (a.dedupe && __ldcg(&a.stamp[idx]) != a.stamp_base + 1u + (uint32_t)i) continue;
        const NodeLoc n = locate(idx);
        __stcg(&a.desc[n.tile * 32 + n.local].y, upd_leaf(a, i));
        // own tile, then every ancestor tile (entered at its bottom row)
        int64_t tile = n.tile;
        int local = n.local, r = n.r;
        uint64_t kroot = ((uint64_t)idx + 1ull) >> n.r;   // heap k of the tile root
        for (int T = n.T; T >= 0; --T) {
            uint32_t bits = 0;
            for (int j = 0; j <= r; ++j) bits |= 1u << (local >> (r - j));
            const uint32_t old = atomicOr(&a.mask[tile], bits);
            if (old == 0u) {
                const int s = atomicAdd(&a.counts[T], 1);
                a.worklist[(int64_t)T * a.max_n + s] = (int32_t)tile;
            }
            if (T == 0) break;
            // parent of the tile root is a bottom-row node of the parent tile
            const uint64_t kp = kroot >> 1;               // heap k, depth 5T-1
            local = 16 | (int)(kp & 15ull);
            r = 4;
            kroot = kp >> 4;
            tile = tile_base(T - 1) + (int64_t)(kroot - (1ull << (5 * (T - 1))));
        }
    }
    cluster.sync();

    const int lane = threadIdx.x & 31;
    const int64_t warp = tid >> 5, nwarps = nthr >> 5;
    for (int T = a.g.n_tile_levels - 1; T >= 0; --T) {
        const int cnt = __ldcg(&a.counts[T]);
        for (int64_t wi = warp; wi < cnt; wi += nwarps)
            repair_tile(a, (int64_t)__ldcg(&a.worklist[(int64_t)T * a.max_n + wi]), T, lane);
        if (T > 0) cluster.sync();
    }
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
