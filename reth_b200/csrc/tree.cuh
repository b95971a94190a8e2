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

// Programmatic dependent launch (a launch with cudaLaunchAttributeProgrammaticStreamSerialization): the next
// kernel of the stream may be scheduled while this one still runs -- pdl_trigger() -- and blocks in
// pdl_wait() until the kernel before it has completed and flushed.  Both are no-ops in an ordinary launch.
// Every kernel waits FIRST and triggers after: at most one grid per stream is resident ahead of its turn
// (triggering before the wait lets the whole queue of the stream cascade onto the device).
// Used for the dependent chain draw -> repair -> draw of strict alternation and Mode G, where the launch
// latency of every kernel is on the critical path.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---- shard roots board (Mode G, see roots_exchange_kernel below)
struct BoardEntry {
    double sum, mn;
    unsigned long long seq, pad;
};
constexpr int kMaxShards = 64;
constexpr int kBoardEntries = 2 * kMaxShards;
constexpr int kErrBoardTimeout = 0x100;

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}


// Publish this shard's root {total, min} with exchange number `seq` into every shard's roots board (plain
// stores over NVLink, the flag with release semantics at system scope); lanes stride over the shards.
__device__ __forceinline__ void board_publish(BoardEntry *const *boards, int n_shards, int shard,
                                              unsigned long long seq, const double2 *agg, int lane)
{
    const double2 r = ldcg2(&agg[1]);
    for (int s = lane; s < n_shards; s += 32) {
        BoardEntry *e = boards[s] + (int)(seq & 1ull) * kMaxShards + shard;
        volatile double *vs = &e->sum, *vm = &e->mn;
        *vs = r.x;
        *vm = r.y;
        st_release_sys(&e->seq, seq);
    }
}
// Wait until shard s's entry of exchange `seq` has arrived in the OWN board (local L2 polls); false on timeout.
__device__ __forceinline__ bool board_wait(const BoardEntry *own, int s, unsigned long long seq,
                                           unsigned long long timeout_ns)
{
    const BoardEntry *e = own + (int)(seq & 1ull) * kMaxShards + s;
    if (ld_acquire_sys(&e->seq) == seq) return true;
    const unsigned long long t0 = global_ns();
    while (ld_acquire_sys(&e->seq) != seq)
        if (global_ns() - t0 > timeout_ns) return false;
    return true;
}

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
    const int64_t *quota;      // global mode with groups > 1: int64[4*groups] = {first, count, packed offset, -} per draw
    int32_t groups;            // independent stratified draws of `batch` in one launch (>= 1)
    int32_t warps_per_draw;    // with quota: warps the launch provides per draw (they stride over the share)
    int32_t out_stride;        // elements between the outputs of consecutive draws
    // loader pipeline: indices / weights are also written straight into pinned host memory mapped into
    // the device address space (no device->host copy operation), and the error flag is mirrored
    int64_t *idx_mirror;
    double *w_mirror;
    int32_t *err_mirror;
    const int32_t *err;
    // Mode G with the exchange folded into the chain: the shard roots of exchange `wait_seq` are taken from
    // this shard's roots board (published by the peers' repair kernels) instead of a snapshot made by
    // roots_exchange_kernel; the first kernel of the draw waits for them and writes the snapshot roots_out.
    const BoardEntry *wait_board;
    unsigned long long wait_seq, wait_timeout_ns;
    double *roots_out;
    int32_t *err_rw;
};

// Wait for every shard's root of exchange a.wait_seq on the own board and return shard s's {total, min}
// through `roots_out` (snapshot) -- one call per launch by a single warp / block; false on timeout.
__device__ __forceinline__ bool board_collect(const SampleArgs &a, int tid, int nthr)
{
    bool ok = true;
    for (int s = tid; s < a.n_shards; s += nthr) {
        if (!board_wait(a.wait_board, s, a.wait_seq, a.wait_timeout_ns)) { ok = false; continue; }
        const BoardEntry *e = a.wait_board + (int)(a.wait_seq & 1ull) * kMaxShards + s;
        const volatile double *vs = &e->sum, *vm = &e->mn;
        a.roots_out[2 * s] = *vs;
        a.roots_out[2 * s + 1] = *vm;
    }
    if (!ok) atomicOr(a.err_rw, kErrBoardTimeout);
    return ok;
}

__device__ __forceinline__ double draw_u(const SampleArgs &a, int64_t i)
{
    return a.uniforms ? a.uniforms[i] : uniform53(a.seed, a.draw_base + (uint64_t)i);
}

// smallest i in [0,batch] with ((i + u_i) * segment) >= bound (targets are
// monotone in i); evaluated with the exact target formula over a 32-wide window.
__device__ __forceinline__ int64_t lower_bound_target(const SampleArgs &a, double segment,
                                                      double bound, int lane, int64_t draw0 = 0)
{
    if (!(bound > 0.0)) return 0;
    double q = bound / segment;
    if (!(q < (double)a.batch + 8.0)) q = (double)a.batch + 8.0;
    const int64_t i0 = (int64_t)q - 15;
    const int64_t j = i0 + lane;
    bool ge = true;
    if (j < 0) ge = false;
    else if (j < a.batch) ge = (((double)j + draw_u(a, draw0 + j)) * segment >= bound);
    const unsigned b = __ballot_sync(0xffffffffu, ge);
    int64_t res = b ? i0 + (__ffs(b) - 1) : i0 + 32;
    if (res < 0) res = 0;
    if (res > a.batch) res = a.batch;
    return res;
}

// _numba_find_index (utils/sumtree.py:34-58) on the tiled layout, by one warp: a coalesced 512-byte tile load,
// five levels walked with shuffles, next tile.  The comparisons and subtractions are the reference's, in the
// reference's order.  Returns the leaf's value.
__device__ __forceinline__ double descend(const SampleArgs &a, double w, int lane, int64_t &idx_out)
{
    const int64_t cap = a.g.capacity;
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
    idx_out = idx;
    return val;
}

// One warp per stratified sample (utils/sumtree.py:70-79 + :34-58), IS weight
// fused (sampler/per_sampler.py:26).
__global__ void __launch_bounds__(256) per_sample_kernel(const SampleArgs a)
{
    pdl_wait();
    pdl_trigger();
    const int lane = threadIdx.x & 31;
    const int64_t W = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (a.err_mirror && W == 0 && lane == 0) *a.err_mirror = *a.err;   // set by earlier launches of the stream

    if (a.quota) {
        // Mode G with the shares already known (global_quota_kernel ran just before: {first, count, packed
        // offset} per draw): the launch covers warps_per_draw warps per draw -- about this shard's share, not
        // the global batch -- and warp j of draw g takes positions first_g + j, + warps_per_draw, ... of it.
        const int64_t grp = W / a.warps_per_draw;
        if (grp >= a.groups) return;
        const int64_t first = a.quota[4 * grp], count = a.quota[4 * grp + 1], off = a.quota[4 * grp + 2];
        double total = 0.0, prefix = 0.0, mn = 0.0;
        for (int s = 0; s < a.n_shards; ++s) {          // global CDF over shard totals in rank order
            const double t = a.all_roots[2 * s], m = a.all_roots[2 * s + 1];
            if (s == a.shard) prefix = total;
            total += t;
            if (t != 0.0 && m != 0.0) mn = (mn == 0.0 || m < mn) ? m : mn;
        }
        const double pmin = (mn != 0.0) ? mn : 1.0;
        const double segment = total / (double)a.batch;
        for (int64_t j = W - grp * a.warps_per_draw; j < count; j += a.warps_per_draw) {
            const int64_t i = first + j;
            double w = ((double)i + draw_u(a, grp * a.batch + i)) * segment;
            w -= prefix;
            int64_t idx;
            const double val = descend(a, w, lane, idx);
            if (lane == 0) {
                const int64_t o = off + j;
                const double wt = pow(val / pmin, -a.beta);
                a.idx_out[o] = idx;
                a.prio_out[o] = val;
                a.w_out[o] = wt;
                if (a.idx_mirror) { a.idx_mirror[o] = idx; a.w_mirror[o] = wt; }
            }
        }
        return;
    }

    if (W >= (int64_t)a.batch * a.groups) return;
    const int64_t grp = W / a.batch, i = W - grp * a.batch;   // draw `grp`, stratum i
    double total, pmin, prefix = 0.0;
    int64_t first = 0;
    if (a.n_shards > 1) {
        // global CDF over shard totals in rank order (Mode G), one draw, share not known yet
        total = 0.0;
        double mn = 0.0, mine = 0.0;
        const bool from_board = a.wait_seq != 0ull;
        const BoardEntry *brd = a.wait_board + (int)(a.wait_seq & 1ull) * kMaxShards;
        if (from_board) {
            // every warp waits on the own board (local L2); warp 0 also leaves the snapshot for the host side
            bool ok = true;
            for (int s = lane; s < a.n_shards; s += 32) ok = board_wait(a.wait_board, s, a.wait_seq, a.wait_timeout_ns) && ok;
            if (!__all_sync(0xffffffffu, ok)) { if (lane == 0) atomicOr(a.err_rw, kErrBoardTimeout); return; }
        }
        for (int s = 0; s < a.n_shards; ++s) {
            double t, m;
            if (from_board) {
                const volatile double *vs = &brd[s].sum, *vm = &brd[s].mn;
                t = *vs; m = *vm;
                if (W == 0 && lane == 0) { a.roots_out[2 * s] = t; a.roots_out[2 * s + 1] = m; }
            } else { t = a.all_roots[2 * s]; m = a.all_roots[2 * s + 1]; }
            if (s == a.shard) { prefix = total; mine = t; }
            total += t;
            if (t != 0.0 && m != 0.0) mn = (mn == 0.0 || m < mn) ? m : mn;
        }
        pmin = (mn != 0.0) ? mn : 1.0;
        const double seg = total / (double)a.batch;
        const bool last = a.shard == a.n_shards - 1;
        if (i == 0) {                          // one warp reports the share {first, count}
            const int64_t f = lower_bound_target(a, seg, prefix, lane);
            int64_t e = last ? (int64_t)a.batch : lower_bound_target(a, seg, prefix + mine, lane);
            if (e < f) e = f;
            if (lane == 0) { a.meta[0] = f; a.meta[1] = e - f; }
        }
        // targets are monotone in i, so membership in [first, end) follows from the warp's own target:
        // i >= first  <=>  target_i >= prefix,   i < end  <=>  not (target_i >= prefix + mine).
        // Warps outside the share (all but 1/n_shards of a global batch) leave after one draw.
        const double tgt = ((double)i + draw_u(a, W)) * seg;
        if ((prefix > 0.0 && !(tgt >= prefix)) || (!last && tgt >= prefix + mine)) return;
        first = lower_bound_target(a, seg, prefix, lane);
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
    int64_t idx;
    const double val = descend(a, w, lane, idx);
    if (lane == 0) {
        const int64_t o = grp * a.out_stride + i - first;
        const double wt = pow(val / pmin, -a.beta);  // per_sampler.py:26
        a.idx_out[o] = idx;
        a.prio_out[o] = val;
        a.w_out[o] = wt;
        if (a.idx_mirror) { a.idx_mirror[o] = idx; a.w_mirror[o] = wt; }
    }
}

// Mode G with a prefetch group: warp g computes the share {first, count} of draw g of this shard
// (the range rule of per_sample_kernel); thread 0 then lays the shares out back to back -- the
// slots of the group become ONE packed batch of sum(count) rows -- and writes
//   quota[4g..] = {first_g, count_g, offset_g, 0},  meta = {0, sum(count)}.
__global__ void __launch_bounds__(1024) global_quota_kernel(const SampleArgs a, int64_t *quota_out,
                                                           int64_t *slot_quota /* per slot {first, count, offset, 0} */)
{
    pdl_wait();
    pdl_trigger();
    __shared__ int64_t sh_first[32], sh_count[32];
    const int lane = threadIdx.x & 31, grp = threadIdx.x >> 5;
    if (a.wait_seq != 0ull) {          // exchange folded into the chain: wait for the peers' roots, snapshot them
        board_collect(a, threadIdx.x, blockDim.x);
        __syncthreads();               // the snapshot (a.all_roots == a.roots_out) is read below by this block
    }
    if (grp < a.groups) {
        double total = 0.0, prefix = 0.0, mine = 0.0;
        for (int s = 0; s < a.n_shards; ++s) {
            const double t = a.all_roots[2 * s];
            if (s == a.shard) { prefix = total; mine = t; }
            total += t;
        }
        const double seg = total / (double)a.batch;
        const int64_t d0 = (int64_t)grp * a.batch;
        const int64_t f = lower_bound_target(a, seg, prefix, lane, d0);
        int64_t e = (a.shard == a.n_shards - 1) ? (int64_t)a.batch : lower_bound_target(a, seg, prefix + mine, lane, d0);
        if (e < f) e = f;
        if (lane == 0) { sh_first[grp] = f; sh_count[grp] = e - f; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int64_t off = 0;
        for (int g = 0; g < a.groups; ++g) {
            const int64_t q[4] = {sh_first[g], sh_count[g], off, 0};
            for (int k = 0; k < 4; ++k) { quota_out[4 * g + k] = q[k]; slot_quota[4 * g + k] = q[k]; }
            off += sh_count[g];
        }
        // one draw: the slot is an ordinary Mode G slot {first, count}; several: ONE packed batch of `off` rows
        a.meta[0] = a.groups == 1 ? sh_first[0] : 0;
        a.meta[1] = off;
    }
}

// ------------------------------------------------------------------ update --
struct UpdateArgs {
    TreeGeom g;
    double2 *desc;
    double2 *agg;
    uint32_t *stamp;      // [capacity]  last-writer-wins arbitration
    uint32_t *mask;       // [n_tiles]   touched-node bits of the running update
    int32_t *counts;      // [kMaxTileLevels] this call
    int32_t *counts_next; // [kMaxTileLevels] zeroed for the next call
    int32_t *err;
    int64_t n;            // <= cluster size * kUpdateThreads (the host chunks larger batches)
    const int64_t *n_dev; // optional: the entry count lives on the device (Mode G slot); this launch
    int64_t n_off;        //           covers entries [n_off, min(n_off + n, *n_dev))
    uint32_t stamp_base;  // entry i carries stamp 1 + stamp_base + i (chunks of one batch keep batch order)
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
    // Mode G: the launch that completes a write-back also publishes the new root to every shard's roots
    // board (exchange number pub_seq != 0), so the next global draw needs no exchange kernel before it
    BoardEntry *const *pub_boards;
    int32_t pub_shards, pub_shard;
    unsigned long long pub_seq;
    int32_t pub_when;     // 0 always; 1 only if the device-side entry count ends inside this launch (or before it);
                          // 2 only if this launch has entries (the catch-all behind launches sized for the expected share)
    // Rows appended right behind a write-back ride in its launch (config 4: a launch of their own costs the repair's
    // fixed ~30 us again): the last tail_n entries of the launch are the slots (tail_base + t) % capacity with float32
    // priorities tail_prio[t].  Needs prio_kind 0 and dedupe 1 (a sampled slot that is being recycled: the row wins).
    int64_t tail_n, tail_base;
    const float *tail_prio;
};

#define RETH_STAMP(k) do { if (a.dbg && brank == 0 && threadIdx.x == 0) a.dbg[k] = clock64(); } while (0)

__device__ __forceinline__ int64_t upd_index(const UpdateArgs &a, int64_t i)
{
    if (a.tail_n > 0 && i >= a.n - a.tail_n) return (a.tail_base + (i - (a.n - a.tail_n))) % a.g.capacity;
    if (a.idx_kind == 0) return ((const int64_t *)a.idx)[i];
    if (a.idx_kind == 1) return (int64_t)((const int32_t *)a.idx)[i];
    return (a.iota_base + i) % a.g.capacity;
}

// (w + 1e-6) ** alpha in the dtype of the incoming array (per_sampler.py:16-17):
// float32 in -> float32 add and a correctly rounded float32 power.
__device__ __forceinline__ double upd_prio_load(const UpdateArgs &a, int64_t i)
{
    if (a.prio_kind == 2) return 1.0;                    // prioritized_buffer.py:36-37
    if (a.tail_n > 0 && i >= a.n - a.tail_n) return (double)a.tail_prio[i - (a.n - a.tail_n)];
    if (a.prio_kind == 0) return (double)((const float *)a.prio)[i];   // exact widening
    return ((const double *)a.prio)[i];
}

__device__ __forceinline__ double upd_leaf(const UpdateArgs &a, double raw)
{
    if (a.prio_kind == 2) return 1.0;
    const bool f32 = a.prio_kind == 0;
    double base, expo = a.alpha;
    if (f32) {
        const float w = (float)raw;
        if (!a.transform) return (double)w;
        base = (double)(w + 1e-6f);
        expo = (double)(float)a.alpha;
    } else {
        const double w = raw;
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

// Warp-cooperative repair of up to K tiles of one tile level, lane = node.  The loads of ALL
// tiles are issued before the first use (one memory round trip per call instead of one per
// tile).  derived == false: exactly the nodes whose mask bit is set are recomputed.
// derived == true (tile levels 0 and 1): a node is also recomputed when a child has ever been
// maintained (child min != 0, which _numba_maintain_node guarantees for every node it visited):
// its ancestors were maintained then too, so recomputing them is idempotent and deep paths need
// no marking atomics on the two top tile levels.  `count` must be warp-uniform.
template <int K>
__device__ __forceinline__ void repair_tiles_warp(const UpdateArgs &a, const int64_t (&tile)[K], int count,
                                                  int T, int lane, bool derived)
{
    double2 ag[K], ex[K], de[K];
    uint32_t m[K];
#pragma unroll
    for (int q = 0; q < K; ++q) {
        if (q >= count) continue;
        ag[q] = ldcg2(&a.agg[tile[q] * 64 + lane]);
        ex[q] = ldcg2(&a.agg[tile[q] * 64 + 32 + lane]);
        de[q] = ldcg2(&a.desc[tile[q] * 32 + lane]);
        m[q] = __ldcg(&a.mask[tile[q]]);
    }
    // level-major so that the K independent dependency chains interleave
    bool wrote[K];
#pragma unroll
    for (int q = 0; q < K; ++q) wrote[q] = false;
#pragma unroll
    for (int l = 4; l >= 0; --l) {
        const int src = ((l == 4) ? 2 * (lane - 16) : 2 * lane) & 31;
#pragma unroll
        for (int q = 0; q < K; ++q) {
            if (q >= count) continue;
            const double sx = (l == 4) ? ex[q].x : ag[q].x, sy = (l == 4) ? ex[q].y : ag[q].y;
            double2 Lc, Rc;
            Lc.x = __shfl_sync(0xffffffffu, sx, src);
            Lc.y = __shfl_sync(0xffffffffu, sy, src);
            Rc.x = __shfl_sync(0xffffffffu, sx, src | 1);
            Rc.y = __shfl_sync(0xffffffffu, sy, src | 1);
            const bool own = (m[q] >> lane) & 1u;
            if ((lane >> l) == 1 && (own || (derived && (Lc.y != 0.0 || Rc.y != 0.0)))) {
                ag[q] = maintain_node(de[q].y, Lc, Rc);
                de[q].x = Lc.x;
                wrote[q] = true;
            }
        }
    }
#pragma unroll
    for (int q = 0; q < K; ++q) {
        if (q >= count) continue;
        if (wrote[q]) {
            __stcg(&a.agg[tile[q] * 64 + lane], ag[q]);
            __stcg(&a.desc[tile[q] * 32 + lane].x, de[q].x);
        }
        if (lane == 1 && T > 0) {
            const int64_t rel = tile[q] - tile_base(T);
            __stcg(&a.agg[(tile_base(T - 1) + (rel >> 5)) * 64 + 32 + (rel & 31)], ag[q]);
        }
        if (lane == 0 && m[q] != 0u) __stcg(&a.mask[tile[q]], 0u);
    }
}

// Batched priority write-back with tree-path repair (utils/sumtree.py:61-67, 24-31) for
// n <= (cluster size x 512) entries, one entry per thread.  One thread-block cluster; stages:
//   A stamp  (dedupe 1) atomicMax(stamp[idx], 1+base+i): the last i of a duplicate index wins; barrier
//   B leaves : winners write val and mark their root-ward path: tile levels >= 3 with a
//              returning atomicOr per tile (the first marker OWNS the tile and keeps it in a
//              register), tile level 2 with a fire-and-forget atomicOr, levels 0/1 not at all
//              (a leaf living there sets its own bit); cluster barrier
//   R sparse : tile levels deepest -> 3 (skipped when nobody owns a tile there): an owner whose
//              tile holds a single root-ward chain repairs it alone (repair_tile_chain), tiles
//              with several paths are repaired by the warp, two per round trip; cluster barrier
//   D dense  : tile level 2 (<= 1024 tiles): CTA c takes the children of the level-1 tiles
//              j = c (mod cluster), a warp repairs its touched tiles four per round trip;
//              __syncthreads; the CTA's level-1 tiles (derived rule); cluster barrier
//   ROOT     : tile 0 by one warp of CTA 0 (derived rule)
// The result equals the reference's sequential loop because sum/min are pure functions of
// the final leaves on the union of touched paths.
// kLv: tile levels compiled in (5 covers capacity <= 2^25 and keeps the per-level bookkeeping of the two
// deepest levels out of the register file; 7 covers the ABI's 2^31); kRegCtas: CTAs of this kernel that
// must fit one SM's register file (2 -> <= 64 registers at 512 threads): the repair is latency-bound
// and waits most of its life, so what matters is how many gather CTAs stay resident beside it.
// kRegCtas == 1 is capped at 104 registers, not the 128 the block size would allow: a 512-thread CTA then
// leaves 12 K registers on its SM, room for the CTAs of a draw kernel that programmatic dependent launch made
// resident ahead of its turn.  At 114 registers an SM holding ONE such waiting CTA could no longer host a
// repair CTA, and with the draw kernel's CTAs spread over every GPC no GPC had 16 free SMs for the cluster:
// two shards on one device then waited for each other until the exchange timed out.
// One round of the repair: entries [base, min(n, base + cluster size x block size)) of the launch.
template <int kLv, int kRegCtas>
__device__ __forceinline__ void update_round(const UpdateArgs &a, const int64_t base, const int64_t n, const bool publish)
{
    cg::cluster_group cluster = cg::this_cluster();
    const int nblk = (int)cluster.num_blocks(), brank = (int)cluster.block_rank();
    const int64_t cap = a.g.capacity;
    const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
    const int kWarps = (int)(blockDim.x >> 5);          // block size is a launch parameter (<= kUpdateThreads)
    const int NT = a.g.n_tile_levels;
    const int64_t per = (int64_t)nblk * blockDim.x;
    // consecutive entries are spread over the CTAs of the cluster
    const int64_t i = base + (int64_t)threadIdx.x * nblk + brank;
    const bool inrange = i < n;
    int64_t idx = 0, nxt = -1;
    double raw = 0.0;
    if (inrange) { idx = upd_index(a, i); raw = upd_prio_load(a, i); }
    if (a.dedupe == 2 && i + 1 < n) nxt = upd_index(a, i + 1);
    bool active = inrange;
    if (active && (idx < 0 || idx >= cap)) { atomicOr(a.err, 1); active = false; }
    const uint32_t my_stamp = 1u + a.stamp_base + (uint32_t)i;
    if (a.dedupe == 1) {
        if (active) atomicMax(&a.stamp[idx], my_stamp);
        RETH_STAMP(1);
        cluster.sync();
    }
    RETH_STAMP(2);
    uint32_t st = 0;
    if (active && a.dedupe == 1) st = __ldcg(&a.stamp[idx]);
    // dedupe 2: indices drawn by per_sample_kernel.  Stratified targets are monotone and
    // _numba_find_index is a monotone step function of its target, so equal indices are
    // always adjacent: the last element of a run is the sequential loop's last writer.
    if (active && a.dedupe == 2 && nxt == idx) active = false;
    const double leaf = active ? upd_leaf(a, raw) : 0.0;            // pow overlaps the stamp load
    RETH_STAMP(12);
    if (active && a.dedupe == 1) {
        if (st != my_stamp) active = false;
        else if (!a.keep_stamps) __stcg(&a.stamp[idx], 0u);       // stamps are zero between launches
    }
    int32_t own_tile[kLv];                             // tile this thread owns at level T (>= 3), else -1
#pragma unroll
    for (int T = 0; T < kLv; ++T) own_tile[T] = -1;
    if (active) {
        const NodeLoc n = locate(idx);
        __stcg(&a.desc[n.tile * 32 + n.local].y, leaf);
        if (n.T <= 1) {
            atomicOr(&a.mask[n.tile], 1u << n.local);             // own bit; ancestors are derived
        } else {
            int64_t tiles[kLv];
            uint32_t bits[kLv], old[kLv];
            int64_t tile = n.tile;
            int local = n.local, r = n.r;
            uint64_t kroot = ((uint64_t)idx + 1ull) >> n.r;       // heap k of the tile root
#pragma unroll
            for (int T = kLv - 1; T >= 2; --T) {
                if (T > n.T) continue;
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
            for (int T = kLv - 1; T >= 3; --T)
                if (T <= n.T) old[T] = atomicOr(&a.mask[tiles[T]], bits[T]);
            atomicOr(&a.mask[tiles[2]], bits[2]);                 // level 2 is swept densely: no owner
#pragma unroll
            for (int T = kLv - 1; T >= 3; --T)
                if (T <= n.T && old[T] == 0u) own_tile[T] = (int32_t)tiles[T];
        }
    }
#pragma unroll
    for (int T = kLv - 1; T >= 3; --T) {
        if (T >= NT) continue;                                    // uniform
        const unsigned bal = __ballot_sync(0xffffffffu, own_tile[T] >= 0);
        if (bal != 0u && lane == __ffs(bal) - 1) atomicOr(&a.counts[0], 1 << T);   // level has work
    }
    RETH_STAMP(3);
    cluster.sync();
    RETH_STAMP(4);

    // ---- R: sparse tile levels
    if (NT > 3) {
        const int flags = __ldcg(&a.counts[0]);                   // bit T: somebody owns a tile at level T
#pragma unroll
        for (int T = kLv - 1; T >= 3; --T) {
            if (T >= NT || !((flags >> T) & 1)) continue;         // uniform over the cluster
            const int32_t t = own_tile[T];
            uint32_t m = 0;
            if (t >= 0) m = __ldcg(&a.mask[t]);
            const bool chain = t >= 0 && m != 0u && __popc(m) == (31 - __clz(31 - __clz(m | 1u))) + 1;
            if (chain) repair_tile_chain(a, (int64_t)t, m, T);
            unsigned rest = __ballot_sync(0xffffffffu, t >= 0 && !chain);
            while (rest) {
                int64_t t2[2] = {0, 0};
                int cnt = 0;
#pragma unroll
                for (int q = 0; q < 2; ++q)
                    if (rest) {
                        const int src = __ffs(rest) - 1;
                        rest &= rest - 1;
                        t2[q] = (int64_t)__shfl_sync(0xffffffffu, t, src);
                        ++cnt;
                    }
                repair_tiles_warp<2>(a, t2, cnt, T, lane, false);
            }
            RETH_STAMP(5 + 2 * (T - 3));
            cluster.sync();
            RETH_STAMP(6 + 2 * (T - 3));
        }
    }

    // ---- D: tile level 2 (dense) and this CTA's level-1 tiles
    const int64_t n1 = (NT >= 2) ? ((a.g.n_tiles - 1 < 32) ? a.g.n_tiles - 1 : 32) : 0;     // tiles at level 1
    const int64_t n2 = (NT >= 3) ? ((a.g.n_tiles - 33 < 1024) ? a.g.n_tiles - 33 : 1024) : 0;   // tiles at level 2
    const int my_parents = (n1 > brank) ? (int)((n1 - brank + nblk - 1) / nblk) : 0;       // level-1 tiles j = brank + q*nblk
    if (NT >= 3) {
        const int n_e = my_parents * 32;                          // candidate level-2 tiles of this CTA
        for (int base = 0; base < n_e; base += 32 * kWarps) {
            const int e = base + wic + kWarps * lane;
            int64_t tile = -1;
            uint32_t m = 0;
            if (e < n_e) {
                const int64_t rel = ((int64_t)brank + (int64_t)(e >> 5) * nblk) * 32 + (e & 31);
                if (rel < n2) { tile = 33 + rel; m = __ldcg(&a.mask[tile]); }
            }
            unsigned todo = __ballot_sync(0xffffffffu, m != 0u);
            constexpr int kK2 = kRegCtas > 1 ? 2 : 4;             // tiles per round trip (register budget)
            while (todo) {
                int64_t tk[kK2];
                int cnt = 0;
#pragma unroll
                for (int q = 0; q < kK2; ++q) {
                    tk[q] = 0;
                    if (todo) {
                        const int src = __ffs(todo) - 1;
                        todo &= todo - 1;
                        tk[q] = __shfl_sync(0xffffffffu, tile, src);
                        ++cnt;
                    }
                }
                repair_tiles_warp<kK2>(a, tk, cnt, 2, lane, false);
            }
        }
        RETH_STAMP(20);
        __syncthreads();
    }
    if (NT >= 2) {
        for (int q = wic; q < my_parents; q += kWarps) {
            const int64_t t1[1] = {1 + (int64_t)brank + (int64_t)q * nblk};
            repair_tiles_warp<1>(a, t1, 1, 1, lane, true);
        }
        RETH_STAMP(23);
        cluster.sync();
        RETH_STAMP(24);
    }
    if (brank == 0 && wic == 0) {
        const int64_t t0[1] = {0};
        repair_tiles_warp<1>(a, t0, 1, 0, lane, true);
        if (publish && base + per >= n) {
            __syncwarp();              // lane 1 stored the root; order it before the other lanes' loads
            board_publish(a.pub_boards, a.pub_shards, a.pub_shard, a.pub_seq, a.agg, lane);
        }
    }
    RETH_STAMP(22);
}

// kRounds == false: one round, n <= cluster size x block size (the host cuts larger batches into launches).
// kRounds == true: the launch takes its n entries in rounds of (cluster size x block size), each a complete repair
// (batch order, hence last-writer-wins, is kept across rounds as across launches); how many rounds run is decided
// on the device.  The loop costs this build registers (spills), so it only serves as the CATCH-ALL behind a Mode G
// write-back: the host sizes the ordinary launches for the share it expects of a batch whose real share count
// lives on the device, and this one launch takes whatever lies beyond -- normally nothing, then it returns at once.
template <int kLv, int kRegCtas, bool kRounds>
__global__ void __maxnreg__(kRegCtas == 1 ? 104 : 64) per_update_kernel(const UpdateArgs a)
{
    cg::cluster_group cluster = cg::this_cluster();
    const int nblk = (int)cluster.num_blocks(), brank = (int)cluster.block_rank();
    const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;

    pdl_wait();
    pdl_trigger();
    RETH_STAMP(0);
    if (brank == 0 && threadIdx.x < kMaxTileLevels /* array size */) a.counts_next[threadIdx.x] = 0;
    int64_t n = a.n;
    bool publish = a.pub_seq != 0ull;
    if (a.n_dev) {
        const int64_t m = __ldcg(a.n_dev) - a.n_off;
        if (a.pub_when == 1) publish = publish && m <= a.n;      // the share ends here: nothing is left for the catch-all
        if (a.pub_when == 2) publish = publish && m > 0;
        n = m < n ? (m < 0 ? 0 : m) : n;
    }
    if (n <= 0) {           // Mode G: nothing of the share falls into this launch: uniform over the cluster
        if (publish && brank == 0 && wic == 0) board_publish(a.pub_boards, a.pub_shards, a.pub_shard, a.pub_seq, a.agg, lane);
        return;
    }
    if (!kRounds) {
        update_round<kLv, kRegCtas>(a, 0, n, publish);
    } else {
        const int64_t per = (int64_t)nblk * blockDim.x;
        for (int64_t base = 0; base < n; base += per) {
            update_round<kLv, kRegCtas>(a, base, n, publish);
            if (base + per < n) {      // another round: its level flags start clear, this round's stores are visible
                if (brank == 0 && threadIdx.x == 0) __stcg(&a.counts[0], 0);
                cluster.sync();
            }
        }
    }
}

// Zero the arbitration stamps of `n` slots (after a policy-slot scatter that needed them).
__global__ void stamp_clear_kernel(uint32_t *stamp, const int64_t *slots, int64_t n, int64_t cap)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && slots[i] >= 0 && slots[i] < cap) stamp[slots[i]] = 0u;
}


// ------------------------------------------------ non-PER samplers (plugin surface) --
// UniformSampler (reth_buffer/reth_buffer/sampler/uniform_sampler.py:6-25): a log of the first
// `capacity` notified slot indices; sample = log[randint(0, tail)], weights all one.
// FIFOSampler (sampler/fifo_sampler.py:8-29): a bounded queue of (index, weight); sample pops the
// oldest entries.  Host code owns the cursors (tail / head / count), kernels move the entries.
struct NoteArgs {
    const void *idx;
    int32_t idx_kind;          // 0 int64[], 1 int32[], 2 iota (iota_base + i) % capacity
    int64_t iota_base, capacity;
    const void *w;
    int32_t w_kind;            // 0 float32[], 1 float64[], 2 none
    int64_t n, skip;           // entries skip..n-1 are written (a push longer than the queue keeps its tail)
    int64_t *dst_idx;
    double *dst_w;
    int64_t dst_pos, dst_cap;  // entry i -> (dst_pos + i) % dst_cap
};

__global__ void sampler_note_kernel(const NoteArgs a)
{
    for (int64_t i = a.skip + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n;
         i += (int64_t)gridDim.x * blockDim.x) {
        int64_t v;
        if (a.idx_kind == 0) v = ((const int64_t *)a.idx)[i];
        else if (a.idx_kind == 1) v = (int64_t)((const int32_t *)a.idx)[i];
        else v = (a.iota_base + i) % a.capacity;
        const int64_t p = (a.dst_pos + i) % a.dst_cap;
        a.dst_idx[p] = v;
        if (a.dst_w) a.dst_w[p] = a.w_kind == 0 ? (double)((const float *)a.w)[i] : ((const double *)a.w)[i];
    }
}

struct PickArgs {
    const int64_t *src_idx;
    const double *src_w;       // nullptr: weights are 1.0 (uniform)
    int64_t src_cap, head;     // fifo: entry i = src[(head + i) % src_cap]
    const int64_t *draws;      // uniform: explicit positions in [0, tail) or nullptr (device generator)
    int64_t tail;
    uint64_t seed, draw_base;
    int32_t uniform;
    int64_t n;
    int64_t *idx_out;
    double *w_out, *prio_out;
    int64_t *meta;             // {first, count} per batch slot
    int32_t batch;
    int32_t *err;
    int64_t *idx_mirror;       // loader pipeline: mapped pinned host copies (see SampleArgs)
    double *w_mirror;
    int32_t *err_mirror;
};

__global__ void sampler_pick_kernel(const PickArgs a)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    if (a.err_mirror && i == 0) *a.err_mirror = *a.err;
    if (i % a.batch == 0) { a.meta[2 * (i / a.batch)] = 0; a.meta[2 * (i / a.batch) + 1] = a.batch; }
    int64_t p;
    if (a.uniform) {
        if (a.draws) {
            p = a.draws[i];
            if (p < 0 || p >= a.tail) { atomicOr(a.err, 1); p = 0; }
        } else {
            p = (int64_t)(uniform53(a.seed, a.draw_base + (uint64_t)i) * (double)a.tail);
            if (p >= a.tail) p = a.tail - 1;
        }
    } else {
        p = (a.head + i) % a.src_cap;
    }
    const int64_t v = a.src_idx[p];
    a.idx_out[i] = v;
    const double w = a.src_w ? a.src_w[p] : 1.0;
    a.w_out[i] = w;
    a.prio_out[i] = w;
    if (a.idx_mirror) { a.idx_mirror[i] = v; a.w_mirror[i] = w; }
}

// ------------------------------------------------ shard roots over NVLink --
// Mode G needs every shard's 16-byte root {total, min} on every rank before a global
// stratified draw.  The message is pure latency, so instead of a host-driven collective each shard
// owns a "roots board" in its HBM that every peer maps (CUDA IPC / peer access) and WRITES its root
// into with plain stores over NVLink; the same kernel then waits on its OWN board (local L2 polls)
// until the entries of the current exchange number arrived.  One launch per step on the sampling
// stream, no host round trip: a device-side all-gather fused between the tree update and the draw.
// Entries are double-buffered by the parity of the exchange number: a peer can publish exchange
// k+1 only after its own wait for k, which needed MY publish of k, so nobody is ever two ahead.
struct RootsArgs {
    const double2 *agg;                 // this shard's tree (root = agg[1])
    BoardEntry *board[kMaxShards];      // board of shard s as mapped here (own board for s == shard)
    double *all_roots;                  // out: double[2*n_shards] = {total_s, min_s} in rank order
    int32_t n_shards, shard, phases;    // phases bit0: publish, bit1: wait + snapshot
    unsigned long long seq;             // exchange number, >= 1
    unsigned long long timeout_ns;
    int32_t *err;
};

// one thread per shard
__global__ void __launch_bounds__(kMaxShards) roots_exchange_kernel(const RootsArgs a)
{
    pdl_wait();
    pdl_trigger();
    const int s = threadIdx.x;
    if (s >= a.n_shards) return;
    const int par = (int)(a.seq & 1ull);
    if (a.phases & 1) {
        const double2 r = ldcg2(&a.agg[1]);
        BoardEntry *e = a.board[s] + par * kMaxShards + a.shard;
        volatile double *vs = &e->sum, *vm = &e->mn;
        *vs = r.x;
        *vm = r.y;
        st_release_sys(&e->seq, a.seq);      // orders the two data stores before the flag, system scope
    }
    if (a.phases & 2) {
        const BoardEntry *e = a.board[a.shard] + par * kMaxShards + s;
        const unsigned long long t0 = global_ns();
        bool ok = true;
        while (ld_acquire_sys(&e->seq) != a.seq) {
            if (global_ns() - t0 > a.timeout_ns) { ok = false; break; }
        }
        if (!ok) { atomicOr(a.err, kErrBoardTimeout); return; }
        const volatile double *vs = &e->sum, *vm = &e->mn;
        a.all_roots[2 * s] = *vs;
        a.all_roots[2 * s + 1] = *vm;
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
