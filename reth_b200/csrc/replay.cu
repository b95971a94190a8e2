// replay.cu -- C ABI (include/reth_replay.h) of the B200-native PER shard.
// Host-side state + launches; kernels live in tree.cuh / copy.cuh.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false ...
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include <unistd.h>

#include <nvtx3/nvToolsExt.h>

#include "../../include/reth_replay.h"
#include "copy.cuh"
#include "lz4f.h"
#include "nstep.cuh"
#include "tree.cuh"

using namespace reth;

#define RETH_EXPORT extern "C" __attribute__((visibility("default")))

namespace {

thread_local std::string g_create_error;

struct BatchSlot {
    uint8_t *cols[kMaxCols] = {};
    int64_t *idx = nullptr;
    double *w = nullptr;
    double *prio = nullptr;
    int64_t *meta = nullptr;   // {first, count}
    int32_t batch = 0;
    bool partial = false;      // Mode G draw: only meta[1] of the `batch` entries belong to this shard
    int shards = 1;            // ... of a global batch drawn over this many shards (expected share batch / shards)
    int packed = 0;            // > 0: first slot of a packed Mode G group of this many draws (rows = meta[1] over all of them)
};

}  // namespace

struct reth_replay {
    reth_replay_config cfg{};
    TreeGeom geom{};
    int n_sms = 148;
    // tree
    double2 *desc = nullptr, *agg = nullptr;
    uint32_t *stamp = nullptr, *mask = nullptr;
    int32_t *counts = nullptr /*[2][kMaxTileLevels]*/, *err = nullptr;
    int64_t max_n = 0;            // largest n of one update launch
    int counts_flip = 0;
    uint64_t draw_counter = 0;
    // staging for host-resident arguments
    void *stage_idx = nullptr, *stage_prio = nullptr;
    double *stage_u = nullptr;
    int64_t *host_pin = nullptr;  // small pinned scratch (err/meta readback)
    // ring + batch slots
    uint8_t *ring[kMaxCols] = {};
    // batch slots are carved from contiguous arrays so that G consecutive slots form one
    // G*max_batch batch (group mode)
    uint8_t *slot_cols[kMaxCols] = {};
    int64_t *slot_idx = nullptr, *slot_meta = nullptr;
    double *slot_w = nullptr, *slot_prio = nullptr;
    std::vector<BatchSlot> slots;
    int64_t tail = 0, size = 0;
    reth_replay_stats st{};
    // step loop
    cudaStream_t s_tree = nullptr, s_gather = nullptr, s_gather2 = nullptr;   // gathers alternate between two streams
    int two_gather_streams = 1;
    std::vector<cudaEvent_t> ev_sampled, ev_gathered;
    cudaEvent_t ev_fork = nullptr, ev_join_t = nullptr, ev_join_g = nullptr;
    cudaStream_t s_append = nullptr;          // large device-source appends: the row copy runs beside the tree update
    cudaEvent_t ev_app_fork = nullptr, ev_app_done = nullptr, ev_ingest = nullptr;
    cudaEvent_t last_gather_ev = nullptr;     // newest gather of the internal streams
    cudaEvent_t last_gather_on[2] = {nullptr, nullptr};   // ... per gather stream (s_gather, s_gather2): a writer of ring
                                              // rows waits for BOTH, the other stream's previous gather may still run
    // n-step assembly at append time (reth_replay_nstep_*): per environment stream, the raw rows that have not
    // left the adder yet (double-buffered), and a staging area for host-resident rows
    struct NStep {
        bool open = false;
        int n_streams = 0, step = 0, r_col = 0, s1_col = 0, done_col = 0;
        double gamma = 0.0;
        int64_t max_rows = 0;
        uint8_t *carry[2][kMaxCols] = {};
        uint8_t *stage[kMaxCols] = {};
        std::vector<int> count, cur;
    } nstep;
    int update_cluster = 8;               // largest cluster the device hosts (16, else 8), or the caller's choice
    int update_cluster_fixed = 0;         // the caller (config / RETH_UPDATE_CLUSTER) chose the cluster size
    int update_threads = kUpdateThreads;  // threads per CTA of the repair kernel
    int update_lean = -1;                 // repair-kernel build: 0 full registers, 1 lean (<= 64), -1 by launch size
    int64_t out_rb[kMaxCols] = {};      // bytes of one row of each BATCH column (4x row_bytes when the gather widens)
    int col_frames[kMaxCols] = {};      // K > 0: the column stores K int32 frame references per row
    int n_frame_cols = 0;
    uint8_t *frames = nullptr;          // frame ring: frame_capacity x frame_bytes
    int64_t frame_tail = 0, frames_appended = 0;
    int64_t *row_oldest = nullptr;      // [capacity] oldest frame (sequence number) each stored row refers to
    int max_dyn_smem = 0;               // opt-in shared memory per block
    bool gather_bulk = false;           // gather through row_gather_bulk_kernel (TMA bulk copies)
    int bulk_grid_cap = 0;              // RETH_BULK_GRID at create time: cap on the persistent grid (tests)
    long long *dbg = nullptr;           // RETH_DEBUG_TIMING=1: stage clocks of the last update launch
    std::vector<cudaEvent_t> ev_time;   // pairs (start, stop) around timed gather launches
    size_t ev_time_used = 0;
    int64_t ev_time_launches = 0;       // gather launches covered by the recorded brackets
    // non-PER sampler plugins (allocated on first use)
    int64_t *uni_log = nullptr, *fifo_idx = nullptr;
    double *fifo_w = nullptr;
    int64_t uni_tail = 0, fifo_head = 0, fifo_len = 0;
    int pipe_kind = 0;                  // sampler of pipe_issue: 0 PER, 1 uniform, 2 fifo
    // asynchronous loader pipeline (reth_replay_pipe_*)
    struct Pipe {
        bool open = false;
        int batch = 0, G = 0, n_groups = 0, upd_entries = 0;
        int64_t E = 0;
        uint64_t seed = 0;
        int64_t *idx_host = nullptr;
        double *w_host = nullptr;
        int32_t *err_host = nullptr;
        uint8_t *upd_host = nullptr;
        uint8_t *upd_dev = nullptr;
        cudaStream_t s_copy = nullptr;
        std::vector<cudaEvent_t> ev_smp, ev_gat, ev_done, ev_upd;
        std::vector<char> upd_used;
        int64_t issued = 0, waited = 0, upd_next = 0;
        int hold = 1;             // delivered groups that stay lent (the newest `hold` of them)
        // experiment (RETH_PIPE_ZEROCOPY=1): kernels write deliveries to / read write-backs from mapped pinned
        // host memory instead of using the copy engines.  Measured slower on B200 (strict alternation
        // 7.7 vs 8.1 M transitions/s, group 8 / one ahead 39.6 vs 43.6 M): PCIe round trips inside the
        // update kernel's dependent chain cost more than the copy operations they save.
        bool zero_copy = false;
    } pipe;
    // shard roots over NVLink (reth_replay_board_*): own board, peers' boards as mapped here, exchange number
    BoardEntry *board = nullptr;
    BoardEntry *peer_board[kMaxShards] = {};
    bool peer_ipc[kMaxShards] = {};   // opened with cudaIpcOpenMemHandle (to close at destroy)
    double *all_roots = nullptr;
    int board_shards = 0, board_shard = 0;
    unsigned long long board_seq = 0;
    BoardEntry **board_ptrs_dev = nullptr;   // peer_board[] on the device (the repair kernel publishes through it)
    unsigned long long wait_seq = 0;         // set around a draw: take the roots of this exchange from the board
    int modeg_fused = 1;                     // Mode G loop: publish folded into the repair kernel, wait into the draw
    bool quota_first = false;                // set by the Mode G loop: the share kernel runs before single draws too
    int64_t *quota = nullptr;     // Mode G prefetch group: {first, count, offset, 0} per draw of the running launch
    int64_t *slot_quota = nullptr; // the same per batch slot (kept for the consumer of the packed batch)
    bool pdl = false;             // tree-chain launches carry the programmatic-dependent-launch attribute (set by step_loop)
    int pipe_reserved = 0;        // batch slots at the end of the ring that a pipe never uses (other sampler topics)
    // set by pipe_issue around its draw: the sampling kernel also writes into mapped pinned host memory
    int64_t *mirror_idx = nullptr;
    double *mirror_w = nullptr;
    int32_t *mirror_err = nullptr;
    std::string error;
};

namespace {

int fail(reth_replay *h, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (h) h->error = buf; else g_create_error = buf;
    return code;
}

#define CK(h, expr)                                                                       \
    do {                                                                                  \
        cudaError_t e__ = (expr);                                                         \
        if (e__ != cudaSuccess)                                                           \
            return fail(h, e__ == cudaErrorMemoryAllocation ? RETH_ERR_OOM : RETH_ERR_CUDA, \
                        "%s: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

template <class T>
cudaError_t dalloc_zero(T **p, size_t count, int64_t *acct)
{
    cudaError_t e = cudaMalloc((void **)p, count * sizeof(T));
    if (e != cudaSuccess) return e;
    *acct += (int64_t)(count * sizeof(T));
    return cudaMemset(*p, 0, count * sizeof(T));
}

inline cudaStream_t S(void *s) { return (cudaStream_t)s; }

// NVTX ranges around the four stages of the path (append / draw / gather / repair), SURVEY.md 5 "Tracing": on with
// RETH_NVTX=1 in the environment (a profiler attaches to them; without one the calls are no-ops of the header-only
// NVTX v3 loader), off by default so that the enqueue path stays free of them.
struct NvtxRange {
    bool on;
    explicit NvtxRange(const char *name)
    {
        static const bool enabled = [] { const char *e = getenv("RETH_NVTX"); return e && atoi(e) != 0; }();
        on = enabled;
        if (on) nvtxRangePushA(name);
    }
    ~NvtxRange() { if (on) nvtxRangePop(); }
};

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

size_t dtype_size(int dt) { return (dt == RETH_F32 || dt == RETH_I32) ? 4 : 8; }

// A gather was enqueued on one of the internal gather streams and `ev` recorded behind it.
void note_gather(reth_replay *h, cudaEvent_t ev, cudaStream_t sg)
{
    h->last_gather_ev = ev;
    h->last_gather_on[sg == h->s_gather2 ? 1 : 0] = ev;
}
void forget_gathers(reth_replay *h)
{
    h->last_gather_ev = nullptr;
    h->last_gather_on[0] = h->last_gather_on[1] = nullptr;
}
// Writers of ring rows (append, append_at, append_frames, n-step append) wait for every gather still in flight:
// a gathered row is never half old, half new (the reference's LMDB rows are replaced atomically,
// server/main_loop.py:55-57).  Consecutive gathers alternate between two streams, so the newest event of EACH
// stream is waited for -- the newest gather overall says nothing about the one before it on the other stream.
int wait_gathers(reth_replay *h, cudaStream_t s)
{
    for (int k = 0; k < 2; ++k)
        if (h->last_gather_on[k]) CK(h, cudaStreamWaitEvent(s, h->last_gather_on[k], 0));
    return RETH_OK;
}

void free_batch_ring(reth_replay *h);
int alloc_batch_ring(reth_replay *h);

void free_all(reth_replay *h)
{
    cudaFree(h->desc); cudaFree(h->agg); cudaFree(h->stamp); cudaFree(h->mask);
    cudaFree(h->counts); cudaFree(h->err);
    cudaFree(h->uni_log); cudaFree(h->fifo_idx); cudaFree(h->fifo_w);
    cudaFree(h->dbg);
    for (int i = 0; i < kMaxShards; ++i)
        if (h->peer_ipc[i] && h->peer_board[i]) cudaIpcCloseMemHandle(h->peer_board[i]);
    cudaFree(h->board); cudaFree(h->all_roots); cudaFree(h->quota); cudaFree(h->board_ptrs_dev);
    cudaFree(h->stage_idx); cudaFree(h->stage_prio);
    if (h->host_pin) cudaFreeHost(h->host_pin);
    for (int c = 0; c < kMaxCols; ++c) cudaFree(h->ring[c]);
    cudaFree(h->frames); cudaFree(h->row_oldest);
    free_batch_ring(h);
    for (int c = 0; c < kMaxCols; ++c) { cudaFree(h->nstep.carry[0][c]); cudaFree(h->nstep.carry[1][c]); cudaFree(h->nstep.stage[c]); }
    for (auto e : h->ev_time) cudaEventDestroy(e);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_app_fork) cudaEventDestroy(h->ev_app_fork);
    if (h->ev_app_done) cudaEventDestroy(h->ev_app_done);
    if (h->ev_ingest) cudaEventDestroy(h->ev_ingest);
    if (h->s_append) cudaStreamDestroy(h->s_append);
    if (h->ev_join_t) cudaEventDestroy(h->ev_join_t);
    if (h->ev_join_g) cudaEventDestroy(h->ev_join_g);
    if (h->s_tree) cudaStreamDestroy(h->s_tree);
    if (h->s_gather) cudaStreamDestroy(h->s_gather);
    if (h->s_gather2) cudaStreamDestroy(h->s_gather2);
}


// The learner-batch ring: ring_depth slots of max_batch rows, carved from one contiguous array per column
// (so G consecutive slots form one G*max_batch batch), their per-slot events, and the uniform staging buffer.
void free_batch_ring(reth_replay *h)
{
    for (int c = 0; c < kMaxCols; ++c) {
        if (h->slot_cols[c]) h->st.hbm_bytes -= (int64_t)h->slots.size() * h->cfg.max_batch * h->out_rb[c];
        cudaFree(h->slot_cols[c]);
        h->slot_cols[c] = nullptr;
    }
    if (h->slot_idx) h->st.hbm_bytes -= (int64_t)h->slots.size() * ((int64_t)h->cfg.max_batch * 24 + 16 + 32);
    cudaFree(h->slot_idx); cudaFree(h->slot_w); cudaFree(h->slot_prio); cudaFree(h->slot_meta);
    cudaFree(h->slot_quota); cudaFree(h->stage_u);
    h->slot_idx = nullptr; h->slot_w = nullptr; h->slot_prio = nullptr; h->slot_meta = nullptr;
    h->slot_quota = nullptr; h->stage_u = nullptr;
    for (auto e : h->ev_sampled) cudaEventDestroy(e);
    for (auto e : h->ev_gathered) cudaEventDestroy(e);
    h->ev_sampled.clear();
    h->ev_gathered.clear();
    h->slots.clear();
    forget_gathers(h);
}

int alloc_batch_ring(reth_replay *h)
{
    int64_t *acct = &h->st.hbm_bytes;
    const size_t depth = (size_t)h->cfg.ring_depth, mb = (size_t)h->cfg.max_batch;
    h->slots.assign(depth, BatchSlot());
    if (h->cfg.store_rows)
        for (int c = 0; c < h->cfg.n_columns; ++c) {
            CK(h, cudaMalloc((void **)&h->slot_cols[c], depth * mb * h->out_rb[c]));
            *acct += (int64_t)(depth * mb * h->out_rb[c]);
        }
    CK(h, dalloc_zero(&h->slot_idx, depth * mb, acct));
    CK(h, dalloc_zero(&h->slot_w, depth * mb, acct));
    CK(h, dalloc_zero(&h->slot_prio, depth * mb, acct));
    CK(h, dalloc_zero(&h->slot_meta, depth * 2, acct));
    CK(h, dalloc_zero(&h->slot_quota, depth * 4, acct));
    if (!h->quota) CK(h, dalloc_zero(&h->quota, (size_t)4 * 32, acct));
    CK(h, cudaMalloc((void **)&h->stage_u, mb * 8));
    for (size_t i = 0; i < depth; ++i) {
        BatchSlot &s = h->slots[i];
        for (int c = 0; c < h->cfg.n_columns; ++c)
            s.cols[c] = h->slot_cols[c] ? h->slot_cols[c] + i * mb * h->out_rb[c] : nullptr;
        s.idx = h->slot_idx + i * mb;
        s.w = h->slot_w + i * mb;
        s.prio = h->slot_prio + i * mb;
        s.meta = h->slot_meta + i * 2;
    }
    h->ev_sampled.resize(depth);
    h->ev_gathered.resize(depth);
    for (size_t i = 0; i < depth; ++i) {
        CK(h, cudaEventCreateWithFlags(&h->ev_sampled[i], cudaEventDisableTiming));
        CK(h, cudaEventCreateWithFlags(&h->ev_gathered[i], cudaEventDisableTiming));
    }
    return RETH_OK;
}
}  // namespace

namespace {
using UpdateKernel = void (*)(const UpdateArgs);
// the instantiation of the repair kernel for this shard: 5 tile levels cover capacity <= 2^25
UpdateKernel update_kernel_of(const reth_replay *h, bool lean, bool rounds = false)
{
    const bool small = h->geom.n_tile_levels <= 5;
    if (rounds) return small ? per_update_kernel<5, 2, true> : per_update_kernel<kMaxTileLevels, 2, true>;
    if (lean) return small ? per_update_kernel<5, 2, false> : per_update_kernel<kMaxTileLevels, 2, false>;
    return small ? per_update_kernel<5, 1, false> : per_update_kernel<kMaxTileLevels, 1, false>;
}

// Shape of one repair launch.  The kernel is latency-bound (3 cluster barriers, 5 dependent stages) and a
// frame gather normally runs beside it, so what it costs the loop is the SM resources it HOLDS while it
// waits (measured, profiles/r02_footprint_sweep.jsonl: prefetch-group loop 51.2 -> 52.8 M transitions/s).
//   n <= 2048 (strict alternation, prefetch groups of <= 4 batches: the repair is on the critical path, the
//     gather of so few rows is shorter than the tree chain): 16 CTAs, full register budget -- fastest launch;
//   n <= 4096 (one prefetch group): 8 CTAs x 512 threads of the lean build (<= 64 registers: half an SM's
//     register file per CTA, so three gather CTAs stay resident on each of only 8 SMs);
//   larger (appends): lean, 16 CTAs, 8192 entries per launch.
int env_int(const char *name, int dflt);
struct UpdateShape { UpdateKernel fn; int cluster; };
UpdateShape update_shape_of(const reth_replay *h, int64_t n, bool critical = false)
{
    // critical: the repair is on a cross-GPU dependent chain (Mode G: every shard's next draw waits for every
    // shard's root), so its latency counts, not the SM resources it holds: widest cluster, full register budget
    if (critical && h->update_lean < 0 && !h->update_cluster_fixed) return {update_kernel_of(h, false), h->update_cluster};
    static const int fast_n = env_int("RETH_UPDATE_FAST_N", 2048);
    int cluster = h->update_cluster;
    if (!h->update_cluster_fixed && n > fast_n && n <= (int64_t)8 * h->update_threads && cluster > 8) cluster = 8;
    const bool lean = h->update_lean >= 0 ? h->update_lean != 0 : n > fast_n;
    return {update_kernel_of(h, lean), cluster};
}
}  // namespace

RETH_EXPORT int reth_replay_abi_version(void) { return 3; }

// lz4 frame codec (host code, lz4f.h)
RETH_EXPORT int64_t reth_lz4f_content_size(const void *src, int64_t n)
{
    if (!src || n < 0) return lz4f::kErrFormat;
    return lz4f::content_size((const uint8_t *)src, n);
}
RETH_EXPORT int64_t reth_lz4f_decompress(const void *src, int64_t n, void *dst, int64_t cap)
{
    if (!src || n < 0 || cap < 0 || (!dst && cap > 0)) return lz4f::kErrFormat;
    return lz4f::decompress((const uint8_t *)src, n, (uint8_t *)dst, cap);
}
RETH_EXPORT int64_t reth_lz4f_compress_bound(int64_t n) { return n < 0 ? lz4f::kErrFormat : lz4f::compress_bound(n); }
RETH_EXPORT int64_t reth_lz4f_compress(const void *src, int64_t n, void *dst, int64_t cap)
{
    if (n < 0 || (!src && n > 0) || !dst) return lz4f::kErrFormat;
    return lz4f::compress((const uint8_t *)src, n, (uint8_t *)dst, cap);
}

// Row messages -> columns (host code, no handle): the n row messages of one reference-format append
// (client/client.py:26-35: every row is its own serialize()d message, same struct, hence same length L and the
// same column offsets) form an (n, L) byte matrix `mat`; column c of row i sits at mat[i*L + lo[c] .. + len[c]).
// One call copies every column of every row to dst[c] + i*len[c] -- C-contiguous columns, e.g. in pinned
// staging memory -- with no interpreter in the loop (ctypes drops the GIL for the duration: several ingest
// threads run this concurrently).  Replaces the per-row deserialize loop of client/numpy_loader.py:35-50 on the
// ingest side.
RETH_EXPORT int64_t reth_pack_scatter_rows(const void *mat, int64_t n, int64_t L, int32_t ncols, const int64_t *lo,
                                          const int64_t *len, void *const *dst)
{
    if (!mat || n < 0 || L < 0 || ncols < 0 || (ncols > 0 && (!lo || !len || !dst))) return -1;
    for (int c = 0; c < ncols; ++c)
        if (lo[c] < 0 || len[c] < 0 || lo[c] + len[c] > L || !dst[c]) return -1;
    const uint8_t *m = (const uint8_t *)mat;
    int64_t bytes = 0;
    for (int64_t i = 0; i < n; ++i)
        for (int c = 0; c < ncols; ++c) {
            memcpy((uint8_t *)dst[c] + i * len[c], m + i * L + lo[c], (size_t)len[c]);
            bytes += len[c];
        }
    return bytes;
}

RETH_EXPORT const char *reth_replay_last_error(const reth_replay_t *h)
{
    return h ? h->error.c_str() : g_create_error.c_str();
}

RETH_EXPORT int reth_replay_create(const reth_replay_config *cfg, reth_replay_t **out)
{
    if (!cfg || !out) return fail(nullptr, RETH_ERR_INVALID, "null argument");
    *out = nullptr;
    if (cfg->capacity < 1) return fail(nullptr, RETH_ERR_INVALID, "capacity must be >= 1");
    if (cfg->capacity > (1ll << 31)) return fail(nullptr, RETH_ERR_INVALID, "capacity > 2^31 unsupported");
    if (cfg->n_columns < 0 || cfg->n_columns > RETH_MAX_COLUMNS)
        return fail(nullptr, RETH_ERR_INVALID, "n_columns out of range");
    if (cfg->max_batch < 1) return fail(nullptr, RETH_ERR_INVALID, "max_batch must be >= 1");
    for (int c = 0; c < cfg->n_columns; ++c)
        if (cfg->row_bytes[c] < 1) return fail(nullptr, RETH_ERR_INVALID, "row_bytes[%d] < 1", c);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(nullptr, RETH_ERR_CUDA, "no CUDA device: the replay engine has no CPU fallback");
    if (cfg->device < 0 || cfg->device >= ndev) return fail(nullptr, RETH_ERR_INVALID, "bad device ordinal");

    reth_replay *h = new (std::nothrow) reth_replay();
    if (!h) return fail(nullptr, RETH_ERR_OOM, "host allocation failed");
    h->cfg = *cfg;
    if (h->cfg.ring_depth < 1) h->cfg.ring_depth = 2;
    h->geom = make_geom(cfg->capacity);
    for (int c = 0; c < cfg->n_columns; ++c) {
        const int cast = cfg->out_cast[c] & ~RETH_COL_FRAMES;
        if (cfg->out_cast[c] < 0 || cast > 2) { delete h; return fail(nullptr, RETH_ERR_INVALID, "out_cast[%d] invalid", c); }
        h->cfg.out_cast[c] = cast;                          // the cast alone from here on
        if (cfg->out_cast[c] & RETH_COL_FRAMES) {
            if (cfg->frame_bytes < 16 || cfg->frame_bytes % 16 || cfg->frame_capacity < 1 || cfg->row_bytes[c] % 4 || !cfg->store_rows) {
                delete h;
                return fail(nullptr, RETH_ERR_INVALID, "frame column %d needs frame_bytes %% 16 == 0, frame_capacity >= 1, row_bytes = 4 * K", c);
            }
            h->col_frames[c] = (int)(cfg->row_bytes[c] / 4);
            h->n_frame_cols++;
            h->out_rb[c] = (int64_t)h->col_frames[c] * cfg->frame_bytes * (cast ? 4 : 1);
        } else {
            h->out_rb[c] = cfg->row_bytes[c] * (cast ? 4 : 1);
        }
    }
    DeviceGuard guard(cfg->device);
    auto bail = [&](int code) { free_all(h); g_create_error = h->error; delete h; return code; };
#define CKC(expr)                                                                          \
    do {                                                                                   \
        cudaError_t e__ = (expr);                                                          \
        if (e__ != cudaSuccess) {                                                          \
            fail(h, e__ == cudaErrorMemoryAllocation ? RETH_ERR_OOM : RETH_ERR_CUDA,       \
                 "%s: %s", #expr, cudaGetErrorString(e__));                                \
            cudaGetLastError();                                                            \
            return bail(e__ == cudaErrorMemoryAllocation ? RETH_ERR_OOM : RETH_ERR_CUDA);  \
        }                                                                                  \
    } while (0)
    cudaDeviceProp prop;
    CKC(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major < 9)
        return bail(fail(h, RETH_ERR_CUDA, "device sm_%d%d: built for sm_100a only", prop.major, prop.minor));
    h->n_sms = prop.multiProcessorCount;
    h->max_dyn_smem = (int)prop.sharedMemPerBlockOptin;
    {
        int mode = cfg->gather_mode;
        if (const char *e = getenv("RETH_GATHER_MODE")) mode = atoi(e);
        bool any_wide = false;
        for (int c = 0; c < cfg->n_columns; ++c)
            if (cfg->row_bytes[c] % 16 == 0 && cfg->row_bytes[c] >= 256) any_wide = true;
        for (int c = 0; c < cfg->n_columns; ++c)
            if (cfg->out_cast[c]) any_wide = false;          // the DMA gather cannot widen (or expand frame references): LDG/STG kernel
        h->gather_bulk = (mode == 2) && any_wide;   // auto = LDG/STG.128: measured faster (profiles/r01_gather_sweep.log)
        if (const char *e = getenv("RETH_BULK_GRID")) h->bulk_grid_cap = atoi(e);
        if (h->gather_bulk)
            CKC(cudaFuncSetAttribute(row_gather_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     h->max_dyn_smem));
    }
    int64_t *acct = &h->st.hbm_bytes;
    const size_t nt = (size_t)h->geom.n_tiles;
    CKC(dalloc_zero(&h->desc, nt * 32, acct));
    CKC(dalloc_zero(&h->agg, nt * 64, acct));
    CKC(dalloc_zero(&h->stamp, (size_t)cfg->capacity, acct));
    CKC(dalloc_zero(&h->mask, nt, acct));
    h->max_n = cfg->max_batch > 16384 ? cfg->max_batch : 16384;
    CKC(dalloc_zero(&h->counts, 2 * kMaxTileLevels, acct));
    CKC(dalloc_zero(&h->err, 4, acct));
    CKC(cudaMalloc(&h->stage_idx, h->max_n * 8));
    CKC(cudaMalloc(&h->stage_prio, h->max_n * 8));
    CKC(cudaMallocHost((void **)&h->host_pin, 64 * sizeof(int64_t)));
    if (getenv("RETH_DEBUG_TIMING")) CKC(dalloc_zero(&h->dbg, 32, acct));
    if (cfg->store_rows)
        for (int c = 0; c < cfg->n_columns; ++c) {
            CKC(cudaMalloc((void **)&h->ring[c], (size_t)cfg->capacity * cfg->row_bytes[c]));
            *acct += cfg->capacity * cfg->row_bytes[c];
        }
    if (h->n_frame_cols) {
        CKC(cudaMalloc((void **)&h->frames, (size_t)cfg->frame_capacity * cfg->frame_bytes));
        *acct += (int64_t)cfg->frame_capacity * cfg->frame_bytes;
        CKC(dalloc_zero(&h->row_oldest, (size_t)cfg->capacity, acct));
    }
    {
        const int rc_ring = alloc_batch_ring(h);
        if (rc_ring) return bail(rc_ring);
    }
    int lo = 0, hi = 0;
    CKC(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CKC(cudaStreamCreateWithPriority(&h->s_tree, cudaStreamNonBlocking, hi));
    CKC(cudaStreamCreateWithPriority(&h->s_gather, cudaStreamNonBlocking, lo));
    CKC(cudaStreamCreateWithPriority(&h->s_gather2, cudaStreamNonBlocking, lo));
    if (const char *e = getenv("RETH_GATHER_STREAMS")) h->two_gather_streams = atoi(e) >= 2;
    CKC(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
    CKC(cudaEventCreateWithFlags(&h->ev_app_fork, cudaEventDisableTiming));
    CKC(cudaEventCreateWithFlags(&h->ev_app_done, cudaEventDisableTiming));
    CKC(cudaEventCreateWithFlags(&h->ev_ingest, cudaEventDisableTiming));
    CKC(cudaStreamCreateWithFlags(&h->s_append, cudaStreamNonBlocking));
    CKC(cudaEventCreateWithFlags(&h->ev_join_t, cudaEventDisableTiming));
    CKC(cudaEventCreateWithFlags(&h->ev_join_g, cudaEventDisableTiming));
    if (const char *e = getenv("RETH_UPDATE_LEAN")) h->update_lean = atoi(e);
    if (const char *e = getenv("RETH_UPDATE_THREADS")) {
        const int t = atoi(e);
        if (t == 128 || t == 256 || t == 512) h->update_threads = t;
    }
    int want_cluster = cfg->update_cluster;
    if (const char *e = getenv("RETH_UPDATE_CLUSTER")) want_cluster = atoi(e);
    if (want_cluster > 16) want_cluster = 16;
    for (int lean = 0; lean < 2; ++lean)
        CKC(cudaFuncSetAttribute(update_kernel_of(h, lean != 0), cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    CKC(cudaFuncSetAttribute(update_kernel_of(h, true, true), cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    if (want_cluster <= 0) {
        // default: up to a 16-CTA cluster (non-portable size) when the device can host one
        cudaLaunchConfig_t lc = {};
        lc.gridDim = dim3(16);
        lc.blockDim = dim3(h->update_threads);
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 16; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        lc.attrs = at;
        lc.numAttrs = 1;
        int nclusters = 0;
        if (cudaOccupancyMaxActiveClusters(&nclusters, update_kernel_of(h, false), &lc) == cudaSuccess && nclusters > 0)
            h->update_cluster = 16;
        else
            h->update_cluster = 8;
        cudaGetLastError();
    } else {
        h->update_cluster = want_cluster;
        h->update_cluster_fixed = 1;
    }
    CKC(cudaDeviceSynchronize());
#undef CKC
    *out = h;
    return RETH_OK;
}

namespace { void pipe_free(reth_replay *h); }

RETH_EXPORT int reth_replay_destroy(reth_replay_t *h)
{
    if (!h) return RETH_OK;
    {
        DeviceGuard guard(h->cfg.device);
        cudaDeviceSynchronize();
        pipe_free(h);
        free_all(h);
    }
    delete h;
    return RETH_OK;
}

// ------------------------------------------------------------ launch helpers --
namespace {

// Bring a host-resident argument array to the device staging buffer.
int to_device(reth_replay *h, const void *p, int loc, size_t bytes, void *staging,
              cudaStream_t s, const void **out)
{
    if (loc == RETH_DEVICE) { *out = p; return RETH_OK; }
    CK(h, cudaMemcpyAsync(staging, p, bytes, cudaMemcpyHostToDevice, s));
    *out = staging;
    return RETH_OK;
}

// Tree update of n entries: one launch of the cluster kernel per (cluster size x block size) entries, in batch
// order on one stream, so last-writer-wins holds across the launches too.
// n_expected > 0 (Mode G: only the first *a.n_dev of the n entries are valid, about n_expected of them): ordinary
// launches cover 1.25 x n_expected + 64 entries and ONE catch-all launch (the build with rounds) takes whatever lies
// beyond -- normally nothing.  A write-back padded to 8 x the share then costs 2 launches on the chain, not 8..16.
int launch_update(reth_replay *h, UpdateArgs &a, cudaStream_t s, int64_t n_expected = 0, bool critical = false)
{
    NvtxRange nvtx("reth:repair");
    if (a.n <= 0) return RETH_OK;
    a.g = h->geom; a.desc = h->desc; a.agg = h->agg; a.stamp = h->stamp; a.mask = h->mask;
    a.err = h->err; a.dbg = h->dbg;
    const int64_t n = a.n;
    const unsigned long long pub_seq = a.pub_seq;
    if (pub_seq) { a.pub_boards = h->board_ptrs_dev; a.pub_shards = h->board_shards; a.pub_shard = h->board_shard; }
    const bool spec = a.n_dev != nullptr && n_expected > 0;
    const UpdateShape shape = update_shape_of(h, spec ? n_expected : n, critical);
    const int64_t per = (int64_t)shape.cluster * h->update_threads;
    int64_t covered = n;                 // entries [0, covered) go through ordinary launches
    if (spec) {
        covered = (int64_t)(1.25 * (double)n_expected) + 64;
        covered = (covered + per - 1) / per * per;
        if (covered > n) covered = n;
    }
    const uint8_t *idx0 = (const uint8_t *)a.idx, *prio0 = (const uint8_t *)a.prio;
    const int64_t iota0 = a.iota_base;
    // appended rows riding behind a write-back: entries [n - tail, n) of the batch
    const int64_t tail = a.tail_n, tail_base0 = a.tail_base, n_main = n - tail;
    const float *tail_prio0 = a.tail_prio;
    if (tail < 0 || tail > n || (tail > 0 && (a.n_dev || a.prio_kind != 0 || a.dedupe != 1 || !tail_prio0)))
        return fail(h, RETH_ERR_INVALID, "update: bad tail entries");
    const size_t is = a.idx_kind == 0 ? 8 : 4, ps = a.prio_kind == 0 ? 4 : 8;
    for (int64_t off = 0; off < n;) {
        const bool rest = off >= covered;                    // the catch-all launch: everything that is left
        const int64_t m = rest ? n - off : (covered - off < per ? covered - off : per);
        a.n = m;
        a.n_off = off;
        // the launch that completes the write-back publishes the new root: with a catch-all, the last ordinary
        // launch when the share ends inside it (the normal case: the publish does not wait for an empty launch),
        // else the catch-all
        a.pub_seq = (off + m >= n || (spec && off + m >= covered)) ? pub_seq : 0ull;
        a.pub_when = !spec || covered >= n ? 0 : (rest ? 2 : 1);
        a.stamp_base = (uint32_t)off;
        if (a.idx_kind != 2) a.idx = idx0 + off * is;
        if (a.prio_kind != 2) a.prio = prio0 + off * ps;
        a.iota_base = iota0 + off;
        if (tail > 0) {
            const int64_t skip = off > n_main ? off - n_main : 0;       // tail entries taken by earlier launches
            a.tail_n = off + m > n_main ? off + m - n_main - skip : 0;
            a.tail_base = tail_base0 + skip;
            a.tail_prio = tail_prio0 + skip;
        }
        a.counts = h->counts + h->counts_flip * kMaxTileLevels;
        a.counts_next = h->counts + (h->counts_flip ^ 1) * kMaxTileLevels;
        h->counts_flip ^= 1;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(shape.cluster);
        cfg.blockDim = dim3(h->update_threads);
        cfg.stream = s;
        cudaLaunchAttribute at[2];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = shape.cluster;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[1].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = h->pdl ? 2 : 1;
        CK(h, cudaLaunchKernelEx(&cfg, rest ? update_kernel_of(h, true, true) : shape.fn, (const UpdateArgs)a));
        h->st.kernel_launches++;
        off += m;
    }
    a.n = n; a.idx = idx0; a.prio = prio0; a.iota_base = iota0; a.pub_seq = pub_seq;
    a.tail_n = tail; a.tail_base = tail_base0; a.tail_prio = tail_prio0;
    return RETH_OK;
}

int set_prio(reth_replay *h, UpdateArgs &a, const void *prio, int prio_dtype, int prio_loc,
             int64_t off, int64_t n, cudaStream_t s)
{
    if (!prio) { a.prio = nullptr; a.prio_kind = 2; return RETH_OK; }
    if (prio_dtype != RETH_F32 && prio_dtype != RETH_F64)
        return fail(h, RETH_ERR_INVALID, "priorities must be float32 or float64");
    const size_t es = dtype_size(prio_dtype);
    a.prio_kind = prio_dtype == RETH_F32 ? 0 : 1;
    return to_device(h, (const uint8_t *)prio + off * es, prio_loc, n * es, h->stage_prio, s, &a.prio);
}

// What a non-zero device-side error flag means.
int device_err(reth_replay *h, int32_t flag)
{
    if (flag & kErrBoardTimeout)
        return fail(h, RETH_ERR_STATE, "roots exchange timed out: a peer shard did not publish its root");
    if (flag & kErrStaleFrame)
        return fail(h, RETH_ERR_STATE, "a sampled transition refers to a frame that has been overwritten: the frame "
                                       "ring (frame_capacity) is too small for the transitions it must outlive");
    return fail(h, RETH_ERR_RANGE, "an index outside [0, capacity) was passed to the shard");
}

// Device-side error flag (index out of range seen by a kernel); checked at sync points.
int check_device_err(reth_replay *h, cudaStream_t s)
{
    int32_t *flag = (int32_t *)h->host_pin;
    CK(h, cudaMemcpyAsync(flag, h->err, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    CK(h, cudaStreamSynchronize(s));
    if (*flag) {
        cudaMemsetAsync(h->err, 0, sizeof(int32_t), s);
        return device_err(h, *flag);
    }
    return RETH_OK;
}

int slot_ok(reth_replay *h, int slot)
{
    if (slot < 0 || slot >= (int)h->slots.size()) return fail(h, RETH_ERR_INVALID, "bad batch slot %d", slot);
    return RETH_OK;
}

}  // namespace

RETH_EXPORT int reth_replay_clear(reth_replay_t *h, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    const size_t nt = (size_t)h->geom.n_tiles;
    CK(h, cudaMemsetAsync(h->desc, 0, nt * 32 * sizeof(double2), s));
    CK(h, cudaMemsetAsync(h->agg, 0, nt * 64 * sizeof(double2), s));
    CK(h, cudaMemsetAsync(h->mask, 0, nt * sizeof(uint32_t), s));
    CK(h, cudaMemsetAsync(h->stamp, 0, (size_t)h->cfg.capacity * sizeof(uint32_t), s));
    CK(h, cudaMemsetAsync(h->counts, 0, 2 * kMaxTileLevels * sizeof(int32_t), s));
    h->tail = 0;
    h->size = 0;
    h->uni_tail = h->fifo_head = h->fifo_len = 0;   // UniformSampler.clear / FIFOSampler.clear
    return RETH_OK;
}

namespace {
// Record the oldest referenced frame of the rows just stored at n slots (see frame_mark_kernel).
int mark_frame_refs(reth_replay *h, const int64_t *slots_dev, int64_t first, int64_t n, cudaStream_t s)
{
    if (!h->n_frame_cols || n <= 0) return RETH_OK;
    MarkArgs a = {};
    for (int c = 0; c < h->cfg.n_columns; ++c)
        if (h->col_frames[c]) {
            a.refs[a.ncols] = h->ring[c]; a.rb[a.ncols] = h->cfg.row_bytes[c]; a.K[a.ncols] = h->col_frames[c];
            a.ncols++;
        }
    a.slots = slots_dev; a.first = first; a.n = n; a.capacity = h->cfg.capacity;
    a.frame_tail = h->frame_tail; a.frames_appended = h->frames_appended; a.frame_capacity = h->cfg.frame_capacity;
    a.row_oldest = h->row_oldest; a.err = h->err;
    frame_mark_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(a);
    CK(h, cudaGetLastError());
    h->st.kernel_launches++;
    return RETH_OK;
}
}  // namespace

RETH_EXPORT int reth_replay_batch_ring_resize(reth_replay_t *h, int max_batch, int ring_depth)
{
    if (!h) return RETH_ERR_INVALID;
    if (max_batch < 1 || ring_depth < 1) return fail(h, RETH_ERR_INVALID, "batch ring: max_batch and ring_depth must be >= 1");
    if (h->pipe.open) return fail(h, RETH_ERR_STATE, "batch ring: close the loader pipe first");
    DeviceGuard guard(h->cfg.device);
    CK(h, cudaDeviceSynchronize());                       // nothing may still read or fill a slot
    free_batch_ring(h);
    h->cfg.max_batch = max_batch;
    h->cfg.ring_depth = ring_depth;
    h->pipe_reserved = 0;
    const int64_t want = max_batch > 16384 ? max_batch : 16384;
    if (want > h->max_n) {                                // staging of host-resident index / priority arguments
        cudaFree(h->stage_idx); cudaFree(h->stage_prio);
        h->stage_idx = h->stage_prio = nullptr;
        h->max_n = want;
        CK(h, cudaMalloc(&h->stage_idx, h->max_n * 8));
        CK(h, cudaMalloc(&h->stage_prio, h->max_n * 8));
    }
    int rc = alloc_batch_ring(h);
    if (rc) return rc;
    CK(h, cudaDeviceSynchronize());
    return RETH_OK;
}

RETH_EXPORT int reth_replay_append_frames(reth_replay_t *h, const void *frames, int frames_loc, int64_t n,
                                          int64_t *first_frame_out, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (!h->frames) return fail(h, RETH_ERR_STATE, "this shard has no frame ring (no frame-reference column)");
    const int64_t cap = h->cfg.frame_capacity, fb = h->cfg.frame_bytes;
    if (n < 0 || n > cap) return fail(h, RETH_ERR_INVALID, "append of %lld frames exceeds the frame ring (%lld)", (long long)n, (long long)cap);
    if (n > 0 && !frames) return fail(h, RETH_ERR_INVALID, "frames is null");
    if (first_frame_out) *first_frame_out = h->frame_tail;
    if (n == 0) return RETH_OK;
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    { const int rcw = wait_gathers(h, s); if (rcw) return rcw; }   // a running gather may still read the slots
    const cudaMemcpyKind kind = frames_loc == RETH_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    const int64_t len1 = n < cap - h->frame_tail ? n : cap - h->frame_tail, len2 = n - len1;
    CK(h, cudaMemcpyAsync(h->frames + h->frame_tail * fb, frames, (size_t)(len1 * fb), kind, s));
    if (len2 > 0) CK(h, cudaMemcpyAsync(h->frames, (const uint8_t *)frames + len1 * fb, (size_t)(len2 * fb), kind, s));
    h->frame_tail = (h->frame_tail + n) % cap;
    h->frames_appended += n;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_frame_read(reth_replay_t *h, int64_t first_frame, int64_t n, void *out_host, void *stream)
{
    if (!h || !out_host) return RETH_ERR_INVALID;
    if (!h->frames) return fail(h, RETH_ERR_STATE, "this shard has no frame ring");
    const int64_t cap = h->cfg.frame_capacity, fb = h->cfg.frame_bytes;
    if (first_frame < 0 || n < 0 || first_frame + n > cap) return fail(h, RETH_ERR_INVALID, "frames %lld..%lld outside the ring", (long long)first_frame, (long long)(first_frame + n));
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    CK(h, cudaMemcpyAsync(out_host, h->frames + first_frame * fb, (size_t)(n * fb), cudaMemcpyDeviceToHost, s));
    CK(h, cudaStreamSynchronize(s));
    return RETH_OK;
}

RETH_EXPORT int reth_replay_append(reth_replay_t *h, const void *const *cols, int cols_loc,
                                   const void *prio, int prio_dtype, int prio_loc, int64_t n,
                                   double alpha, int transform, int64_t *first_slot_out, void *stream)
{
    NvtxRange nvtx("reth:append");
    if (!h) return RETH_ERR_INVALID;
    if (n < 0 || n > h->cfg.capacity)   // fifo_policy.py:12
        return fail(h, RETH_ERR_INVALID, "append of %lld rows exceeds capacity", (long long)n);
    if (h->cfg.store_rows && h->cfg.n_columns > 0 && !cols) return fail(h, RETH_ERR_INVALID, "cols is null");
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    const int64_t cap = h->cfg.capacity;
    const int64_t first_slot = h->tail;
    if (first_slot_out) *first_slot_out = h->tail;
    // a gather still running on the internal streams must not see half-written rows (the reference's
    // LMDB rows are replaced atomically, main_loop.py:55-57)
    if (h->cfg.store_rows && n > 0) { const int rcw = wait_gathers(h, s); if (rcw) return rcw; }
    bool forked = false;
    for (int64_t off = 0; off < n; off += h->max_n) {
        const int64_t m = (n - off < h->max_n) ? n - off : h->max_n;
        const int64_t start = h->tail;
        if (h->cfg.store_rows && h->cfg.n_columns > 0) {
            const int64_t len1 = (m < cap - start) ? m : cap - start, len2 = m - len1;
            if (cols_loc == RETH_DEVICE) {
                AppendArgs a = {};
                a.ncols = h->cfg.n_columns; a.start = start; a.n = m; a.capacity = cap;
                int64_t bytes = 0;
                for (int c = 0; c < a.ncols; ++c) {
                    a.rb[c] = h->cfg.row_bytes[c];
                    a.src[c] = (const uint8_t *)cols[c] + off * a.rb[c];
                    a.ring[c] = h->ring[c];
                    bytes += m * a.rb[c];
                }
                int64_t blocks = (bytes / 16 + kCopyThreads * 4 - 1) / (kCopyThreads * 4);
                const int64_t cap_blocks = (int64_t)h->n_sms * 8;
                if (blocks > cap_blocks) blocks = cap_blocks;
                if (blocks < 1) blocks = 1;
                cudaStream_t sc = s;
                if (bytes >= (32ll << 20)) {           // big copy: run it beside the tree update of the same rows
                    if (!forked) {
                        CK(h, cudaEventRecord(h->ev_app_fork, s));
                        CK(h, cudaStreamWaitEvent(h->s_append, h->ev_app_fork, 0));
                        forked = true;
                    }
                    sc = h->s_append;
                }
                static const int use_memcpy = [] { const char *e = getenv("RETH_APPEND_MEMCPY"); return e ? atoi(e) : 0; }();
                if (use_memcpy && bytes >= (32ll << 20)) {
                    // experiment: the driver's device-to-device copy per column span instead of the append kernel
                    for (int c = 0; c < a.ncols; ++c) {
                        CK(h, cudaMemcpyAsync(a.ring[c] + start * a.rb[c], a.src[c], (size_t)(len1 * a.rb[c]), cudaMemcpyDeviceToDevice, sc));
                        if (len2 > 0)
                            CK(h, cudaMemcpyAsync(a.ring[c], a.src[c] + len1 * a.rb[c], (size_t)(len2 * a.rb[c]), cudaMemcpyDeviceToDevice, sc));
                    }
                } else {
                    // 16-byte-aligned spans: one CTA per contiguous 16 KB tile; otherwise the grid-strided kernel
                    AppendTiledArgs t = {};
                    bool aligned = true;
                    for (int c = 0; c < a.ncols && aligned; ++c) {
                        const int64_t rb = a.rb[c];
                        uint8_t *d1 = a.ring[c] + start * rb;
                        const uint8_t *s1 = a.src[c], *s2 = a.src[c] + len1 * rb;
                        if ((((uintptr_t)d1 | (uintptr_t)s1 | (uint64_t)(len1 * rb)) & 15) ||
                            (len2 > 0 && (((uintptr_t)a.ring[c] | (uintptr_t)s2 | (uint64_t)(len2 * rb)) & 15))) { aligned = false; break; }
                        if (len1 > 0) {
                            AppendSeg &g = t.seg[t.nseg++];
                            g.dst = d1; g.src = s1; g.units = len1 * rb / 16; g.tile_base = t.tiles;
                            t.tiles += (g.units + kAppendTileUnits - 1) / kAppendTileUnits;
                        }
                        if (len2 > 0) {
                            AppendSeg &g = t.seg[t.nseg++];
                            g.dst = a.ring[c]; g.src = s2; g.units = len2 * rb / 16; g.tile_base = t.tiles;
                            t.tiles += (g.units + kAppendTileUnits - 1) / kAppendTileUnits;
                        }
                    }
                    static const int tiled = [] { const char *e = getenv("RETH_APPEND_TILED"); return e ? atoi(e) : 1; }();
                    if (aligned && tiled && t.tiles > 0) {
                        const int64_t maxg = (int64_t)h->n_sms * 8 * 64;
                        ring_append_tiled_kernel<<<(unsigned)(t.tiles < maxg ? t.tiles : maxg), kCopyThreads, 0, sc>>>(t);
                    } else {
                        ring_append_kernel<<<(unsigned)blocks, kCopyThreads, 0, sc>>>(a);
                    }
                    CK(h, cudaGetLastError());
                    h->st.kernel_launches++;
                }
            } else {
                for (int c = 0; c < h->cfg.n_columns; ++c) {
                    const int64_t rb = h->cfg.row_bytes[c];
                    const uint8_t *src = (const uint8_t *)cols[c] + off * rb;
                    CK(h, cudaMemcpyAsync(h->ring[c] + start * rb, src, len1 * rb, cudaMemcpyHostToDevice, s));
                    if (len2 > 0)
                        CK(h, cudaMemcpyAsync(h->ring[c], src + len1 * rb, len2 * rb, cudaMemcpyHostToDevice, s));
                }
            }
        }
        UpdateArgs a = {};
        a.n = m; a.idx = nullptr; a.idx_kind = 2; a.iota_base = start;
        a.alpha = alpha; a.transform = transform; a.dedupe = 0;
        int rc = set_prio(h, a, prio, prio_dtype, prio_loc, off, m, s);
        if (rc) return rc;
        rc = launch_update(h, a, s);
        if (rc) return rc;
        h->tail = (start + m) % cap;
        h->size = (h->size + m < cap) ? h->size + m : cap;
        h->st.appended_rows += m;
    }
    if (forked) {
        CK(h, cudaEventRecord(h->ev_app_done, h->s_append));
        CK(h, cudaStreamWaitEvent(s, h->ev_app_done, 0));
    }
    if (h->cfg.store_rows && n > 0) {
        const int rcm = mark_frame_refs(h, nullptr, first_slot, n, s);
        if (rcm) return rcm;
    }
    h->st.tail = h->tail; h->st.size = h->size;
    return RETH_OK;
}

// ------------------------------------------------------------ n-step assembly --
RETH_EXPORT int reth_replay_nstep_open(reth_replay_t *h, int n_streams, int step, double gamma, int r_col, int s1_col,
                                       int done_col, int64_t max_rows)
{
    if (!h) return RETH_ERR_INVALID;
    auto &ns = h->nstep;
    if (ns.open) return fail(h, RETH_ERR_STATE, "n-step assembly is already open");
    const int nc = h->cfg.n_columns;
    if (!h->cfg.store_rows || nc < 1) return fail(h, RETH_ERR_STATE, "n-step assembly needs a shard that stores rows");
    if (n_streams < 1 || step < 1 || max_rows < 1) return fail(h, RETH_ERR_INVALID, "n_streams, step and max_rows must be >= 1");
    if (r_col < 0 || r_col >= nc || s1_col < 0 || s1_col >= nc || done_col < 0 || done_col >= nc || r_col == s1_col ||
        r_col == done_col || s1_col == done_col)
        return fail(h, RETH_ERR_INVALID, "reward / s1 / done columns must be three different columns of the row struct");
    if (h->cfg.row_bytes[r_col] != 4 && h->cfg.row_bytes[r_col] != 8)
        return fail(h, RETH_ERR_INVALID, "the reward column must be one float32 or float64 per row");
    DeviceGuard guard(h->cfg.device);
    for (int c = 0; c < nc; ++c) {
        const size_t rb = (size_t)h->cfg.row_bytes[c];
        for (int b = 0; b < 2; ++b) {
            CK(h, cudaMalloc((void **)&ns.carry[b][c], (size_t)n_streams * step * rb));
            h->st.hbm_bytes += (int64_t)((size_t)n_streams * step * rb);
        }
        CK(h, cudaMalloc((void **)&ns.stage[c], (size_t)max_rows * rb));
        h->st.hbm_bytes += (int64_t)((size_t)max_rows * rb);
    }
    ns.n_streams = n_streams; ns.step = step; ns.gamma = gamma; ns.r_col = r_col; ns.s1_col = s1_col;
    ns.done_col = done_col; ns.max_rows = max_rows;
    ns.count.assign(n_streams, 0);
    ns.cur.assign(n_streams, 0);
    ns.open = true;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_nstep_reset(reth_replay_t *h, int stream_id)
{
    if (!h) return RETH_ERR_INVALID;
    auto &ns = h->nstep;
    if (!ns.open) return fail(h, RETH_ERR_STATE, "n-step assembly is not open");
    if (stream_id < -1 || stream_id >= ns.n_streams) return fail(h, RETH_ERR_INVALID, "bad stream id %d", stream_id);
    for (int k = 0; k < ns.n_streams; ++k)
        if (stream_id < 0 || k == stream_id) ns.count[k] = 0;      // a fresh NStepAdder: buffered rows are dropped
    return RETH_OK;
}

RETH_EXPORT int reth_replay_append_nstep(reth_replay_t *h, int stream_id, const void *const *cols, int cols_loc,
                                         const void *prio, int prio_dtype, int prio_loc, int64_t n, double alpha,
                                         int transform, int64_t *first_slot_out, int64_t *emitted_out, void *stream)
{
    NvtxRange nvtx("reth:append_nstep");
    if (!h) return RETH_ERR_INVALID;
    auto &ns = h->nstep;
    if (!ns.open) return fail(h, RETH_ERR_STATE, "n-step assembly is not open");
    if (stream_id < 0 || stream_id >= ns.n_streams) return fail(h, RETH_ERR_INVALID, "bad stream id %d", stream_id);
    if (n < 0 || (n > 0 && !cols)) return fail(h, RETH_ERR_INVALID, "bad rows");
    if (prio && prio_dtype != RETH_F32 && prio_dtype != RETH_F64) return fail(h, RETH_ERR_INVALID, "priorities must be float32 or float64");
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    const int nc = h->cfg.n_columns;
    const int64_t cap = h->cfg.capacity;
    if (first_slot_out) *first_slot_out = h->tail;
    int64_t emitted = 0;
    if (n > 0) { const int rcw = wait_gathers(h, s); if (rcw) return rcw; }
    const size_t ps = prio ? dtype_size(prio_dtype) : 0;
    for (int64_t off = 0; off < n; off += ns.max_rows) {
        const int64_t m = (n - off < ns.max_rows) ? n - off : ns.max_rows;
        NStepArgs a = {};
        a.ncols = nc; a.step = ns.step; a.r_col = ns.r_col; a.s1_col = ns.s1_col; a.done_col = ns.done_col;
        a.r_f64 = h->cfg.row_bytes[ns.r_col] == 8;
        a.carry_count = ns.count[stream_id];
        a.n = m; a.tail = h->tail; a.capacity = cap; a.gamma = ns.gamma;
        const int64_t L = a.carry_count + m;
        a.E = L > ns.step ? L - ns.step : 0;
        if (a.E > cap) return fail(h, RETH_ERR_INVALID, "n-step append of %lld rows exceeds capacity", (long long)a.E);
        const int cur = ns.cur[stream_id];
        for (int c = 0; c < nc; ++c) {
            NStepCol &col = a.col[c];
            const int64_t rb = h->cfg.row_bytes[c];
            col.rb = rb;
            col.carry = ns.carry[cur][c] + (size_t)stream_id * ns.step * rb;
            col.carry_next = ns.carry[cur ^ 1][c] + (size_t)stream_id * ns.step * rb;
            col.ring = h->ring[c];
            const uint8_t *src = (const uint8_t *)cols[c] + off * rb;
            if (cols_loc == RETH_DEVICE) col.fresh = src;
            else {
                CK(h, cudaMemcpyAsync(ns.stage[c], src, (size_t)(m * rb), cudaMemcpyHostToDevice, s));
                col.fresh = ns.stage[c];
            }
            col.part_bytes = 8192;
            col.parts = rb > 8192 ? (int32_t)((rb + 8191) / 8192) : 1;
            col.part_base = a.parts_per_row;
            a.parts_per_row += col.parts;
        }
        const int64_t maxg = (int64_t)h->n_sms * 8 * 16;
        if (a.E > 0) {
            const int64_t items = a.E * a.parts_per_row;
            nstep_emit_kernel<<<(unsigned)(items < maxg ? items : maxg), kCopyThreads, 0, s>>>(a);
            CK(h, cudaGetLastError());
            h->st.kernel_launches++;
        }
        if (a.E > 0) { const int rcm = mark_frame_refs(h, nullptr, h->tail, a.E, s); if (rcm) return rcm; }
        const int keep = (int)(L < ns.step ? L : ns.step);
        {
            const int64_t items = (int64_t)keep * a.parts_per_row;
            nstep_carry_kernel<<<(unsigned)(items < maxg ? items : maxg), kCopyThreads, 0, s>>>(a, keep);
            CK(h, cudaGetLastError());
            h->st.kernel_launches++;
        }
        if (a.E > 0) {
            // the row that push p of this call sends out of the adder carries prio[p]: emitted row e left at push
            // e + step - carry_count, so the emitted rows' priorities are one contiguous slice
            const int64_t p0 = off + (ns.step - a.carry_count);
            for (int64_t uo = 0; uo < a.E; uo += h->max_n) {
                const int64_t um = a.E - uo < h->max_n ? a.E - uo : h->max_n;
                UpdateArgs u = {};
                u.n = um; u.idx = nullptr; u.idx_kind = 2; u.iota_base = h->tail + uo;
                u.alpha = alpha; u.transform = transform; u.dedupe = 0;
                int rc = set_prio(h, u, prio ? (const uint8_t *)prio + (p0 + uo) * ps : nullptr, prio_dtype, prio_loc, 0, um, s);
                if (rc) return rc;
                rc = launch_update(h, u, s);
                if (rc) return rc;
            }
            h->tail = (h->tail + a.E) % cap;
            h->size = (h->size + a.E < cap) ? h->size + a.E : cap;
            h->st.appended_rows += a.E;
            emitted += a.E;
        }
        ns.count[stream_id] = keep;
        ns.cur[stream_id] = cur ^ 1;
    }
    h->st.tail = h->tail; h->st.size = h->size;
    if (emitted_out) *emitted_out = emitted;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_update(reth_replay_t *h, const void *idx, int idx_dtype, int idx_loc,
                                   const void *prio, int prio_dtype, int prio_loc, int64_t n,
                                   double alpha, int transform, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (n < 0 || (n > 0 && !idx)) return fail(h, RETH_ERR_INVALID, "bad indices");
    if (idx_dtype != RETH_I32 && idx_dtype != RETH_I64)
        return fail(h, RETH_ERR_INVALID, "indices must be int32 or int64");
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    const size_t es = dtype_size(idx_dtype);
    // chunks keep batch order: later chunks run later on the stream, so
    // last-writer-wins holds across chunk boundaries too
    for (int64_t off = 0; off < n; off += h->max_n) {
        const int64_t m = (n - off < h->max_n) ? n - off : h->max_n;
        UpdateArgs a = {};
        a.n = m; a.idx_kind = idx_dtype == RETH_I64 ? 0 : 1;
        a.alpha = alpha; a.transform = transform; a.dedupe = 1;
        int rc = to_device(h, (const uint8_t *)idx + off * es, idx_loc, m * es, h->stage_idx, s, &a.idx);
        if (rc) return rc;
        rc = set_prio(h, a, prio, prio_dtype, prio_loc, off, m, s);
        if (rc) return rc;
        rc = launch_update(h, a, s);
        if (rc) return rc;
        h->st.updated_rows += m;
    }
    return RETH_OK;
}

RETH_EXPORT int reth_replay_update_slot(reth_replay_t *h, int slot, const void *prio, int prio_dtype,
                                        int prio_loc, double alpha, int transform, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (slot_ok(h, slot)) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    BatchSlot &b = h->slots[slot];
    if (b.batch <= 0) return fail(h, RETH_ERR_STATE, "slot %d holds no sampled batch", slot);
    UpdateArgs a = {};
    a.n = b.batch; a.idx = b.idx; a.idx_kind = 0;
    if (b.partial) a.n_dev = b.meta + 1;                      // Mode G: this shard's share of the global batch
    a.alpha = alpha; a.transform = transform; a.dedupe = 2;   // own stratified draw: duplicates are adjacent
    int rc = set_prio(h, a, prio, prio_dtype, prio_loc, 0, b.batch, S(stream));
    if (rc) return rc;
    rc = launch_update(h, a, S(stream), b.partial ? b.batch / b.shards + 1 : 0);
    if (rc) return rc;
    h->st.updated_rows += b.batch;
    return RETH_OK;
}

// ------------------------------------------------------------ sample / gather --
namespace {

int env_int(const char *name, int dflt)
{
    const char *e = getenv(name);
    return e ? atoi(e) : dflt;
}

int gather_units_target()
{
    static int v = [] {
        const char *e = getenv("RETH_GATHER_UNITS");
        int x = e ? atoi(e) : 1024;
        return x < 64 ? 64 : x;
    }();
    return v;
}

// experiment switch: plain wide columns through the warp-per-part kernel as well
int gather_warp_mode()
{
    static int v = [] { const char *e = getenv("RETH_GATHER_WARP"); return e ? atoi(e) : 0; }();
    return v;
}

// Work decomposition of a row move over `batch` rows (gather or scatter).
void build_row_args(reth_replay *h, RowArgs &a, uint8_t *const *batch_cols, const int64_t *idx, int batch,
                    int units_target = 0, bool widen = false)
{
    if (units_target <= 0) {
        units_target = gather_units_target();
        // small launches (strict alternation, a Mode G share): the tree chain is what is critical there, and its
        // cluster launch waits for gather CTAs to retire -- shorter-lived gather CTAs (8 KB items) cost the chain less
        // (strict loop 25.1 -> 23.4 us per iteration, scripts/strict_probe.py)
        static const bool fixed = getenv("RETH_GATHER_UNITS") != nullptr;
        if (!fixed && batch <= 1024 && units_target > 512) units_target = 512;
    }
    a = RowArgs{};
    a.ncols = h->cfg.n_columns;
    a.idx = idx;
    a.batch = batch;
    a.capacity = h->cfg.capacity;
    a.err = h->err;
    int64_t items = 0;
    bool narrow = false;
    for (int c = 0; c < a.ncols; ++c) {
        RowCol &col = a.col[c];
        col.ring = h->ring[c];
        col.batch = batch_cols[c];
        col.rb = h->cfg.row_bytes[c];
        col.cast = widen ? h->cfg.out_cast[c] : 0;
        const bool wide = (col.rb % 16 == 0) && col.rb >= 256 && ((uintptr_t)col.batch % 16 == 0);
        if (widen && (h->col_frames[c] || ((col.cast || gather_warp_mode()) && wide))) {
            // gather: frame-reference columns and widened wide columns go through frame_gather_kernel
            col.chunks = -1;
            col.cast = 0;
            continue;
        }
        if (wide) {
            const int64_t upr = col.rb / 16;
            col.chunks = (int32_t)((upr + units_target - 1) / units_target);
            col.units = (int32_t)((upr + col.chunks - 1) / col.chunks);
            col.item_base = items;
            col.chunk_base = a.chunks_per_row;
            a.chunks_per_row += col.chunks;
            items += (int64_t)batch * col.chunks;
        } else {
            col.chunks = 0;
            narrow = true;
        }
    }
    a.wide_items = items;
    a.narrow_groups = narrow ? (batch + kCopyThreads - 1) / kCopyThreads : 0;
}

template <bool kScatter>
int launch_row_move(reth_replay *h, const RowArgs &a, cudaStream_t s)
{
    bool widen = false;
    for (int c = 0; c < a.ncols; ++c) widen |= a.col[c].cast != 0;
    int64_t total = a.wide_items + a.narrow_groups;
    if (total <= 0) return RETH_OK;
    static const int waves = [] { const char *e = getenv("RETH_GATHER_WAVES"); int v = e ? atoi(e) : 64; return v < 1 ? 1 : v; }();
    const int64_t maxg = (int64_t)h->n_sms * 6 * waves;     // resident CTAs x waves; beyond that CTAs stride over the items
    unsigned grid = (unsigned)(total < maxg ? total : maxg);
    if (!kScatter && !widen && a.meta && a.share_div > 1) {
        // Mode G share: CTAs beyond the valid items leave at once, but 0.4 us per thousand of them adds up
        // (48 k surplus CTAs = 22 us): size the grid for twice the expected share, a larger share strides
        // (stratified targets: a shard's share is its fraction of the global total +- 1 row per draw, so the
        // slack only has to cover unequal shard totals; a larger share strides)
        static const double slack = [] { const char *e = getenv("RETH_SHARE_GRID"); return e ? atof(e) : 1.25; }();
        int64_t want = (int64_t)(slack * (double)(total / a.share_div)) + 64;
        if (want < (int64_t)h->n_sms * 6 * 4) want = (int64_t)h->n_sms * 6 * 4;
        if (want < (int64_t)grid) grid = (unsigned)want;
    }
    if (!kScatter && widen) row_move_kernel<false, true><<<grid, kCopyThreads, 0, s>>>(a);
    else if (!kScatter && a.meta) row_move_kernel<false, false, true><<<grid, kCopyThreads, 0, s>>>(a);   // row count on the device
    else row_move_kernel<kScatter><<<grid, kCopyThreads, 0, s>>>(a);
    CK(h, cudaGetLastError());
    h->st.kernel_launches++;
    return RETH_OK;
}

// Frame-reference columns of a gather (frame_gather_kernel); no-op when the struct has none.
int launch_frame_gather(reth_replay *h, uint8_t *const *batch_cols, const int64_t *idx, int batch,
                        const int64_t *meta, int dst_row_from_meta, cudaStream_t s)
{
    FrameArgs a = {};
    a.idx = idx; a.batch = batch; a.meta = meta; a.dst_row_from_meta = dst_row_from_meta;
    a.capacity = h->cfg.capacity; a.err = h->err;
    for (int c = 0; c < h->cfg.n_columns; ++c) {
        const int64_t rb = h->cfg.row_bytes[c];
        const bool direct = !h->col_frames[c] && (h->cfg.out_cast[c] || gather_warp_mode()) && rb % 16 == 0 && rb >= 256 &&
                            ((uintptr_t)batch_cols[c] % 16 == 0);
        if (!h->col_frames[c] && !direct) continue;
        FrameCol &col = a.col[a.ncols++];
        col.refs = h->ring[c]; col.batch = batch_cols[c]; col.rb = rb; col.cast = h->cfg.out_cast[c];
        if (direct) {
            col.direct = 1; col.src = h->ring[c]; col.fb = rb; col.fcap = h->cfg.capacity; col.K = 1;
        } else {
            col.src = h->frames; col.fb = h->cfg.frame_bytes; col.fcap = h->cfg.frame_capacity; col.K = h->col_frames[c];
        }
        col.split = col.K >= 8 ? 1 : (8 + col.K - 1) / col.K;      // K * split >= the 8 warps of a CTA
    }
    a.row_oldest = h->row_oldest; a.frames_appended = h->frames_appended; a.frame_capacity = h->cfg.frame_capacity;
    const int64_t items = (int64_t)batch * a.ncols;
    if (items <= 0) return RETH_OK;
    const int64_t maxg = (int64_t)h->n_sms * 6 * 64;
    frame_gather_kernel<<<(unsigned)(items < maxg ? items : maxg), kCopyThreads, 0, s>>>(a);
    CK(h, cudaGetLastError());
    h->st.kernel_launches++;
    return RETH_OK;
}


// TMA bulk-copy gather (row_gather_bulk_kernel): persistent, one CTA per SM.
int launch_gather_bulk(reth_replay *h, uint8_t *const *batch_cols, const int64_t *idx, int batch,
                       const int64_t *meta, cudaStream_t s)
{
    static const int units_cap = env_int("RETH_BULK_UNITS", 3584);     // 57 344 B: >= 4 stages per SM
    static const int ctas_per_sm = env_int("RETH_BULK_CTAS", 1);
    static const int ef_loads = env_int("RETH_BULK_EF_LOADS", 1), ef_stores = env_int("RETH_BULK_EF_STORES", 0);
    BulkArgs b = {};
    build_row_args(h, b.r, batch_cols, idx, batch, units_cap);
    b.r.meta = meta;
    int max_units = 0;
    for (int c = 0; c < b.r.ncols; ++c)
        if (b.r.col[c].chunks > 0 && b.r.col[c].units > max_units) max_units = b.r.col[c].units;
    b.stage_bytes = (max_units * 16 + 127) & ~127;
    const int fixed = 128 + kBulkSlotCache * 8;
    const int budget = h->max_dyn_smem / ctas_per_sm - (ctas_per_sm > 1 ? 1024 : 0) - fixed;
    int stages = b.stage_bytes > 0 ? budget / b.stage_bytes : 2;
    if (stages > 8) stages = 8;
    if (stages < 2) return fail(h, RETH_ERR_INVALID, "bulk gather: a %d-byte stage does not fit shared memory", b.stage_bytes);
    b.stages = stages;
    b.evict_first_loads = ef_loads;
    b.evict_first_stores = ef_stores;
    const size_t smem = (size_t)stages * b.stage_bytes + fixed;
    int64_t grid = (int64_t)h->n_sms * ctas_per_sm;
    const int64_t work = b.r.wide_items > 0 ? b.r.wide_items : 1;
    if (grid > work) grid = work;
    if (h->bulk_grid_cap > 0 && grid > h->bulk_grid_cap) grid = h->bulk_grid_cap;
    row_gather_bulk_kernel<<<(unsigned)grid, kBulkThreads, smem, s>>>(b);
    CK(h, cudaGetLastError());
    h->st.kernel_launches++;
    return RETH_OK;
}

int do_sample(reth_replay *h, int batch, const double *uniforms, int uniforms_loc, uint64_t seed,
              double beta, int slot, const double *all_roots, int n_shards, int shard, cudaStream_t s,
              int groups = 1)
{
    NvtxRange nvtx("reth:draw");
    if (slot_ok(h, slot)) return RETH_ERR_INVALID;
    if (batch < 1 || batch > h->cfg.max_batch)
        return fail(h, RETH_ERR_INVALID, "batch %d outside [1, max_batch=%d]", batch, h->cfg.max_batch);
    if (groups < 1 || slot + groups > (int)h->slots.size())
        return fail(h, RETH_ERR_INVALID, "slots %d..%d outside the batch ring", slot, slot + groups - 1);
    if (groups > 1 && (batch != h->cfg.max_batch || (uniforms && uniforms_loc != RETH_DEVICE)))
        return fail(h, RETH_ERR_INVALID, "multi-slot sampling: batch == max_batch, device uniforms");
    if (groups > 32 && n_shards > 1) return fail(h, RETH_ERR_INVALID, "Mode G: at most 32 draws per launch");
    BatchSlot &b = h->slots[slot];
    SampleArgs a = {};
    a.g = h->geom; a.desc = h->desc; a.agg = h->agg; a.batch = batch;
    if (uniforms) {
        const void *u = nullptr;
        int rc = to_device(h, uniforms, uniforms_loc, (size_t)batch * groups * 8, h->stage_u, s, &u);
        if (rc) return rc;
        a.uniforms = (const double *)u;
    }
    a.seed = seed; a.draw_base = h->draw_counter;
    h->draw_counter += (uint64_t)batch * groups;
    a.groups = groups; a.out_stride = h->cfg.max_batch;
    a.beta = beta; a.min_mode = h->cfg.min_mode;
    a.idx_out = b.idx; a.w_out = b.w; a.prio_out = b.prio; a.meta = b.meta;
    a.idx_mirror = h->mirror_idx; a.w_mirror = h->mirror_w; a.err_mirror = h->mirror_err; a.err = h->err;
    a.all_roots = all_roots; a.n_shards = n_shards; a.shard = shard;
    if (h->wait_seq && n_shards > 1) {
        a.wait_board = h->board; a.wait_seq = h->wait_seq; a.roots_out = h->all_roots; a.err_rw = h->err;
        a.wait_timeout_ns = (unsigned long long)env_int("RETH_BOARD_TIMEOUT_MS", 2000) * 1000000ull;
    }
    const bool with_quota = n_shards > 1 && (groups > 1 || h->quota_first);
    if (with_quota) {
        // the shares of the `groups` draws are computed first (and packed back to back from slot `slot` on): the
        // draw kernel is then launched for about this shard's share instead of the whole global batch
        // (both buffers exist since create / the last ring resize: no allocation -- and no legacy-stream memset,
        // which would wait for every blocking stream of the process -- on the sampling path)
        global_quota_kernel<<<1, 32 * groups, 0, s>>>(a, h->quota, h->slot_quota + 4 * slot);
        CK(h, cudaGetLastError());
        h->st.kernel_launches++;
        a.quota = h->quota;
    }
    const int wpb = 4;
    int64_t warps = (int64_t)batch * groups;
    if (with_quota) {
        // stratified targets: the share is this shard's fraction of the global total +- 1; slack for unequal
        // shards, beyond which the warps stride
        int64_t wpd = (int64_t)(1.25 * (double)batch / n_shards) + 32;
        if (wpd > batch) wpd = batch;
        a.warps_per_draw = (int32_t)wpd;
        warps = wpd * groups;
    }
    if (h->pdl) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)((warps + wpb - 1) / wpb));
        cfg.blockDim = dim3(wpb * 32);
        cfg.stream = s;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        CK(h, cudaLaunchKernelEx(&cfg, per_sample_kernel, (const SampleArgs)a));
    } else {
        per_sample_kernel<<<(unsigned)((warps + wpb - 1) / wpb), wpb * 32, 0, s>>>(a);
    }
    CK(h, cudaGetLastError());
    h->st.kernel_launches++;
    h->st.sampled_rows += (int64_t)batch * groups;
    for (int g = 0; g < groups; ++g) {
        h->slots[slot + g].batch = batch;
        h->slots[slot + g].partial = n_shards > 1;
        h->slots[slot + g].shards = n_shards;
        h->slots[slot + g].packed = (n_shards > 1 && groups > 1 && g == 0) ? groups : 0;
    }
    return RETH_OK;
}

// groups > 1: slots slot..slot+groups-1 hold full batches (batch == max_batch), i.e. one
// contiguous groups*batch index array and one contiguous output per column
int do_gather(reth_replay *h, int slot, bool use_meta, cudaStream_t s, int groups = 1)
{
    NvtxRange nvtx("reth:gather");
    BatchSlot &b = h->slots[slot];
    if (!h->cfg.store_rows || h->cfg.n_columns == 0) return RETH_OK;
    const int64_t *meta = (use_meta && (groups == 1 || b.packed == groups)) ? b.meta : nullptr;
    int rc;
    if (h->gather_bulk) {
        rc = launch_gather_bulk(h, b.cols, b.idx, b.batch * groups, meta, s);
    } else {
        RowArgs a;
        build_row_args(h, a, b.cols, b.idx, b.batch * groups, 0, true);
        a.meta = meta;
        a.share_div = b.partial ? b.shards : 0;
        rc = launch_row_move<false>(h, a, s);
        if (!rc) rc = launch_frame_gather(h, b.cols, b.idx, b.batch * groups, meta, 0, s);
    }
    if (rc) return rc;
    h->st.gathered_rows += (int64_t)b.batch * groups;
    return RETH_OK;
}

int group_ok(reth_replay *h, int first_slot, int n_slots)
{
    if (n_slots < 1 || first_slot < 0 || first_slot + n_slots > (int)h->slots.size())
        return fail(h, RETH_ERR_INVALID, "slots %d..%d outside the batch ring", first_slot, first_slot + n_slots - 1);
    for (int g = 0; g < n_slots; ++g) {
        const BatchSlot &b = h->slots[first_slot + g];
        if (b.batch <= 0) return fail(h, RETH_ERR_STATE, "slot %d holds no sampled batch", first_slot + g);
        if (n_slots > 1 && b.batch != h->cfg.max_batch)
            return fail(h, RETH_ERR_STATE, "multi-slot calls need batch == max_batch");
    }
    return RETH_OK;
}

}  // namespace

RETH_EXPORT int reth_replay_sample(reth_replay_t *h, int batch, const double *uniforms, int uniforms_loc,
                                   uint64_t seed, double beta, int slot, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    return do_sample(h, batch, uniforms, uniforms_loc, seed, beta, slot, nullptr, 1, 0, S(stream));
}

RETH_EXPORT int reth_replay_sample_global(reth_replay_t *h, int global_batch, const double *all_roots_dev,
                                          int n_shards, int shard, const double *uniforms, int uniforms_loc,
                                          uint64_t seed, double beta, int slot, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (!all_roots_dev || n_shards < 1 || shard < 0 || shard >= n_shards)
        return fail(h, RETH_ERR_INVALID, "bad shard arguments");
    DeviceGuard guard(h->cfg.device);
    // n_shards == 1 still takes the global path when asked (keeps one code path testable)
    return do_sample(h, global_batch, uniforms, uniforms_loc, seed, beta, slot, all_roots_dev,
                     n_shards > 1 ? n_shards : 1, shard, S(stream));
}

RETH_EXPORT int reth_replay_sample_global_multi(reth_replay_t *h, int global_batch, const double *all_roots_dev,
                                                int n_shards, int shard, int first_slot, int n_slots,
                                                const double *uniforms_dev, uint64_t seed, double beta, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (!all_roots_dev || n_shards < 2 || shard < 0 || shard >= n_shards || n_slots < 1)
        return fail(h, RETH_ERR_INVALID, "bad shard arguments");
    DeviceGuard guard(h->cfg.device);
    return do_sample(h, global_batch, uniforms_dev, RETH_DEVICE, seed, beta, first_slot, all_roots_dev,
                     n_shards, shard, S(stream), n_slots);
}

RETH_EXPORT int reth_replay_slot_quota(reth_replay_t *h, int first_slot, int n_slots, int64_t *out, void *stream)
{
    if (!h || !out) return RETH_ERR_INVALID;
    if (n_slots < 1 || first_slot < 0 || first_slot + n_slots > (int)h->slots.size())
        return fail(h, RETH_ERR_INVALID, "slots %d..%d outside the batch ring", first_slot, first_slot + n_slots - 1);
    DeviceGuard guard(h->cfg.device);
    CK(h, cudaMemcpyAsync(out, h->slot_quota + 4 * first_slot, (size_t)n_slots * 32, cudaMemcpyDeviceToHost, S(stream)));
    CK(h, cudaStreamSynchronize(S(stream)));
    return RETH_OK;
}

RETH_EXPORT int reth_replay_gather(reth_replay_t *h, int slot, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (slot_ok(h, slot)) return RETH_ERR_INVALID;
    if (h->slots[slot].batch <= 0) return fail(h, RETH_ERR_STATE, "slot %d holds no sampled batch", slot);
    DeviceGuard guard(h->cfg.device);
    return do_gather(h, slot, true, S(stream));
}

RETH_EXPORT int reth_replay_sample_multi(reth_replay_t *h, int batch, int first_slot, int n_slots,
                                         const double *uniforms_dev, uint64_t seed, double beta, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    return do_sample(h, batch, uniforms_dev, RETH_DEVICE, seed, beta, first_slot, nullptr, 1, 0, S(stream), n_slots);
}

RETH_EXPORT int reth_replay_gather_multi(reth_replay_t *h, int first_slot, int n_slots, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (group_ok(h, first_slot, n_slots)) return h->error.find("holds no") != std::string::npos ? RETH_ERR_STATE : RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    return do_gather(h, first_slot, true, S(stream), n_slots);
}

RETH_EXPORT int reth_replay_update_slots(reth_replay_t *h, int first_slot, int n_slots, const void *prio,
                                         int prio_dtype, int prio_loc, double alpha, int transform, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (group_ok(h, first_slot, n_slots)) return h->error.find("holds no") != std::string::npos ? RETH_ERR_STATE : RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    const int64_t n = (int64_t)h->slots[first_slot].batch * n_slots;
    if (n > h->max_n && prio && prio_loc != RETH_DEVICE)       // host priorities pass through the staging buffer
        return fail(h, RETH_ERR_INVALID, "group of %lld rows with host priorities exceeds %lld", (long long)n, (long long)h->max_n);
    UpdateArgs a = {};
    a.n = n; a.idx = h->slots[first_slot].idx; a.idx_kind = 0;
    a.alpha = alpha; a.transform = transform;
    // one draw: duplicates adjacent; several draws: the same slot may recur anywhere -> stamps
    a.dedupe = n_slots == 1 ? 2 : 1;
    if (h->slots[first_slot].packed == n_slots || (n_slots == 1 && h->slots[first_slot].partial))
        a.n_dev = h->slots[first_slot].meta + 1;          // Mode G: the share count lives on the device
    int rc = set_prio(h, a, prio, prio_dtype, prio_loc, 0, n, S(stream));
    if (rc) return rc;
    rc = launch_update(h, a, S(stream), a.n_dev ? n / h->slots[first_slot].shards + 1 : 0);
    if (rc) return rc;
    h->st.updated_rows += n;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_append_at(reth_replay_t *h, const int64_t *slots, int slots_loc,
                                      const void *const *cols, int cols_loc,
                                      const void *prio, int prio_dtype, int prio_loc, int64_t n,
                                      double alpha, int transform, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (n < 0 || (n > 0 && !slots)) return fail(h, RETH_ERR_INVALID, "bad slots");
    if (n > h->max_n) return fail(h, RETH_ERR_INVALID, "append_at: n > %lld", (long long)h->max_n);
    if (n == 0) return RETH_OK;
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    UpdateArgs a = {};
    a.n = n; a.idx_kind = 0; a.alpha = alpha; a.transform = transform; a.dedupe = 1;
    a.keep_stamps = 1;
    int rc = to_device(h, slots, slots_loc, n * 8, h->stage_idx, s, &a.idx);
    if (rc) return rc;
    rc = set_prio(h, a, prio, prio_dtype, prio_loc, 0, n, s);
    if (rc) return rc;
    rc = launch_update(h, a, s);   // also arbitrates duplicates: stamp[slot] == 1+i for the winner
    if (rc) return rc;
    if (h->cfg.store_rows && h->cfg.n_columns > 0) {
        if (!cols) return fail(h, RETH_ERR_INVALID, "cols is null");
        if (cols_loc != RETH_DEVICE)
            return fail(h, RETH_ERR_INVALID, "append_at needs device-resident rows (stage them first)");
        uint8_t *src[kMaxCols];
        for (int c = 0; c < h->cfg.n_columns; ++c) src[c] = (uint8_t *)cols[c];
        { const int rcw = wait_gathers(h, s); if (rcw) return rcw; }   // no gather may read a slot while it is rewritten
        RowArgs r;
        build_row_args(h, r, src, (const int64_t *)a.idx, (int)n);
        r.stamp = h->stamp;
        rc = launch_row_move<true>(h, r, s);
        if (rc) return rc;
        rc = mark_frame_refs(h, (const int64_t *)a.idx, 0, n, s);
        if (rc) return rc;
    }
    stamp_clear_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(h->stamp, (const int64_t *)a.idx, n,
                                                                  h->cfg.capacity);
    CK(h, cudaGetLastError());
    h->st.kernel_launches++;
    h->st.appended_rows += n;
    h->size = (h->size + n < h->cfg.capacity) ? h->size + n : h->cfg.capacity;
    h->st.size = h->size;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_slot_ptrs(reth_replay_t *h, int slot, void **cols_out, int64_t **indices_out,
                                      double **weights_out, double **prios_out, int32_t *batch_out)
{
    if (!h) return RETH_ERR_INVALID;
    if (slot_ok(h, slot)) return RETH_ERR_INVALID;
    BatchSlot &b = h->slots[slot];
    if (cols_out) for (int c = 0; c < h->cfg.n_columns; ++c) cols_out[c] = b.cols[c];
    if (indices_out) *indices_out = b.idx;
    if (weights_out) *weights_out = b.w;
    if (prios_out) *prios_out = b.prio;
    if (batch_out) *batch_out = b.batch;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_slot_read(reth_replay_t *h, int slot, int64_t *indices, double *weights,
                                      void *const *cols, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (slot_ok(h, slot)) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    BatchSlot &b = h->slots[slot];
    const size_t n = (size_t)b.batch;
    if (indices) CK(h, cudaMemcpyAsync(indices, b.idx, n * 8, cudaMemcpyDeviceToHost, s));
    if (weights) CK(h, cudaMemcpyAsync(weights, b.w, n * 8, cudaMemcpyDeviceToHost, s));
    if (cols && h->cfg.store_rows)
        for (int c = 0; c < h->cfg.n_columns; ++c)
            if (cols[c])
                CK(h, cudaMemcpyAsync(cols[c], b.cols[c], n * h->out_rb[c], cudaMemcpyDeviceToHost, s));
    return check_device_err(h, s);
}

RETH_EXPORT int reth_replay_group_read(reth_replay_t *h, int first_slot, int n_slots, int64_t *indices,
                                       double *weights, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (group_ok(h, first_slot, n_slots)) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    const size_t n = (size_t)h->slots[first_slot].batch * n_slots;
    if (indices) CK(h, cudaMemcpyAsync(indices, h->slots[first_slot].idx, n * 8, cudaMemcpyDeviceToHost, s));
    if (weights) CK(h, cudaMemcpyAsync(weights, h->slots[first_slot].w, n * 8, cudaMemcpyDeviceToHost, s));
    return check_device_err(h, s);
}

RETH_EXPORT int reth_replay_slot_meta(reth_replay_t *h, int slot, int64_t *first_count, void *stream)
{
    if (!h || !first_count) return RETH_ERR_INVALID;
    if (slot_ok(h, slot)) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    CK(h, cudaMemcpyAsync(h->host_pin + 8, h->slots[slot].meta, 16, cudaMemcpyDeviceToHost, S(stream)));
    CK(h, cudaStreamSynchronize(S(stream)));
    first_count[0] = h->host_pin[8];
    first_count[1] = h->host_pin[9];
    return RETH_OK;
}

RETH_EXPORT int reth_replay_tree_read(reth_replay_t *h, double *sum, double *mn, double *val, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    const int64_t cap = h->cfg.capacity;
    double *d[3] = {nullptr, nullptr, nullptr};
    double *host[3] = {sum, mn, val};
    for (int k = 0; k < 3; ++k)
        if (host[k]) CK(h, cudaMalloc((void **)&d[k], (size_t)cap * 8));
    tree_read_kernel<<<(unsigned)((cap + 255) / 256), 256, 0, s>>>(h->geom, h->desc, h->agg, d[0], d[1], d[2]);
    int rc = RETH_OK;
    if (cudaGetLastError() != cudaSuccess) rc = fail(h, RETH_ERR_CUDA, "tree_read_kernel launch failed");
    for (int k = 0; k < 3 && rc == RETH_OK; ++k)
        if (host[k] && cudaMemcpyAsync(host[k], d[k], (size_t)cap * 8, cudaMemcpyDeviceToHost, s) != cudaSuccess)
            rc = fail(h, RETH_ERR_CUDA, "tree_read copy failed");
    cudaStreamSynchronize(s);
    for (int k = 0; k < 3; ++k) cudaFree(d[k]);
    if (rc) return rc;
    return check_device_err(h, s);
}

// ---- tree_read without allocation (no counterpart in the reference).
// reth_replay_tree_read() above takes its device scratch (three capacity-sized arrays in heap order) from cudaMalloc
// on every call; callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid it)
// get the same kernel writing into a file-scope device array instead.  The array is shared by every call on a GPU and
// keeps the last call's contents: tree_read_kernel writes every element [0, capacity) of each requested array, so
// nothing has to start at zero; two calls that overlap on the same GPU race on it.
constexpr int64_t kNoAllocMaxCapacity = RETH_NOALLOC_MAX_CAPACITY;
static __device__ __align__(256) double g_tree_scratch[3 * kNoAllocMaxCapacity];     // 1.5 MB on every GPU that loads the library

namespace {
// The array's name is a host-side stand-in in host code, not a device address: resolve it once per device, with that
// device current, and hand the kernel the resolved pointer.
int tree_scratch_ptr(reth_replay *h, double **out)
{
    static double *ptr[64] = {};
    const int dev = h->cfg.device;
    if (dev < 0 || dev >= 64) return fail(h, RETH_ERR_INVALID, "device ordinal %d out of range", dev);
    if (!ptr[dev]) {
        void *p = nullptr;
        CK(h, cudaGetSymbolAddress(&p, g_tree_scratch));      // the caller holds a DeviceGuard for h->cfg.device
        if (!p) return fail(h, RETH_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address");
        ptr[dev] = (double *)p;
    }
    *out = ptr[dev];
    return RETH_OK;
}
}  // namespace

RETH_EXPORT int reth_replay_tree_read_noalloc(reth_replay_t *h, double *sum, double *mn, double *val, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    const int64_t cap = h->cfg.capacity;
    if (cap > kNoAllocMaxCapacity)
        return fail(h, RETH_ERR_INVALID, "tree_read_noalloc: capacity %lld exceeds the static scratch (at most %lld nodes); "
                                         "use reth_replay_tree_read", (long long)cap, (long long)kNoAllocMaxCapacity);
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    double *base = nullptr;
    int rc = tree_scratch_ptr(h, &base);
    if (rc) return rc;
    double *host[3] = {sum, mn, val};
    double *d[3] = {nullptr, nullptr, nullptr};
    for (int k = 0; k < 3; ++k)
        if (host[k]) d[k] = base + k * kNoAllocMaxCapacity;
    tree_read_kernel<<<(unsigned)((cap + 255) / 256), 256, 0, s>>>(h->geom, h->desc, h->agg, d[0], d[1], d[2]);
    CK(h, cudaGetLastError());
    for (int k = 0; k < 3; ++k)
        if (host[k]) CK(h, cudaMemcpyAsync(host[k], d[k], (size_t)cap * 8, cudaMemcpyDeviceToHost, s));
    return check_device_err(h, s);                             // synchronises the stream
}

RETH_EXPORT int reth_replay_tree_root(reth_replay_t *h, double *total, double *minimum, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    double *r = (double *)(h->host_pin + 16);
    CK(h, cudaMemcpyAsync(r, &h->agg[1], 16, cudaMemcpyDeviceToHost, S(stream)));
    CK(h, cudaStreamSynchronize(S(stream)));
    if (total) *total = r[0];
    if (minimum) {
        if (h->cfg.min_mode == RETH_MIN_SERVICE) *minimum = (r[0] != 0.0) ? r[1] : 1.0;
        else *minimum = (r[1] != 0.0) ? r[1] : 1.0;
    }
    return RETH_OK;
}

RETH_EXPORT int reth_replay_root_dev(reth_replay_t *h, double **root_dev_out)
{
    if (!h || !root_dev_out) return RETH_ERR_INVALID;
    *root_dev_out = (double *)&h->agg[1];
    return RETH_OK;
}

RETH_EXPORT int reth_replay_ring_read(reth_replay_t *h, int column, int64_t first_slot, int64_t n,
                                      void *out_host, void *stream)
{
    if (!h || !out_host) return RETH_ERR_INVALID;
    if (!h->cfg.store_rows || column < 0 || column >= h->cfg.n_columns)
        return fail(h, RETH_ERR_INVALID, "bad column");
    if (first_slot < 0 || n < 0 || first_slot + n > h->cfg.capacity)
        return fail(h, RETH_ERR_INVALID, "slot range outside the ring");
    DeviceGuard guard(h->cfg.device);
    const int64_t rb = h->cfg.row_bytes[column];
    CK(h, cudaMemcpyAsync(out_host, h->ring[column] + first_slot * rb, (size_t)(n * rb),
                          cudaMemcpyDeviceToHost, S(stream)));
    CK(h, cudaStreamSynchronize(S(stream)));
    return RETH_OK;
}

RETH_EXPORT int reth_replay_stats_get(reth_replay_t *h, reth_replay_stats *out)
{
    if (!h || !out) return RETH_ERR_INVALID;
    h->st.tail = h->tail;
    h->st.size = h->size;
    *out = h->st;
    return RETH_OK;
}


// ------------------------------------------------------- non-PER sampler plugins --
namespace {

// entries [skip, n) of (idx, w) -> dst[(dst_pos + i) % dst_cap]; host arrays go through the staging buffers
int note_entries(reth_replay *h, const void *idx, int idx_dtype, int idx_loc, int64_t first_slot, const void *w,
                 int w_dtype, int w_loc, int64_t n, int64_t skip, int64_t *dst_idx, double *dst_w, int64_t dst_pos,
                 int64_t dst_cap, cudaStream_t s)
{
    if (idx && idx_dtype != RETH_I32 && idx_dtype != RETH_I64) return fail(h, RETH_ERR_INVALID, "indices must be int32 or int64");
    if (w && w_dtype != RETH_F32 && w_dtype != RETH_F64) return fail(h, RETH_ERR_INVALID, "weights must be float32 or float64");
    const size_t is = idx ? dtype_size(idx_dtype) : 0, ws = w ? dtype_size(w_dtype) : 0;
    for (int64_t off = skip; off < n; off += h->max_n) {
        const int64_t m = (n - off < h->max_n) ? n - off : h->max_n;
        NoteArgs a = {};
        a.capacity = h->cfg.capacity;
        a.n = m; a.skip = 0;
        a.dst_idx = dst_idx; a.dst_w = w ? dst_w : nullptr;
        a.dst_pos = (dst_pos + off) % dst_cap; a.dst_cap = dst_cap;
        if (idx) {
            a.idx_kind = idx_dtype == RETH_I64 ? 0 : 1;
            int rc = to_device(h, (const uint8_t *)idx + off * is, idx_loc, m * is, h->stage_idx, s, &a.idx);
            if (rc) return rc;
        } else {
            a.idx_kind = 2; a.iota_base = first_slot + off;
        }
        a.w_kind = 2;
        if (w) {
            a.w_kind = w_dtype == RETH_F32 ? 0 : 1;
            int rc = to_device(h, (const uint8_t *)w + off * ws, w_loc, m * ws, h->stage_prio, s, &a.w);
            if (rc) return rc;
        }
        sampler_note_kernel<<<(unsigned)((m + 255) / 256), 256, 0, s>>>(a);
        CK(h, cudaGetLastError());
        h->st.kernel_launches++;
    }
    return RETH_OK;
}

int pick_into_slots(reth_replay *h, PickArgs &a, int batch, int first_slot, int n_slots, cudaStream_t s)
{
    BatchSlot &b = h->slots[first_slot];
    a.n = (int64_t)batch * n_slots;
    a.idx_out = b.idx; a.w_out = b.w; a.prio_out = b.prio; a.meta = b.meta; a.batch = batch; a.err = h->err;
    a.idx_mirror = h->mirror_idx; a.w_mirror = h->mirror_w; a.err_mirror = h->mirror_err;
    sampler_pick_kernel<<<(unsigned)((a.n + 255) / 256), 256, 0, s>>>(a);
    CK(h, cudaGetLastError());
    h->st.kernel_launches++;
    h->st.sampled_rows += a.n;
    for (int g = 0; g < n_slots; ++g) { h->slots[first_slot + g].batch = batch; h->slots[first_slot + g].partial = false; }
    return RETH_OK;
}

int slots_ok(reth_replay *h, int batch, int first_slot, int n_slots)
{
    if (batch < 1 || batch > h->cfg.max_batch) return fail(h, RETH_ERR_INVALID, "batch %d outside [1, max_batch=%d]", batch, h->cfg.max_batch);
    if (n_slots < 1 || first_slot < 0 || first_slot + n_slots > (int)h->slots.size())
        return fail(h, RETH_ERR_INVALID, "slots %d..%d outside the batch ring", first_slot, first_slot + n_slots - 1);
    if (n_slots > 1 && batch != h->cfg.max_batch) return fail(h, RETH_ERR_INVALID, "multi-slot draws need batch == max_batch");
    return RETH_OK;
}

int do_sample_uniform(reth_replay *h, int batch, int first_slot, int n_slots, const int64_t *draws, int draws_loc,
                      uint64_t seed, cudaStream_t s)
{
    if (slots_ok(h, batch, first_slot, n_slots)) return RETH_ERR_INVALID;
    if (h->uni_tail <= 0) return fail(h, RETH_ERR_STATE, "uniform sampler: nothing has been notified yet");
    PickArgs a = {};
    a.uniform = 1; a.src_idx = h->uni_log; a.tail = h->uni_tail; a.seed = seed; a.draw_base = h->draw_counter;
    h->draw_counter += (uint64_t)batch * n_slots;
    if (draws) {
        if ((int64_t)batch * n_slots > h->max_n) return fail(h, RETH_ERR_INVALID, "too many explicit draws");
        const void *d = nullptr;
        int rc = to_device(h, draws, draws_loc, (size_t)batch * n_slots * 8, h->stage_idx, s, &d);
        if (rc) return rc;
        a.draws = (const int64_t *)d;
    }
    return pick_into_slots(h, a, batch, first_slot, n_slots, s);
}

int do_sample_fifo(reth_replay *h, int batch, int first_slot, int n_slots, cudaStream_t s)
{
    if (slots_ok(h, batch, first_slot, n_slots)) return RETH_ERR_INVALID;
    const int64_t n = (int64_t)batch * n_slots;
    if (!(h->fifo_len > n)) return fail(h, RETH_ERR_STATE, "fifo sampler: %lld queued, need more than %lld", (long long)h->fifo_len, (long long)n);
    PickArgs a = {};
    a.uniform = 0; a.src_idx = h->fifo_idx; a.src_w = h->fifo_w; a.src_cap = h->cfg.capacity; a.head = h->fifo_head;
    int rc = pick_into_slots(h, a, batch, first_slot, n_slots, s);
    if (rc) return rc;
    h->fifo_head = (h->fifo_head + n) % h->cfg.capacity;
    h->fifo_len -= n;
    return RETH_OK;
}

}  // namespace

RETH_EXPORT int reth_replay_uniform_note(reth_replay_t *h, const void *idx, int idx_dtype, int idx_loc,
                                         int64_t first_slot, int64_t n, void *stream)
{
    if (!h || n < 0) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    const int64_t cap = h->cfg.capacity;
    if (!h->uni_log) CK(h, dalloc_zero(&h->uni_log, (size_t)cap, &h->st.hbm_bytes));
    if (h->uni_tail >= cap) return RETH_OK;                       // uniform_sampler.py:18-19
    const int64_t m = n < cap - h->uni_tail ? n : cap - h->uni_tail;
    int rc = note_entries(h, idx, idx_dtype, idx_loc, first_slot, nullptr, 0, 0, m, 0, h->uni_log, nullptr,
                          h->uni_tail, cap + 1, S(stream));     // dst_cap > any position: no wrap
    if (rc) return rc;
    h->uni_tail += m;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_sample_uniform(reth_replay_t *h, int batch, int first_slot, int n_slots,
                                           const int64_t *draws, int draws_loc, uint64_t seed, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    return do_sample_uniform(h, batch, first_slot, n_slots, draws, draws_loc, seed, S(stream));
}

RETH_EXPORT int reth_replay_fifo_push(reth_replay_t *h, const void *idx, int idx_dtype, int idx_loc, int64_t first_slot,
                                      const void *w, int w_dtype, int w_loc, int64_t n, void *stream)
{
    if (!h || n < 0 || !w) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    const int64_t cap = h->cfg.capacity;
    if (!h->fifo_idx) {
        CK(h, dalloc_zero(&h->fifo_idx, (size_t)cap, &h->st.hbm_bytes));
        CK(h, dalloc_zero(&h->fifo_w, (size_t)cap, &h->st.hbm_bytes));
    }
    // deque(maxlen=capacity).appendleft: a full queue drops its oldest entry; a push longer than the
    // queue keeps only its last `capacity` entries
    const int64_t skip = n > cap ? n - cap : 0;
    int rc = note_entries(h, idx, idx_dtype, idx_loc, first_slot, w, w_dtype, w_loc, n, skip, h->fifo_idx, h->fifo_w,
                          (h->fifo_head + h->fifo_len) % cap, cap, S(stream));
    if (rc) return rc;
    const int64_t total = h->fifo_len + n;
    if (total > cap) { h->fifo_head = (h->fifo_head + (total - cap)) % cap; h->fifo_len = cap; }
    else h->fifo_len = total;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_sample_fifo(reth_replay_t *h, int batch, int first_slot, int n_slots, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    return do_sample_fifo(h, batch, first_slot, n_slots, S(stream));
}

RETH_EXPORT int reth_replay_sampler_state(reth_replay_t *h, int64_t *out)
{
    if (!h || !out) return RETH_ERR_INVALID;
    out[0] = h->uni_tail; out[1] = h->fifo_len; out[2] = h->fifo_head; out[3] = 0;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_pipe_sampler(reth_replay_t *h, int kind)
{
    if (!h || kind < 0 || kind > 2) return RETH_ERR_INVALID;
    h->pipe_kind = kind;
    return RETH_OK;
}

// ------------------------------------------------------- asynchronous pipeline --
namespace {

void pipe_free(reth_replay *h)
{
    auto &p = h->pipe;
    if (p.idx_host) cudaFreeHost(p.idx_host);
    if (p.w_host) cudaFreeHost(p.w_host);
    if (p.err_host) cudaFreeHost(p.err_host);
    if (p.upd_host) cudaFreeHost(p.upd_host);
    cudaFree(p.upd_dev);
    for (auto *v : {&p.ev_smp, &p.ev_gat, &p.ev_done, &p.ev_upd}) {
        for (auto e : *v) cudaEventDestroy(e);
        v->clear();
    }
    if (p.s_copy) cudaStreamDestroy(p.s_copy);
    forget_gathers(h);
    p = reth_replay::Pipe();
}

// give_up_lent: the caller hands the lent group back with this call (pipe_next), so its slots may
// be drawn into before the wait for the next delivery -- the device then never idles between groups.
int pipe_issue(reth_replay *h, double beta, bool give_up_lent = false)
{
    auto &p = h->pipe;
    if (!p.open) return fail(h, RETH_ERR_STATE, "no pipe is open");
    int64_t lent = p.waited < p.hold ? p.waited : p.hold;
    if (give_up_lent && lent == p.hold) lent -= 1;          // fewer delivered than `hold`: nothing to hand back yet
    if (p.issued - p.waited + lent >= p.n_groups)
        return fail(h, RETH_ERR_STATE, "all %d group slots are issued or lent", p.n_groups);
    const int g = (int)(p.issued % p.n_groups), slot = g * p.G;
    const size_t E = (size_t)p.E;
    int rc;
    if (p.zero_copy) { h->mirror_idx = p.idx_host + g * E; h->mirror_w = p.w_host + g * E; h->mirror_err = p.err_host + g; }
    if (h->pipe_kind == 1) rc = do_sample_uniform(h, p.batch, slot, p.G, nullptr, RETH_DEVICE, p.seed, h->s_tree);
    else if (h->pipe_kind == 2) rc = do_sample_fifo(h, p.batch, slot, p.G, h->s_tree);
    else rc = do_sample(h, p.batch, nullptr, RETH_DEVICE, p.seed, beta, slot, nullptr, 1, 0, h->s_tree, p.G);
    h->mirror_idx = nullptr; h->mirror_w = nullptr; h->mirror_err = nullptr;
    if (rc) return rc;
    CK(h, cudaEventRecord(p.ev_smp[g], h->s_tree));
    if (!p.zero_copy) {
        CK(h, cudaStreamWaitEvent(p.s_copy, p.ev_smp[g], 0));
        CK(h, cudaMemcpyAsync(p.idx_host + g * E, h->slots[slot].idx, E * 8, cudaMemcpyDeviceToHost, p.s_copy));
        CK(h, cudaMemcpyAsync(p.w_host + g * E, h->slots[slot].w, E * 8, cudaMemcpyDeviceToHost, p.s_copy));
        CK(h, cudaMemcpyAsync(p.err_host + g, h->err, sizeof(int32_t), cudaMemcpyDeviceToHost, p.s_copy));
    }
    cudaStream_t sg = (h->two_gather_streams && (p.issued & 1)) ? h->s_gather2 : h->s_gather;
    CK(h, cudaStreamWaitEvent(sg, p.ev_smp[g], 0));
    rc = do_gather(h, slot, false, sg, p.G);
    if (rc) return rc;
    CK(h, cudaEventRecord(p.ev_gat[g], sg));
    note_gather(h, p.ev_gat[g], sg);
    if (!p.zero_copy) {
        CK(h, cudaStreamWaitEvent(p.s_copy, p.ev_gat[g], 0));
        CK(h, cudaEventRecord(p.ev_done[g], p.s_copy));
    }
    p.issued++;
    return RETH_OK;
}

int pipe_wait(reth_replay *h, int32_t *group_out)
{
    auto &p = h->pipe;
    if (!p.open) return fail(h, RETH_ERR_STATE, "no pipe is open");
    if (p.issued <= p.waited) return fail(h, RETH_ERR_STATE, "pipe_wait: no group has been issued");
    const int g = (int)(p.waited % p.n_groups);
    // zero copy: the gather stream waited for the draw, so its event covers the indices in host memory too
    CK(h, cudaEventSynchronize(p.zero_copy ? p.ev_gat[g] : p.ev_done[g]));
    p.waited++;
    if (group_out) *group_out = g;
    if (p.err_host[g]) {
        cudaMemsetAsync(h->err, 0, sizeof(int32_t), h->s_tree);
        return device_err(h, p.err_host[g]);
    }
    return RETH_OK;
}

}  // namespace

RETH_EXPORT int reth_replay_pipe_open(reth_replay_t *h, int batch, int group, uint64_t seed, reth_pipe_info *out)
{
    if (!h) return RETH_ERR_INVALID;
    auto &p = h->pipe;
    if (p.open) return fail(h, RETH_ERR_STATE, "a pipe is already open");
    const int depth = (int)h->slots.size() - h->pipe_reserved;
    if (group < 1 || batch < 1 || batch > h->cfg.max_batch || (group > 1 && batch != h->cfg.max_batch))
        return fail(h, RETH_ERR_INVALID, "pipe: batch %d / group %d invalid (group > 1 needs batch == max_batch)", batch, group);
    if (depth / group < 2) return fail(h, RETH_ERR_INVALID, "pipe: ring_depth %d holds fewer than 2 groups of %d", depth, group);
    if ((int64_t)batch * group > h->max_n) return fail(h, RETH_ERR_INVALID, "pipe: group of %d rows exceeds %lld", batch * group, (long long)h->max_n);
    DeviceGuard guard(h->cfg.device);
    p.batch = batch; p.G = group; p.n_groups = depth / group; p.E = (int64_t)batch * group;
    p.upd_entries = 2 * p.n_groups + 2;
    p.seed = seed;
    const size_t E = (size_t)p.E;
    auto bail = [&](int code) { pipe_free(h); return code; };
#define CKP(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) return bail(fail(h, RETH_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e__))); } while (0)
    if (const char *e = getenv("RETH_PIPE_ZEROCOPY")) p.zero_copy = atoi(e) != 0;
    // mapped + portable: with unified addressing the device uses the same pointers
    const unsigned hf = cudaHostAllocMapped | cudaHostAllocPortable;
    CKP(cudaHostAlloc((void **)&p.idx_host, p.n_groups * E * 8, hf));
    CKP(cudaHostAlloc((void **)&p.w_host, p.n_groups * E * 8, hf));
    CKP(cudaHostAlloc((void **)&p.err_host, p.n_groups * sizeof(int32_t), hf));
    memset(p.err_host, 0, p.n_groups * sizeof(int32_t));
    CKP(cudaHostAlloc((void **)&p.upd_host, p.upd_entries * E * 16, hf));
    if (p.zero_copy) {
        void *dp = nullptr;
        if (cudaHostGetDevicePointer(&dp, p.idx_host, 0) != cudaSuccess || dp != (void *)p.idx_host) {
            cudaGetLastError();
            p.zero_copy = false;              // no unified addressing for pinned memory: copy engines instead
        }
    }
    CKP(cudaMalloc((void **)&p.upd_dev, E * 16));
    CKP(cudaStreamCreateWithFlags(&p.s_copy, cudaStreamNonBlocking));
    for (auto *v : {&p.ev_smp, &p.ev_gat, &p.ev_done})
        for (int i = 0; i < p.n_groups; ++i) {
            cudaEvent_t e;
            CKP(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            v->push_back(e);
        }
    for (int i = 0; i < p.upd_entries; ++i) {
        cudaEvent_t e;
        CKP(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        p.ev_upd.push_back(e);
    }
#undef CKP
    p.upd_used.assign(p.upd_entries, 0);
    p.issued = p.waited = p.upd_next = 0;
    p.open = true;
    if (out) {
        out->batch = batch; out->group = group; out->n_groups = p.n_groups; out->upd_entries = p.upd_entries;
        out->entry_elems = p.E; out->idx_host = p.idx_host; out->w_host = p.w_host; out->upd_host = p.upd_host;
    }
    return RETH_OK;
}

RETH_EXPORT int reth_replay_pipe_close(reth_replay_t *h)
{
    if (!h) return RETH_ERR_INVALID;
    if (!h->pipe.open) return RETH_OK;
    DeviceGuard guard(h->cfg.device);
    cudaStreamSynchronize(h->s_tree);
    cudaStreamSynchronize(h->s_gather);
    cudaStreamSynchronize(h->s_gather2);
    cudaStreamSynchronize(h->pipe.s_copy);
    pipe_free(h);
    return RETH_OK;
}

RETH_EXPORT int reth_replay_pipe_stream(reth_replay_t *h, void **stream_out)
{
    if (!h || !stream_out) return RETH_ERR_INVALID;
    *stream_out = (void *)h->s_tree;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_pipe_issue(reth_replay_t *h, double beta)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    return pipe_issue(h, beta);
}

RETH_EXPORT int reth_replay_pipe_wait(reth_replay_t *h, int32_t *group_out)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    return pipe_wait(h, group_out);
}

RETH_EXPORT int reth_replay_pipe_next(reth_replay_t *h, double beta, int32_t *group_out)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    auto &p = h->pipe;
    if (!p.open) return fail(h, RETH_ERR_STATE, "no pipe is open");
    if (p.issued <= p.waited) return fail(h, RETH_ERR_STATE, "pipe_next: no group has been issued");
    // the group lent so far is handed back by this call: draw into its slots FIRST, so that the new
    // draw + gather queue up behind the group in flight, and only then block for the delivery
    int rc = pipe_issue(h, beta, true);
    if (rc) return rc;
    return pipe_wait(h, group_out);
}

RETH_EXPORT int reth_replay_pipe_reserve(reth_replay_t *h, int n_slots)
{
    if (!h) return RETH_ERR_INVALID;
    if (h->pipe.open) return fail(h, RETH_ERR_STATE, "pipe_reserve must precede pipe_open");
    if (n_slots < 0 || n_slots > (int)h->slots.size() - 2) return fail(h, RETH_ERR_INVALID, "cannot reserve %d of %d batch slots", n_slots, (int)h->slots.size());
    h->pipe_reserved = n_slots;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_pipe_hold(reth_replay_t *h, int groups)
{
    if (!h) return RETH_ERR_INVALID;
    auto &p = h->pipe;
    if (!p.open) return fail(h, RETH_ERR_STATE, "no pipe is open");
    if (groups < 1 || groups >= p.n_groups) return fail(h, RETH_ERR_INVALID, "hold %d outside [1, n_groups - 1 = %d]", groups, p.n_groups - 1);
    if (p.issued) return fail(h, RETH_ERR_STATE, "pipe_hold must precede the first pipe_issue");
    p.hold = groups;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_pipe_advance(reth_replay_t *h, double beta)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    return pipe_issue(h, beta, true);
}

RETH_EXPORT int reth_replay_pipe_ready(reth_replay_t *h, int32_t *ready_out)
{
    if (!h || !ready_out) return RETH_ERR_INVALID;
    auto &p = h->pipe;
    if (!p.open) return fail(h, RETH_ERR_STATE, "no pipe is open");
    *ready_out = 0;
    if (p.issued <= p.waited) return RETH_OK;
    DeviceGuard guard(h->cfg.device);
    const int gq = (int)(p.waited % p.n_groups);
    const cudaError_t e = cudaEventQuery(p.zero_copy ? p.ev_gat[gq] : p.ev_done[gq]);
    if (e == cudaSuccess) *ready_out = 1;
    else if (e != cudaErrorNotReady) return fail(h, RETH_ERR_CUDA, "cudaEventQuery: %s", cudaGetErrorString(e));
    return RETH_OK;
}

RETH_EXPORT int reth_replay_pipe_update_begin(reth_replay_t *h, int32_t *entry_out)
{
    if (!h || !entry_out) return RETH_ERR_INVALID;
    auto &p = h->pipe;
    if (!p.open) return fail(h, RETH_ERR_STATE, "no pipe is open");
    const int e = (int)(p.upd_next % p.upd_entries);
    if (p.upd_used[e]) {
        DeviceGuard guard(h->cfg.device);
        CK(h, cudaEventSynchronize(p.ev_upd[e]));     // its previous host->device copy has drained
        p.upd_used[e] = 0;
    }
    *entry_out = e;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_pipe_update_commit(reth_replay_t *h, int entry, int64_t n, int prio_dtype, double alpha,
                                               int transform)
{
    if (!h) return RETH_ERR_INVALID;
    auto &p = h->pipe;
    if (!p.open) return fail(h, RETH_ERR_STATE, "no pipe is open");
    if (entry != (int)(p.upd_next % p.upd_entries)) return fail(h, RETH_ERR_STATE, "commit of entry %d out of order", entry);
    if (n < 0 || n > p.E) return fail(h, RETH_ERR_INVALID, "commit of %lld elements exceeds the entry (%lld)", (long long)n, (long long)p.E);
    if (prio_dtype != RETH_F32 && prio_dtype != RETH_F64) return fail(h, RETH_ERR_INVALID, "priorities must be float32 or float64");
    if (n == 0) return RETH_OK;
    DeviceGuard guard(h->cfg.device);
    const size_t E = (size_t)p.E, ps = dtype_size(prio_dtype);
    const uint8_t *src = p.upd_host + (size_t)entry * E * 16;
    const uint8_t *in = p.zero_copy ? src : p.upd_dev;   // zero copy: the kernel reads the staging entry itself
    if (!p.zero_copy) {
        if ((size_t)n == E) {
            CK(h, cudaMemcpyAsync(p.upd_dev, src, E * 8 + E * ps, cudaMemcpyHostToDevice, h->s_tree));
        } else {
            CK(h, cudaMemcpyAsync(p.upd_dev, src, (size_t)n * 8, cudaMemcpyHostToDevice, h->s_tree));
            CK(h, cudaMemcpyAsync(p.upd_dev + E * 8, src + E * 8, (size_t)n * ps, cudaMemcpyHostToDevice, h->s_tree));
        }
        CK(h, cudaEventRecord(p.ev_upd[entry], h->s_tree));
    }
    p.upd_used[entry] = 1;
    p.upd_next++;
    UpdateArgs a = {};
    a.n = n; a.idx = in; a.idx_kind = 0;
    a.prio = in + E * 8; a.prio_kind = prio_dtype == RETH_F32 ? 0 : 1;
    a.alpha = alpha; a.transform = transform; a.dedupe = 1;
    int rc = launch_update(h, a, h->s_tree);
    if (p.zero_copy) CK(h, cudaEventRecord(p.ev_upd[entry], h->s_tree));   // the entry is free once the kernel has read it
    if (rc) return rc;
    h->st.updated_rows += n;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_update_async(reth_replay_t *h, const int64_t *idx_host, const void *prio_host,
                                         int prio_dtype, int64_t n, double alpha, int transform)
{
    if (!h) return RETH_ERR_INVALID;
    if (!h->pipe.open) return fail(h, RETH_ERR_STATE, "no pipe is open");
    if (n < 0 || (n > 0 && (!idx_host || !prio_host))) return fail(h, RETH_ERR_INVALID, "bad arguments");
    if (prio_dtype != RETH_F32 && prio_dtype != RETH_F64) return fail(h, RETH_ERR_INVALID, "priorities must be float32 or float64");
    const size_t ps = dtype_size(prio_dtype);
    for (int64_t off = 0; off < n; off += h->pipe.E) {
        const int64_t m = n - off < h->pipe.E ? n - off : h->pipe.E;
        int32_t e = 0;
        int rc = reth_replay_pipe_update_begin(h, &e);
        if (rc) return rc;
        uint8_t *dst = h->pipe.upd_host + (size_t)e * h->pipe.E * 16;
        memcpy(dst, idx_host + off, (size_t)m * 8);
        memcpy(dst + h->pipe.E * 8, (const uint8_t *)prio_host + off * ps, (size_t)m * ps);
        rc = reth_replay_pipe_update_commit(h, e, m, prio_dtype, alpha, transform);
        if (rc) return rc;
    }
    return RETH_OK;
}

RETH_EXPORT int reth_replay_pipe_drain(reth_replay_t *h)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    CK(h, cudaStreamSynchronize(h->s_tree));
    CK(h, cudaStreamSynchronize(h->s_gather));
    CK(h, cudaStreamSynchronize(h->s_gather2));
    if (h->pipe.open) CK(h, cudaStreamSynchronize(h->pipe.s_copy));
    return RETH_OK;
}

// ------------------------------------------------------------ cross-process ring --
struct reth_ipc_view {
    reth_ipc_export ex{};
    int open_device = 0;          // device that was current when the ring was mapped
    uint8_t *cols[kMaxCols] = {};
    int64_t *idx = nullptr;
    double *w = nullptr;
};

RETH_EXPORT int reth_replay_ipc_export(reth_replay_t *h, reth_ipc_export *out)
{
    if (!h || !out) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    memset(out, 0, sizeof *out);
    out->n_columns = h->cfg.store_rows ? h->cfg.n_columns : 0;
    out->ring_depth = (int32_t)h->slots.size();
    out->max_batch = h->cfg.max_batch;
    out->device = h->cfg.device;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    for (int c = 0; c < out->n_columns; ++c) {
        out->row_bytes[c] = h->out_rb[c];               // rows of the BATCH ring
        CK(h, cudaIpcGetMemHandle((cudaIpcMemHandle_t *)out->col_handle[c], h->slot_cols[c]));
    }
    CK(h, cudaIpcGetMemHandle((cudaIpcMemHandle_t *)out->idx_handle, h->slot_idx));
    CK(h, cudaIpcGetMemHandle((cudaIpcMemHandle_t *)out->w_handle, h->slot_w));
    return RETH_OK;
}

RETH_EXPORT int reth_ipc_open(const reth_ipc_export *in, reth_ipc_view_t **out)
{
    if (!in || !out) return RETH_ERR_INVALID;
    *out = nullptr;
    if (in->n_columns < 0 || in->n_columns > RETH_MAX_COLUMNS) return fail(nullptr, RETH_ERR_INVALID, "bad descriptor");
    reth_ipc_view *v = new (std::nothrow) reth_ipc_view();
    if (!v) return fail(nullptr, RETH_ERR_OOM, "host allocation failed");
    v->ex = *in;
    // the ring is mapped for the CALLER's current device: when that is not the exporting GPU,
    // cudaIpcMemLazyEnablePeerAccess sets up the peer (NVLink) mapping
    if (cudaGetDevice(&v->open_device) != cudaSuccess) v->open_device = in->device;
    auto open1 = [&](const unsigned char *hb, void **p) {
        cudaIpcMemHandle_t mh;
        memcpy(&mh, hb, sizeof mh);
        return cudaIpcOpenMemHandle(p, mh, cudaIpcMemLazyEnablePeerAccess);
    };
    cudaError_t e = cudaSuccess;
    for (int c = 0; c < in->n_columns && e == cudaSuccess; ++c) e = open1(in->col_handle[c], (void **)&v->cols[c]);
    if (e == cudaSuccess) e = open1(in->idx_handle, (void **)&v->idx);
    if (e == cudaSuccess) e = open1(in->w_handle, (void **)&v->w);
    if (e != cudaSuccess) {
        fail(nullptr, RETH_ERR_CUDA, "cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
        cudaGetLastError();
        reth_ipc_close(v);
        return RETH_ERR_CUDA;
    }
    *out = v;
    return RETH_OK;
}

RETH_EXPORT int reth_ipc_slot_ptrs(reth_ipc_view_t *v, int slot, void **cols_out, int64_t **indices_out,
                                   double **weights_out)
{
    if (!v || slot < 0 || slot >= v->ex.ring_depth) return RETH_ERR_INVALID;
    const size_t mb = (size_t)v->ex.max_batch;
    if (cols_out)
        for (int c = 0; c < v->ex.n_columns; ++c) cols_out[c] = v->cols[c] + (size_t)slot * mb * v->ex.row_bytes[c];
    if (indices_out) *indices_out = v->idx + (size_t)slot * mb;
    if (weights_out) *weights_out = v->w + (size_t)slot * mb;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_gather_to(reth_replay_t *h, int slot, reth_ipc_view_t *dst, int dst_slot, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (slot_ok(h, slot)) return RETH_ERR_INVALID;
    BatchSlot &b = h->slots[slot];
    if (b.batch <= 0) return fail(h, RETH_ERR_STATE, "slot %d holds no sampled batch", slot);
    const int ncols = h->cfg.store_rows ? h->cfg.n_columns : 0;
    // dst == NULL: this shard's own ring (the learner rank's own slice of the global batch)
    const int d_depth = dst ? dst->ex.ring_depth : (int)h->slots.size();
    const int d_mb = dst ? dst->ex.max_batch : h->cfg.max_batch;
    if (dst_slot < 0 || dst_slot >= d_depth || d_mb < b.batch)
        return fail(h, RETH_ERR_INVALID, "destination slot %d does not hold a batch of %d", dst_slot, b.batch);
    if (!dst && dst_slot == slot) return fail(h, RETH_ERR_INVALID, "assembling into the slot being read");
    if (dst) {
        if (dst->ex.n_columns != ncols) return fail(h, RETH_ERR_INVALID, "destination ring has another row struct");
        for (int c = 0; c < ncols; ++c)
            if (dst->ex.row_bytes[c] != h->out_rb[c])
                return fail(h, RETH_ERR_INVALID, "destination ring has another row struct");
    }
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    const size_t mb = (size_t)d_mb;
    if (ncols > 0) {
        uint8_t *cols[kMaxCols] = {};
        for (int c = 0; c < ncols; ++c)
            cols[c] = (dst ? dst->cols[c] : h->slot_cols[c]) + (size_t)dst_slot * mb * h->out_rb[c];
        RowArgs a;
        build_row_args(h, a, cols, b.idx, b.batch, 0, true);
        a.meta = b.meta;
        a.dst_row_from_meta = 1;
        a.share_div = b.partial ? b.shards : 0;
        int rc = launch_row_move<false>(h, a, s);
        if (!rc) rc = launch_frame_gather(h, cols, b.idx, b.batch, b.meta, 1, s);
        if (rc) return rc;
        h->st.gathered_rows += b.batch;
    }
    slice_copy_kernel<<<(b.batch + 255) / 256, 256, 0, s>>>(b.idx, b.w, b.meta,
                                                          (dst ? dst->idx : h->slot_idx) + (size_t)dst_slot * mb,
                                                          (dst ? dst->w : h->slot_w) + (size_t)dst_slot * mb);
    CK(h, cudaGetLastError());
    h->st.kernel_launches++;
    if (!dst) h->slots[dst_slot].batch = b.batch;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_gather_from(reth_replay_t *h, int dst_slot, reth_replay_t *src, int src_slot, void *stream)
{
    if (!h || !src) return RETH_ERR_INVALID;
    if (slot_ok(h, dst_slot)) return RETH_ERR_INVALID;
    if (src_slot < 0 || src_slot >= (int)src->slots.size()) return fail(h, RETH_ERR_INVALID, "source slot %d outside the index shard's ring", src_slot);
    if (src->cfg.device != h->cfg.device || src->cfg.capacity != h->cfg.capacity)
        return fail(h, RETH_ERR_INVALID, "the index shard must live on the same device and span the same slots");
    const BatchSlot &b = src->slots[src_slot];
    if (b.batch <= 0) return fail(h, RETH_ERR_STATE, "source slot %d holds no sampled batch", src_slot);
    if (b.batch > h->cfg.max_batch) return fail(h, RETH_ERR_INVALID, "batch of %d exceeds max_batch %d", b.batch, h->cfg.max_batch);
    if (src == h && src_slot == dst_slot) return reth_replay_gather(h, dst_slot, stream);
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    BatchSlot &d = h->slots[dst_slot];
    if (h->cfg.store_rows && h->cfg.n_columns > 0) {
        RowArgs a;
        build_row_args(h, a, d.cols, b.idx, b.batch, 0, true);
        a.meta = b.meta;
        a.dst_row_from_meta = 1;
        a.share_div = b.partial ? b.shards : 0;
        int rc = launch_row_move<false>(h, a, s);
        if (!rc) rc = launch_frame_gather(h, d.cols, b.idx, b.batch, b.meta, 1, s);
        if (rc) return rc;
        h->st.gathered_rows += b.batch;
    }
    slice_copy_kernel<<<(b.batch + 255) / 256, 256, 0, s>>>(b.idx, b.w, b.meta, d.idx, d.w, d.meta);
    CK(h, cudaGetLastError());
    h->st.kernel_launches++;
    d.batch = b.batch;
    d.partial = b.partial;
    d.shards = b.shards;
    return RETH_OK;
}

RETH_EXPORT int reth_ipc_close(reth_ipc_view_t *v)
{
    if (!v) return RETH_OK;
    {
        DeviceGuard guard(v->open_device);
        for (int c = 0; c < kMaxCols; ++c)
            if (v->cols[c]) cudaIpcCloseMemHandle(v->cols[c]);
        if (v->idx) cudaIpcCloseMemHandle(v->idx);
        if (v->w) cudaIpcCloseMemHandle(v->w);
    }
    delete v;
    return RETH_OK;
}

// ------------------------------------------------------ shard roots over NVLink --
namespace {

int board_alloc(reth_replay *h)
{
    if (h->board) return RETH_OK;
    CK(h, dalloc_zero(&h->board, (size_t)kBoardEntries, &h->st.hbm_bytes));
    CK(h, dalloc_zero(&h->all_roots, (size_t)2 * kMaxShards, &h->st.hbm_bytes));
    return RETH_OK;
}

int launch_roots_exchange(reth_replay *h, int phases, cudaStream_t s)
{
    if (h->board_shards < 1) return fail(h, RETH_ERR_STATE, "roots exchange before reth_replay_board_connect");
    const int timeout_ms = env_int("RETH_BOARD_TIMEOUT_MS", 2000);
    if (phases & 1) h->board_seq++;
    if (h->board_seq == 0) return fail(h, RETH_ERR_STATE, "roots exchange: wait before the first publish");
    RootsArgs a = {};
    a.agg = h->agg;
    for (int i = 0; i < h->board_shards; ++i) a.board[i] = h->peer_board[i];
    a.all_roots = h->all_roots;
    a.n_shards = h->board_shards; a.shard = h->board_shard; a.phases = phases;
    a.seq = h->board_seq;
    a.timeout_ns = (unsigned long long)timeout_ms * 1000000ull;
    a.err = h->err;
    if (h->pdl) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(1);
        cfg.blockDim = dim3(kMaxShards);
        cfg.stream = s;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        CK(h, cudaLaunchKernelEx(&cfg, roots_exchange_kernel, (const RootsArgs)a));
    } else {
        roots_exchange_kernel<<<1, kMaxShards, 0, s>>>(a);
    }
    CK(h, cudaGetLastError());
    h->st.kernel_launches++;
    return RETH_OK;
}

}  // namespace

RETH_EXPORT int reth_replay_board_export(reth_replay_t *h, reth_board_export *out)
{
    if (!h || !out) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    int rc = board_alloc(h);
    if (rc) return rc;
    memset(out, 0, sizeof *out);
    out->device = h->cfg.device;
    out->pid = (int32_t)getpid();
    out->local_ptr = (uint64_t)(uintptr_t)h->board;
    CK(h, cudaIpcGetMemHandle((cudaIpcMemHandle_t *)out->handle, h->board));
    return RETH_OK;
}

RETH_EXPORT int reth_replay_board_connect(reth_replay_t *h, const reth_board_export *boards, int n_shards, int shard)
{
    if (!h || !boards) return RETH_ERR_INVALID;
    if (n_shards < 1 || n_shards > kMaxShards || shard < 0 || shard >= n_shards)
        return fail(h, RETH_ERR_INVALID, "board_connect: shard %d of %d (at most %d shards)", shard, n_shards, kMaxShards);
    if (h->board_shards) return fail(h, RETH_ERR_STATE, "the roots board is already connected");
    DeviceGuard guard(h->cfg.device);
    int rc = board_alloc(h);
    if (rc) return rc;
    for (int i = 0; i < n_shards; ++i) {
        const reth_board_export &b = boards[i];
        if (i == shard) { h->peer_board[i] = h->board; continue; }
        if (b.pid == (int32_t)getpid()) {
            // same process: the pointer is valid here; another device needs peer access (NVLink)
            if (b.device != h->cfg.device) {
                cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else if (e != cudaSuccess)
                    return fail(h, RETH_ERR_CUDA, "peer access %d -> %d: %s", h->cfg.device, b.device, cudaGetErrorString(e));
            }
            h->peer_board[i] = (BoardEntry *)(uintptr_t)b.local_ptr;
        } else {
            cudaIpcMemHandle_t mh;
            memcpy(&mh, b.handle, sizeof mh);
            void *p = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&p, mh, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                cudaGetLastError();
                for (int k = 0; k < i; ++k) {               // undo the mappings made so far
                    if (h->peer_ipc[k]) cudaIpcCloseMemHandle(h->peer_board[k]);
                    h->peer_ipc[k] = false;
                    h->peer_board[k] = nullptr;
                }
                return fail(h, RETH_ERR_CUDA, "cudaIpcOpenMemHandle (board of shard %d): %s", i, cudaGetErrorString(e));
            }
            h->peer_board[i] = (BoardEntry *)p;
            h->peer_ipc[i] = true;
        }
    }
    if (!h->board_ptrs_dev) CK(h, cudaMalloc((void **)&h->board_ptrs_dev, sizeof(BoardEntry *) * kMaxShards));
    CK(h, cudaMemcpy(h->board_ptrs_dev, h->peer_board, sizeof(BoardEntry *) * kMaxShards, cudaMemcpyHostToDevice));
    if (const char *e = getenv("RETH_MODEG_FUSED")) h->modeg_fused = atoi(e);
    h->board_shards = n_shards;
    h->board_shard = shard;
    return RETH_OK;
}

RETH_EXPORT int reth_replay_roots_exchange(reth_replay_t *h, int phases, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (phases < 1 || phases > 3) return fail(h, RETH_ERR_INVALID, "phases must be 1 (publish), 2 (wait) or 3 (both)");
    DeviceGuard guard(h->cfg.device);
    return launch_roots_exchange(h, phases, S(stream));
}

RETH_EXPORT int reth_replay_all_roots_dev(reth_replay_t *h, double **all_roots_dev_out)
{
    if (!h || !all_roots_dev_out) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    int rc = board_alloc(h);
    if (rc) return rc;
    *all_roots_dev_out = h->all_roots;
    return RETH_OK;
}

// ------------------------------------------------------------------ step loop --
namespace {

// Device-resident rows -> FIFO slots start, start+1, ... of the ring (at most two spans per column), on stream sc.
int launch_ring_copy(reth_replay *h, const void *const *cols, int64_t start, int64_t m, cudaStream_t sc)
{
    const int64_t cap = h->cfg.capacity;
    const int64_t len1 = (m < cap - start) ? m : cap - start, len2 = m - len1;
    AppendTiledArgs t = {};
    AppendArgs a = {};
    a.ncols = h->cfg.n_columns; a.start = start; a.n = m; a.capacity = cap;
    bool aligned = true;
    int64_t bytes = 0;
    for (int c = 0; c < a.ncols; ++c) {
        const int64_t rb = h->cfg.row_bytes[c];
        a.rb[c] = rb; a.src[c] = (const uint8_t *)cols[c]; a.ring[c] = h->ring[c];
        bytes += m * rb;
        uint8_t *d1 = h->ring[c] + start * rb;
        const uint8_t *s1 = a.src[c], *s2 = a.src[c] + len1 * rb;
        if ((((uintptr_t)d1 | (uintptr_t)s1 | (uint64_t)(len1 * rb)) & 15) ||
            (len2 > 0 && (((uintptr_t)h->ring[c] | (uintptr_t)s2 | (uint64_t)(len2 * rb)) & 15))) aligned = false;
        if (!aligned) continue;
        if (len1 > 0) {
            AppendSeg &g = t.seg[t.nseg++];
            g.dst = d1; g.src = s1; g.units = len1 * rb / 16; g.tile_base = t.tiles;
            t.tiles += (g.units + kAppendTileUnits - 1) / kAppendTileUnits;
        }
        if (len2 > 0) {
            AppendSeg &g = t.seg[t.nseg++];
            g.dst = h->ring[c]; g.src = s2; g.units = len2 * rb / 16; g.tile_base = t.tiles;
            t.tiles += (g.units + kAppendTileUnits - 1) / kAppendTileUnits;
        }
    }
    if (aligned && t.tiles > 0) {
        const int64_t maxg = (int64_t)h->n_sms * 8 * 64;
        ring_append_tiled_kernel<<<(unsigned)(t.tiles < maxg ? t.tiles : maxg), kCopyThreads, 0, sc>>>(t);
    } else {
        int64_t blocks = (bytes / 16 + kCopyThreads * 4 - 1) / (kCopyThreads * 4);
        const int64_t cap_blocks = (int64_t)h->n_sms * 8;
        blocks = blocks > cap_blocks ? cap_blocks : (blocks < 1 ? 1 : blocks);
        ring_append_kernel<<<(unsigned)blocks, kCopyThreads, 0, sc>>>(a);
    }
    CK(h, cudaGetLastError());
    h->st.kernel_launches++;
    return RETH_OK;
}

struct LoopIngest {
    const void *const *cols;      // device-resident rows of the producer: rows_per_group rows per column
    const float *prio;            // their |TD| priorities (device float32)
    int64_t rows_per_group;
};

int step_loop_impl(reth_replay_t *h, int steps, int batch, const float *prio_dev, const double *uniforms_dev,
                   uint64_t seed, double alpha, double beta, int flags, const LoopIngest *ing, void *stream);

}  // namespace

RETH_EXPORT int reth_replay_step_loop(reth_replay_t *h, int steps, int batch, const float *prio_dev,
                                      const double *uniforms_dev, uint64_t seed, double alpha, double beta,
                                      int flags, void *stream)
{
    return step_loop_impl(h, steps, batch, prio_dev, uniforms_dev, seed, alpha, beta, flags, nullptr, stream);
}

RETH_EXPORT int reth_replay_step_loop_ingest(reth_replay_t *h, int steps, int batch, const float *prio_dev,
                                             const double *uniforms_dev, uint64_t seed, double alpha, double beta,
                                             int flags, const void *const *ingest_cols_dev, const float *ingest_prio_dev,
                                             int64_t rows_per_group, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (rows_per_group < 0 || rows_per_group > h->cfg.capacity || (rows_per_group > 0 && (!ingest_cols_dev || !ingest_prio_dev)))
        return fail(h, RETH_ERR_INVALID, "bad ingest arguments");
    if ((flags & (2 | 64)) || !h->cfg.store_rows)
        return fail(h, RETH_ERR_INVALID, "the ingest loop needs row storage and runs neither serial nor in Mode G");
    LoopIngest ing = {ingest_cols_dev, ingest_prio_dev, rows_per_group};
    return step_loop_impl(h, steps, batch, prio_dev, uniforms_dev, seed, alpha, beta, flags, rows_per_group > 0 ? &ing : nullptr, stream);
}

namespace {
int step_loop_impl(reth_replay_t *h, int steps, int batch, const float *prio_dev, const double *uniforms_dev,
                   uint64_t seed, double alpha, double beta, int flags, const LoopIngest *ing, void *stream)
{
    if (!h) return RETH_ERR_INVALID;
    if (steps < 0 || !prio_dev) return fail(h, RETH_ERR_INVALID, "bad step_loop arguments");
    DeviceGuard guard(h->cfg.device);
    cudaStream_t s = S(stream);
    const bool no_gather = flags & 1, serial = flags & 2, timed = flags & 4;
    const bool no_sample = flags & 8, no_update = flags & 16;   // kernel micro-timing only
    const bool one_gather_stream = flags & 32;
    const bool global = flags & 64;                   // Mode G: roots exchange over NVLink + global stratified draw
    const int depth = (int)h->slots.size();
    int G = (flags >> 8) & 0xff;                      // batches per tree launch (prefetch group)
    if (G < 1) G = 1;
    if (global && (G > 32 || h->board_shards < 1 || no_sample || no_update || uniforms_dev))
        return fail(h, RETH_ERR_INVALID, "Mode G loop: connected roots board, group <= 32, device generator, sample and update on");
    if (G > 1 && (batch != h->cfg.max_batch || depth % G != 0 || depth < 2 * G || steps % G != 0))
        return fail(h, RETH_ERR_INVALID, "group %d needs batch == max_batch, ring_depth a multiple >= 2*group, steps %% group == 0", G);
    cudaStream_t st = serial ? s : h->s_tree;
    // the chain draw -> repair -> (exchange ->) draw is the critical path only without prefetch groups
    static const int pdl_env = env_int("RETH_PDL", 1);
    struct PdlScope { reth_replay *h; ~PdlScope() { h->pdl = false; } } pdl_scope{h};
    // (measured: strict alternation 18.4 -> 20.4 M transitions/s; Mode G loop at 2 GPUs 34 -> 32 us per step when
    // the share is small, but 50 -> 54 us when the step is gather-bound: grids resident ahead of their turn take
    // CTA slots from the gather, so Mode G uses it only for shares of at most 512 rows)
    h->pdl = pdl_env && G == 1 && !serial && (!global || batch / h->board_shards <= 512);
    if (!serial) {
        CK(h, cudaEventRecord(h->ev_fork, s));
        CK(h, cudaStreamWaitEvent(st, h->ev_fork, 0));
        CK(h, cudaStreamWaitEvent(h->s_gather, h->ev_fork, 0));
        CK(h, cudaStreamWaitEvent(h->s_gather2, h->ev_fork, 0));
        if (ing) CK(h, cudaStreamWaitEvent(h->s_append, h->ev_fork, 0));
    }
    bool copy_pending = false;            // an ingest copy is in flight on s_append: the next gather waits for it
    // timing diagnosis only (scripts/c4_probe.py): bit 0 the copy does not wait for gathers, bit 1 gathers do not wait for
    // the copy (both UNSAFE: rows may tear), bit 2 no leaf update for the new rows, bit 3 no row copy, bit 4 the leaves in
    // a launch of their own (as before they rode in the write-back's)
    const char *probe_env = ing ? getenv("RETH_C4_PROBE") : nullptr;
    const int probe = probe_env ? atoi(probe_env) : 0;
    for (int k = 0; k < steps; k += G) {
        const int slot = k % depth;
        // consecutive gathers alternate streams: the tail of one overlaps the ramp-up of the next
        cudaStream_t sg = serial ? s : ((h->two_gather_streams && !one_gather_stream && ((k / G) & 1)) ? h->s_gather2 : h->s_gather);
        // the slots' index buffers are re-drawn: their previous gather must have consumed them
        if (!serial && !no_gather && k >= depth) CK(h, cudaStreamWaitEvent(st, h->ev_gathered[slot], 0));
        int rc = RETH_OK;
        if (global) {
            // the first exchange of the call is its own kernel; after that the repair kernel of iteration k has
            // published exchange board_seq and the draw of iteration k+1 waits for it on the board itself
            // (the waiting is done by the single-CTA share kernel that precedes every draw of the loop: letting all
            // warps of the draw kernel poll the board deadlocks several shards sharing one device -- the spinning
            // CTAs keep a peer's repair cluster off the SMs -- and wins nothing on one shard per GPU)
            const bool fused = k > 0 && h->modeg_fused;
            if (!fused) rc = launch_roots_exchange(h, 3, st);
            h->wait_seq = fused ? h->board_seq : 0ull;
            h->quota_first = true;
            if (!rc) rc = do_sample(h, batch, nullptr, RETH_DEVICE, seed, beta, slot, h->all_roots,
                                    h->board_shards, h->board_shard, st, G);
            h->wait_seq = 0ull;
            h->quota_first = false;
        } else if (!no_sample || h->slots[slot].batch != batch)
            rc = do_sample(h, batch, uniforms_dev ? uniforms_dev + (size_t)k * batch : nullptr, RETH_DEVICE,
                           seed, beta, slot, nullptr, 1, 0, st, G);
        if (rc) return rc;
        if (!no_gather) {
            if (!serial) {
                CK(h, cudaEventRecord(h->ev_sampled[slot], st));
                CK(h, cudaStreamWaitEvent(sg, h->ev_sampled[slot], 0));
            }
            // timing: with all gathers on one stream ONE bracket spans the first..last gather of this call
            // (a timing event per launch would serialise the stream and stretch every launch); with two
            // gather streams every launch gets its own bracket
            const bool span = timed && (one_gather_stream || serial);
            const bool open_bracket = timed && (!span || k == 0), close_bracket = timed && (!span || k + G >= steps);
            if (open_bracket) {
                while (h->ev_time.size() < h->ev_time_used + 2) {
                    cudaEvent_t e;
                    CK(h, cudaEventCreate(&e));
                    h->ev_time.push_back(e);
                }
                CK(h, cudaEventRecord(h->ev_time[h->ev_time_used], sg));
            }
            if (copy_pending && !(probe & 2)) CK(h, cudaStreamWaitEvent(sg, h->ev_ingest, 0));   // rows the draw may have picked are in place
            rc = do_gather(h, slot, global, sg, G);
            if (rc) return rc;
            if (timed) h->ev_time_launches++;
            if (close_bracket) {
                CK(h, cudaEventRecord(h->ev_time[h->ev_time_used + 1], sg));
                h->ev_time_used += 2;
            }
            if (!serial) { CK(h, cudaEventRecord(h->ev_gathered[slot], sg)); note_gather(h, h->ev_gathered[slot], sg); }
        }
        if (no_update) continue;
        UpdateArgs a = {};
        a.n = (int64_t)batch * G; a.idx = h->slots[slot].idx; a.idx_kind = 0;
        a.prio = prio_dev + (size_t)k * batch; a.prio_kind = 0;
        a.alpha = alpha; a.transform = 1; a.dedupe = G == 1 ? 2 : 1;
        // config 4: the LEAVES of the actor rows that arrived during this group ride in the write-back's launch, as its
        // last entries (batch order: a sampled slot that is being recycled ends up with the new row's priority, as in
        // the call-by-call order write-back, then append).  A launch of their own cost 33 us of the tree stream per
        // group against 42 us for the draw and the write-back together (profiles/r02_c4_probe.jsonl).
        const bool fold = ing && !(probe & 4) && !(probe & 16);
        if (fold) { a.n += ing->rows_per_group; a.tail_n = ing->rows_per_group; a.tail_base = h->tail; a.tail_prio = ing->prio; a.dedupe = 1; }
        if (global) {
            a.n_dev = h->slots[slot].meta + 1;                  // this shard's share of the global batch
            if (h->modeg_fused && k + G < steps) a.pub_seq = ++h->board_seq;
        }
        static const int modeg_critical = env_int("RETH_MODEG_CRITICAL", 1);
        rc = launch_update(h, a, st, global ? (a.n / h->board_shards > 0 ? a.n / h->board_shards : 1) : 0,
                           global && modeg_critical);
        if (rc) return rc;
        h->st.updated_rows += (int64_t)batch * G;
        if (ing) {
            // Actor rows that arrived during this group (config 4): their LEAVES are on the tree stream already (last
            // entries of the write-back's launch) -- the chain draw -> repair -> draw never waits for a gather -- while
            // the ROW copy runs on the ingest stream behind every gather enqueued so far (those may still read the slots
            // being recycled), and every later gather waits for the copy (a later draw may pick the new slots).
            const int64_t n = ing->rows_per_group, cap = h->cfg.capacity, start = h->tail;
            rc = (probe & 1) ? RETH_OK : wait_gathers(h, h->s_append);
            if (!rc && !(probe & 8)) rc = launch_ring_copy(h, ing->cols, start, n, h->s_append);
            if (!rc) rc = mark_frame_refs(h, nullptr, start, n, h->s_append);
            if (rc) return rc;
            CK(h, cudaEventRecord(h->ev_ingest, h->s_append));
            copy_pending = true;
            UpdateArgs u = {};
            u.n = n; u.idx = nullptr; u.idx_kind = 2; u.iota_base = start;
            u.prio = ing->prio; u.prio_kind = 0;
            u.alpha = alpha; u.transform = 1; u.dedupe = 0;
            if (!(probe & 4) && !fold) rc = launch_update(h, u, st);
            if (rc) return rc;
            h->tail = (start + n) % cap;
            h->size = (h->size + n < cap) ? h->size + n : cap;
            h->st.appended_rows += n;
        }
    }
    if (ing && !serial) {
        CK(h, cudaEventRecord(h->ev_app_done, h->s_append));
        CK(h, cudaStreamWaitEvent(s, h->ev_app_done, 0));
        // whoever gathers next (another call of the loop, a loader pipe) may have drawn the rows of the last copy
        CK(h, cudaStreamWaitEvent(h->s_gather, h->ev_app_done, 0));
        CK(h, cudaStreamWaitEvent(h->s_gather2, h->ev_app_done, 0));
    }
    if (!serial) {
        CK(h, cudaEventRecord(h->ev_join_t, st));
        CK(h, cudaStreamWaitEvent(s, h->ev_join_t, 0));
        CK(h, cudaEventRecord(h->ev_join_g, h->s_gather));
        CK(h, cudaStreamWaitEvent(s, h->ev_join_g, 0));
        CK(h, cudaEventRecord(h->ev_join_g, h->s_gather2));
        CK(h, cudaStreamWaitEvent(s, h->ev_join_g, 0));
    }
    return RETH_OK;
}

}  // namespace

RETH_EXPORT int reth_replay_gather_time(reth_replay_t *h, double *total_ms, int64_t *launches)
{
    if (!h) return RETH_ERR_INVALID;
    DeviceGuard guard(h->cfg.device);
    CK(h, cudaDeviceSynchronize());
    double tot = 0;
    for (size_t i = 0; i + 1 < h->ev_time_used; i += 2) {
        float ms = 0;
        CK(h, cudaEventElapsedTime(&ms, h->ev_time[i], h->ev_time[i + 1]));
        tot += ms;
    }
    if (total_ms) *total_ms = tot;
    if (launches) *launches = h->ev_time_launches;
    h->ev_time_used = 0;
    h->ev_time_launches = 0;
    return RETH_OK;
}

// Diagnostic: SM-clock stamps of the stage boundaries of the last update launch
// (only with RETH_DEBUG_TIMING=1 in the environment at create time). out: int64[32].
RETH_EXPORT int reth_replay_debug_stamps(reth_replay_t *h, int64_t *out)
{
    if (!h || !out) return RETH_ERR_INVALID;
    if (!h->dbg) return fail(h, RETH_ERR_STATE, "RETH_DEBUG_TIMING was not set at create time");
    DeviceGuard guard(h->cfg.device);
    CK(h, cudaDeviceSynchronize());
    CK(h, cudaMemcpy(out, h->dbg, 32 * sizeof(int64_t), cudaMemcpyDeviceToHost));
    return RETH_OK;
}
