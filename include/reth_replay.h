/*
 * reth_replay.h -- C ABI of the B200-native prioritized-experience-replay shard.
 *
 * One handle = one replay shard on one GPU: an HBM-resident columnar ring of
 * transitions, a warp-tiled float64 sum/min tree over the ring slots, and a
 * ring of contiguous learner-batch buffers.  Every call is stream-ordered on
 * the caller's cudaStream_t (passed as void*; NULL = the legacy default
 * stream); a handle is NOT re-entrant (SURVEY.md 8b "Threading").
 *
 * Each entry point names the reference (sosp2021/Reth) interface it replaces;
 * paths are relative to the reference checkout.  Semantics that callers of the
 * reference rely on are kept: slot == tree node index == arrival order modulo
 * capacity (cache_policy/fifo_policy.py:11-18), sequential last-writer-wins
 * updates (utils/sumtree.py:61-67), stratified proportional sampling over the
 * in-order traversal of the implicit heap (utils/sumtree.py:34-58,70-79),
 * importance weights (val/min)^-beta (sampler/per_sampler.py:24-28).
 *
 * Return value: 0 on success, a negative reth_status on error; the message is
 * available from reth_replay_last_error().  Nothing throws across this ABI.
 */
#ifndef RETH_REPLAY_H_
#define RETH_REPLAY_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RETH_MAX_COLUMNS 16

typedef struct reth_replay reth_replay_t;

typedef enum reth_status {
    RETH_OK = 0,
    RETH_ERR_INVALID = -1,  /* bad argument (the reference asserts: client/client.py:22-25,38) */
    RETH_ERR_CUDA = -2,     /* CUDA runtime error, see last_error */
    RETH_ERR_OOM = -3,
    RETH_ERR_RANGE = -4,    /* an index outside [0, capacity) reached a kernel */
    RETH_ERR_STATE = -5     /* call not valid in this state (e.g. sample on empty tree) */
} reth_status;

/* where a pointer argument lives */
typedef enum reth_loc { RETH_HOST = 0, RETH_DEVICE = 1 } reth_loc;
/* element type of a priority / index array */
typedef enum reth_dtype { RETH_F32 = 0, RETH_F64 = 1, RETH_I32 = 2, RETH_I64 = 3 } reth_dtype;

/* tree min() flavour: utils/sumtree.py:109-110 (service) vs
 * reth/reth/buffer/sumtree.py:112-113 (in-process twin) */
typedef enum reth_min_mode { RETH_MIN_SERVICE = 0, RETH_MIN_INPROCESS = 1 } reth_min_mode;

typedef struct reth_replay_config {
    int64_t capacity;                       /* ring slots == tree nodes (start_per(capacity,..), __init__.py:30) */
    int32_t n_columns;                      /* row struct discovered from the first row in the reference
                                               (client/torch_cuda_loader.py:106-112); explicit here */
    int64_t row_bytes[RETH_MAX_COLUMNS];    /* bytes of one row of each column */
    int32_t max_batch;                      /* largest sample()/update() batch (batch_size of start_per) */
    int32_t ring_depth;                     /* learner-batch slots (TorchCudaLoader buffer_size, :70) */
    int32_t device;                         /* CUDA device ordinal */
    int32_t min_mode;                       /* reth_min_mode */
    int32_t update_cluster;                 /* CTAs per cluster of the tree-repair kernel (0 = default: 16, or 8 if the device cannot host 16) */
    int32_t gather_mode;                    /* 0 = auto (= 1, the faster one on B200), 1 = LDG/STG.128,
                                               2 = TMA bulk copies (cp.async.bulk + mbarrier ring) */
    int32_t store_rows;                     /* 0 = tree only (no row storage; config 5 "tree-only") */
    int32_t frame_bytes;                    /* frame de-duplication: bytes of one frame (multiple of 16) and */
    int32_t frame_capacity;                 /* frames in the frame ring; used by RETH_COL_FRAMES columns */
    int32_t reserved[3];
    /* per column, what the GATHER emits into the learner batch (ring storage is unchanged):
     * 0 raw bytes; 1 uint8 -> float32; 2 uint8 -> float32 / 255.  Replaces the learner's first pass
     * over the batch, ensure_tensor(s0, torch.float, ...) in reth/reth/algorithm/dqn/dqn_solver.py:71-75
     * (reth/reth/algorithm/util.py:5-12) and ScaledFloatFrame (reth/reth/env/util.py:219): frames stay
     * uint8 in HBM (4x the capacity of the float32 rows test/apex-dqn/worker.py:47-51 stores) and the
     * batch column is float32[row_bytes] per row. */
    int32_t out_cast[RETH_MAX_COLUMNS];
} reth_replay_config;
typedef enum reth_cast { RETH_CAST_NONE = 0, RETH_CAST_U8_F32 = 1, RETH_CAST_U8_F32_UNIT = 2 } reth_cast;
/* out_cast[c] |= RETH_COL_FRAMES: column c is a FRAME-REFERENCE column.  Its stored row is K = row_bytes[c] / 4
 * int32 slots of the shard's frame ring (reth_replay_append_frames); the gather expands it to the K frames
 * themselves, K * frame_bytes per batch row (x4 as float32 with a cast).  Single frames are then stored once
 * instead of once per stack position and per s0/s1 (the reference's FrameStack concatenates k frames into every
 * observation, reth/reth/env/util.py:179-209; its LazyFrames remedy is commented out at :207): an Atari
 * transition costs 7 KB of HBM instead of 56 KB, so a GPU holds 8x the transitions. */
#define RETH_COL_FRAMES 0x100

/* counters for the throughput report (SURVEY.md 5 "Metrics") */
typedef struct reth_replay_stats {
    int64_t appended_rows, sampled_rows, updated_rows, gathered_rows;
    int64_t kernel_launches;                /* launches of this library's own kernels */
    int64_t tail;                           /* next FIFO slot */
    int64_t size;                           /* min(appended, capacity) */
    int64_t hbm_bytes;                      /* device memory held by the shard */
} reth_replay_stats;

/* ---- lifetime: replaces main_loop bring-up of storage + sampler state
 *      (server/main_loop.py:110-216; LMDB init :48-53; PERSampler.__init__ per_sampler.py:6-14) */
int reth_replay_create(const reth_replay_config *cfg, reth_replay_t **out);
int reth_replay_destroy(reth_replay_t *h);
const char *reth_replay_last_error(const reth_replay_t *h);   /* h may be NULL: create() errors */
int reth_replay_abi_version(void);

/* PERSampler.clear / NumbaSumTree.clear (per_sampler.py:19-22, utils/sumtree.py:98-101)
 * + FIFO cursor reset (NumpyBuffer.clear, reth/reth/buffer/buffer.py:95-97) */
int reth_replay_clear(reth_replay_t *h, void *stream);

/* Re-shape the learner-batch ring: ring_depth slots of max_batch rows (TorchCudaLoader(buffer_size=...),
 * client/torch_cuda_loader.py:70,114-135 allocates its ring per loader; here the ring belongs to the shard,
 * so a shard that served batch-512 loaders can go on to serve Mode G global batches of n_shards x 512 beside
 * the same row store).  Synchronises the device; every slot is empty afterwards; not while a pipe is open. */
int reth_replay_batch_ring_resize(reth_replay_t *h, int max_batch, int ring_depth);

/* ---- append: replaces Client.append -> append_loop -> FIFOPolicy.get_indices ->
 *      LMDB put -> PERSampler.update (client/client.py:21-35, server/main_loop.py:44-58,
 *      cache_policy/fifo_policy.py:11-18, sampler/per_sampler.py:34-35) and the in-process
 *      NumpyBuffer.append_batch + sumtree.update (reth/reth/buffer/prioritized_buffer.py:35-42).
 *  cols[c]: n rows of row_bytes[c], C-contiguous, all on `cols_loc`.
 *  prio: n priorities (|TD|) of prio_dtype (F32/F64) on prio_loc, or NULL = raw 1.0
 *        (prioritized_buffer.py:36-37).
 *  transform != 0: leaf = (w + 1e-6)^alpha in the dtype of prio (per_sampler.py:16-17);
 *  transform == 0: leaf = w as given (prioritized_buffer.py:29-33).
 *  Slots are tail, tail+1, ... mod capacity; first_slot_out (host, may be NULL) gets `tail`. */
int reth_replay_append(reth_replay_t *h, const void *const *cols, int cols_loc,
                       const void *prio, int prio_dtype, int prio_loc, int64_t n,
                       double alpha, int transform, int64_t *first_slot_out, void *stream);

/* Frame ring of a shard with RETH_COL_FRAMES columns: n frames of frame_bytes, stored at frame slots
 * tail, tail+1, ... mod frame_capacity; first_frame_out (host, may be NULL) gets `tail`.  The caller writes
 * those slots into the frame-reference columns of the transitions it appends afterwards.  A frame slot is
 * reused after frame_capacity later frames: size the ring for the frames `capacity` transitions span
 * (one frame per step plus one per episode reset, times the environments that interleave).  Liveness is
 * tracked: every stored row remembers the oldest frame it refers to, and a gather that meets a transition
 * which has outlived one of its frames reports RETH_ERR_STATE at the next synchronising call instead of
 * handing the learner a stack rebuilt from another time or environment. */
int reth_replay_append_frames(reth_replay_t *h, const void *frames, int frames_loc, int64_t n,
                              int64_t *first_frame_out, void *stream);
/* parity / debug: frames [first_frame, first_frame + n) of the ring -> host */
int reth_replay_frame_read(reth_replay_t *h, int64_t first_frame, int64_t n, void *out_host, void *stream);

/* ---- n-step assembly at append time: replaces NStepAdder.push on the actor (reth/reth/utils/nstep_adder.py:11-28,
 *      driven per environment by test/apex-dqn/worker.py:35,53).  The actor sends its RAW 1-step rows; the shard
 *      keeps, per environment stream, the rows that have not left the adder yet and writes every finished n-step
 *      transition straight into its FIFO slot: reward = r_t + g r_{t+1} + ... (the reference's running product
 *      t_gamma *= gamma and its operation order; float32 rewards accumulate in float32 like numpy does on the
 *      worker's 0-d float32 array, float64 in float64), s1 = the s1 of the last row folded in, folding stops
 *      behind a row whose done flag is set, every other column is the row's own (its own done flag included).
 *      Works on full rows and on frame-reference columns alike (s1 is then k frame slots).
 *  nstep_open: n_streams environment streams, `step` = n, reward / s1 / done column indices of the row struct;
 *      max_rows = most rows of one append_nstep call that are staged at once (host-resident input).
 *  append_nstep: n raw rows of stream `stream_id` in step order.  The row that push p of this call sends out
 *      of the adder takes prio[p] (pushes that emit nothing ignore theirs; NULL = raw 1.0), as the worker
 *      computes its loss on the rows the adder returned (worker.py:53-60).  *emitted_out = rows stored
 *      (= pushes so far - step, clipped at 0, minus what earlier calls stored).  Rows still buffered when a
 *      stream ends are never stored (nstep_adder.py has no flush); nstep_reset(stream) = a fresh adder
 *      (stream_id -1: all streams). */
int reth_replay_nstep_open(reth_replay_t *h, int n_streams, int step, double gamma, int r_col, int s1_col,
                           int done_col, int64_t max_rows);
int reth_replay_nstep_reset(reth_replay_t *h, int stream_id);
int reth_replay_append_nstep(reth_replay_t *h, int stream_id, const void *const *cols, int cols_loc,
                             const void *prio, int prio_dtype, int prio_loc, int64_t n, double alpha,
                             int transform, int64_t *first_slot_out, int64_t *emitted_out, void *stream);

/* Plugin cache policies (cache_policy/base_policy.py:4-7): rows go to caller-chosen slots.
 * slots: int64[n] on slots_loc; duplicates resolve to the last row in batch order. */
int reth_replay_append_at(reth_replay_t *h, const int64_t *slots, int slots_loc,
                          const void *const *cols, int cols_loc,
                          const void *prio, int prio_dtype, int prio_loc, int64_t n,
                          double alpha, int transform, void *stream);

/* ---- update_priorities: replaces Client.update_priorities -> PERSampler.update ->
 *      _numba_update/_numba_maintain (client/client.py:37-39, per_sampler.py:34-35,
 *      utils/sumtree.py:61-67,24-31,5-21).  idx_dtype I32 or I64.  Sequential semantics:
 *      duplicate indices are last-writer-wins in batch order. */
int reth_replay_update(reth_replay_t *h, const void *idx, int idx_dtype, int idx_loc,
                       const void *prio, int prio_dtype, int prio_loc, int64_t n,
                       double alpha, int transform, void *stream);

/* Same, for the indices the shard itself drew into batch slot `slot` (no index copy). */
int reth_replay_update_slot(reth_replay_t *h, int slot, const void *prio, int prio_dtype,
                            int prio_loc, double alpha, int transform, void *stream);

/* ---- sample: replaces PERSampler.sample -> NumbaSumTree.sample -> _numba_sample /
 *      _numba_find_index + IS weights (per_sampler.py:24-28, utils/sumtree.py:70-79,34-58,109-110).
 *  uniforms: double[batch] in [0,1) on uniforms_loc (parity mode: the reference draws these
 *  from numba's private generator), or NULL = counter-based device generator keyed by
 *  (seed, internal draw counter).  Results land in batch slot `slot`:
 *  indices int64[batch], weights float64[batch], priorities float64[batch]. */
int reth_replay_sample(reth_replay_t *h, int batch, const double *uniforms, int uniforms_loc,
                       uint64_t seed, double beta, int slot, void *stream);

/* ---- gather: replaces the loader row loop (client/numpy_loader.py:35-50,
 *      client/torch_cuda_loader.py:49-61) and NumpyBuffer.select (reth/reth/buffer/buffer.py:92-93):
 *      rows ring[c][indices[i]] -> contiguous batch column c of slot `slot`. */
int reth_replay_gather(reth_replay_t *h, int slot, void *stream);

/* ---- prefetch groups: the reference pre-samples (sampler HWM 1, sampler_loop.py:14) and its
 *      loader keeps a ring of batches in flight (TorchCudaLoader buffer_size, torch_cuda_loader.py:70),
 *      so batch k+1..k+G are drawn before update k lands.  These calls do that for n_slots
 *      consecutive slots in ONE launch each (batch must equal max_batch when n_slots > 1):
 *      n_slots independent stratified draws from the same tree state; one gather over
 *      n_slots*batch rows; one write-back of n_slots*batch priorities (sequential semantics
 *      over the concatenation, i.e. a later slot wins a duplicate). uniforms_dev: device
 *      double[n_slots*batch] or NULL; prio: n_slots*batch values. */
int reth_replay_sample_multi(reth_replay_t *h, int batch, int first_slot, int n_slots,
                             const double *uniforms_dev, uint64_t seed, double beta, void *stream);
int reth_replay_gather_multi(reth_replay_t *h, int first_slot, int n_slots, void *stream);
int reth_replay_update_slots(reth_replay_t *h, int first_slot, int n_slots, const void *prio,
                             int prio_dtype, int prio_loc, double alpha, int transform, void *stream);

/* indices (and weights) of n_slots consecutive full slots in one device->host copy each:
 * host int64[n_slots*batch] / double[n_slots*batch], either may be NULL.  Synchronises. */
int reth_replay_group_read(reth_replay_t *h, int first_slot, int n_slots, int64_t *indices,
                           double *weights, void *stream);

/* Device pointers of a batch slot (lent, as TorchCudaLoader lends ring slots,
 * torch_cuda_loader.py:152-163).  Any out pointer may be NULL. */
int reth_replay_slot_ptrs(reth_replay_t *h, int slot, void **cols_out /*[n_columns]*/,
                          int64_t **indices_out, double **weights_out, double **prios_out,
                          int32_t *batch_out);

/* Copy a slot's indices / weights (and optionally batch columns) to host memory and
 * synchronise the stream: what NumpyLoader.sample returns (numpy_loader.py:27-51).
 * Any of indices/weights/cols may be NULL. */
int reth_replay_slot_read(reth_replay_t *h, int slot, int64_t *indices, double *weights,
                          void *const *cols, void *stream);

/* ---- parity / debug: the three arrays of NumbaSumTree in heap order (utils/sumtree.py:85-88).
 *      host double[capacity] each; any may be NULL.  Synchronises. */
int reth_replay_tree_read(reth_replay_t *h, double *sum, double *min, double *val, void *stream);
/* The same readback without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  tree_read() takes its device scratch from cudaMalloc on every call; this variant runs the same
 * kernel into a static device array of 3 x RETH_NOALLOC_MAX_CAPACITY doubles (1.5 MB held by every GPU that loads the
 * library, until exit) and calls no cudaMalloc / cudaFree / allocating library routine.  Shards with capacity >
 * RETH_NOALLOC_MAX_CAPACITY are rejected with RETH_ERR_INVALID.  Every call on a GPU shares the array, which keeps
 * the last call's contents (the kernel overwrites all it reads back, nothing needs clearing): calls that overlap on
 * the same GPU race -- serialise them.  Same outputs as tree_read(), bit for bit.  Synchronises. */
#define RETH_NOALLOC_MAX_CAPACITY 65536
int reth_replay_tree_read_noalloc(reth_replay_t *h, double *sum, double *min, double *val, void *stream);
/* root: sum() and min() as the reference defines them (utils/sumtree.py:106-110) */
int reth_replay_tree_root(reth_replay_t *h, double *total, double *minimum, void *stream);
/* Read rows straight from the ring (host out), for storage parity tests. */
int reth_replay_ring_read(reth_replay_t *h, int column, int64_t first_slot, int64_t n,
                          void *out_host, void *stream);

int reth_replay_stats_get(reth_replay_t *h, reth_replay_stats *out);

/* ---- multi-shard (SURVEY.md 8e "Mode G"): stratified targets over the GLOBAL CDF of
 *      n_shards shards; this shard serves the targets inside its range.
 *  root_dev(): device pointer to this shard's {total, min} pair (double[2]), the 16 bytes
 *  each rank contributes to the all-gather.
 *  sample_global(): all_roots = device double[2*n_shards] (the all-gather result, rank
 *  order).  Draws `global_batch` stratified targets; the ones that fall in this shard are
 *  a contiguous range [first, first+count) of the global batch; they are written to the
 *  front of slot `slot` and {first,count} to the slot's meta (slot_meta()).  IS weights
 *  use the global minimum.  The reference has no counterpart (it polls shards
 *  round-robin, test/apex-dqn/trainer.py:33-36). */
int reth_replay_root_dev(reth_replay_t *h, double **root_dev_out);
int reth_replay_sample_global(reth_replay_t *h, int global_batch, const double *all_roots_dev,
                              int n_shards, int shard, const double *uniforms, int uniforms_loc,
                              uint64_t seed, double beta, int slot, void *stream);
/* host int64[2] = {first, count} of the last sample_global into `slot`; synchronises. */
int reth_replay_slot_meta(reth_replay_t *h, int slot, int64_t *first_count, void *stream);


/* ---- the other sampler plugins of start_server(samplers=[...]) (server/main_loop.py:135-159) on the
 *      same shard: batches are drawn into the same batch slots and gathered by the same kernel.
 *  UniformSampler (sampler/uniform_sampler.py:6-25): uniform_note() is its update(): the first
 *  `capacity` notified indices are logged in arrival order (appends AND priority updates notify, as
 *  in the reference where both travel the update channel, main_loop.py:58); the reference raises
 *  IndexError when a message crosses the capacity of the log; the C ABI drops the excess (the server keeps running,
 *  where the reference's sampler process would die) and the plugin class reth_b200.sampler.UniformSampler raises
 *  IndexError after the log has been filled, like the reference.
 *  append_note != 0 on reth_replay_append-style FIFO slots: idx == NULL, slots first_slot.. mod capacity.
 *  sample_uniform(): positions = draws[n] in [0, tail) (parity: the reference uses
 *  np.random.choice(tail, batch)) or NULL = device generator; weights are 1.
 *  FIFOSampler (sampler/fifo_sampler.py:8-29): fifo_push() is its update() (a bounded queue,
 *  deque(maxlen=capacity).appendleft: the oldest entry is dropped when full); sample_fifo() pops the
 *  `batch` oldest entries, RETH_ERR_STATE unless len > batch (ready_sample, :17-18).
 *  Both write n_slots consecutive batch slots in one launch (batch == max_batch when n_slots > 1). */
int reth_replay_uniform_note(reth_replay_t *h, const void *idx, int idx_dtype, int idx_loc, int64_t first_slot,
                             int64_t n, void *stream);
int reth_replay_sample_uniform(reth_replay_t *h, int batch, int first_slot, int n_slots, const int64_t *draws,
                               int draws_loc, uint64_t seed, void *stream);
int reth_replay_fifo_push(reth_replay_t *h, const void *idx, int idx_dtype, int idx_loc, int64_t first_slot,
                          const void *w, int w_dtype, int w_loc, int64_t n, void *stream);
int reth_replay_sample_fifo(reth_replay_t *h, int batch, int first_slot, int n_slots, void *stream);
/* host int64[4] = {uniform tail, fifo length, fifo head, 0} */
int reth_replay_sampler_state(reth_replay_t *h, int64_t *out);
/* which sampler reth_replay_pipe_issue() draws with: 0 PER (default), 1 uniform, 2 fifo */
int reth_replay_pipe_sampler(reth_replay_t *h, int kind);

/* ---- asynchronous loader pipeline (the end-to-end path with HOST index / priority buffers).
 *      Replaces, in one process, what the reference spreads over sampler_loop (pre-sampled batch
 *      behind a HWM-1 PUSH socket, server/sampler_loop.py:14,25-28,37-39), the TorchCudaLoader
 *      worker processes + device ring + interprocess events (client/torch_cuda_loader.py:20-66,
 *      94-135,152-163) and the fire-and-forget update channel (client/client.py:37-39 ->
 *      server/main_loop.py:79-88 -> sampler_loop.py:29-36).
 *  A "group" is `group` consecutive batch slots drawn by one launch; the batch ring holds
 *  n_groups = ring_depth / group group slots.  pipe_issue() enqueues, without blocking, on the
 *  shard's own streams:  sample `group` batches -> {indices, weights} D2H into pinned host rings
 *  || gather of group*batch rows.  pipe_wait() blocks until the OLDEST issued group is complete
 *  and lends it: its device batch slots and host index rows stay valid until the next
 *  pipe_wait() (the borrow rule of TorchCudaLoader.sample, torch_cuda_loader.py:152-163).
 *  At most n_groups - 1 groups can be issued-but-not-released at a time.
 *  Priority write-back: pipe_update_begin() reserves an entry of a pinned staging ring (entry e
 *  is the 16*entry_elems bytes at upd_host + e*16*entry_elems: int64 indices first, priorities
 *  (float32 or float64, packed) from byte 8*entry_elems), pipe_update_commit() enqueues ONE host->device copy and
 *  the tree update on the same stream as the sampling, so updates and draws are totally ordered
 *  by call order (the reference's single SUB socket).  pipe_stream() is that stream: pass it
 *  as `stream` to reth_replay_append() while a pipe is open to keep appends in the same order. */
typedef struct reth_pipe_info {
    int32_t batch, group, n_groups, upd_entries;
    int64_t entry_elems;        /* elements per group slot / staging entry = group * batch */
    int64_t *idx_host;          /* pinned int64 [n_groups][entry_elems]: delivered indices */
    double *w_host;             /* pinned double[n_groups][entry_elems]: delivered IS weights */
    void *upd_host;             /* pinned bytes [upd_entries][16*entry_elems] (see above) */
} reth_pipe_info;
int reth_replay_pipe_open(reth_replay_t *h, int batch, int group, uint64_t seed, reth_pipe_info *out);
int reth_replay_pipe_close(reth_replay_t *h);
int reth_replay_pipe_stream(reth_replay_t *h, void **stream_out);
int reth_replay_pipe_issue(reth_replay_t *h, double beta);
/* group_out: group slot g -> batch slots g*group .. g*group+group-1, host rows idx_host[g], w_host[g] */
int reth_replay_pipe_wait(reth_replay_t *h, int32_t *group_out);
/* steady-state call: hands the lent group back, pipe_issue()s a new group into its slots and then
 * pipe_wait()s for the oldest group in flight.  Issuing BEFORE blocking keeps a draw + gather
 * queued behind the group in flight, so the device does not idle while the host wakes up. */
int reth_replay_pipe_next(reth_replay_t *h, double beta, int32_t *group_out);
/* Keep the last n_slots batch slots of the ring out of the pipe (they serve other sampler topics
 * through gather_from()); call before pipe_open. */
int reth_replay_pipe_reserve(reth_replay_t *h, int n_slots);
/* Lending depth: the newest `groups` delivered groups stay valid (default 1; call before the first
 * pipe_issue).  A consumer in ANOTHER process asks for group k+1 when it starts on group k, so the
 * request that reaches the owner hands back group k-1: two groups are lent at a time there. */
int reth_replay_pipe_hold(reth_replay_t *h, int groups);
/* hand the oldest lent group back and pipe_issue() a new group into its slots; does not block */
int reth_replay_pipe_advance(reth_replay_t *h, double beta);
/* *ready_out = 1 when pipe_wait() would return without blocking */
int reth_replay_pipe_ready(reth_replay_t *h, int32_t *ready_out);
int reth_replay_pipe_update_begin(reth_replay_t *h, int32_t *entry_out);
int reth_replay_pipe_update_commit(reth_replay_t *h, int entry, int64_t n, int prio_dtype, double alpha,
                                   int transform);
/* convenience for C callers: begin + memcpy + commit */
int reth_replay_update_async(reth_replay_t *h, const int64_t *idx_host, const void *prio_host, int prio_dtype,
                             int64_t n, double alpha, int transform);
/* wait until everything enqueued on the pipe's streams has finished */
int reth_replay_pipe_drain(reth_replay_t *h);

/* ---- batch ring across processes: the reference hands pre-gathered batches to a learner in
 *      another process through CUDA IPC (client/torch_cuda_loader.py:58,63,114-135,157-160:
 *      shared CUDA tensors + interprocess events).  ipc_export() fills a plain-bytes descriptor of
 *      this shard's batch ring (one cudaIpcMemHandle per column array + indices + weights); a
 *      consumer process maps it with reth_ipc_open() FOR ITS CURRENT DEVICE (the same GPU, or a peer
 *      GPU reached over NVLink) and reads slot pointers.
 *      Completion is signalled by the owner process (it replies after pipe_wait()), so no
 *      interprocess event is needed. */
typedef struct reth_ipc_export {
    int32_t n_columns, ring_depth, max_batch, device;
    int64_t row_bytes[RETH_MAX_COLUMNS];
    unsigned char col_handle[RETH_MAX_COLUMNS][64];   /* cudaIpcMemHandle_t */
    unsigned char idx_handle[64], w_handle[64];
} reth_ipc_export;
typedef struct reth_ipc_view reth_ipc_view_t;
int reth_replay_ipc_export(reth_replay_t *h, reth_ipc_export *out);
int reth_ipc_open(const reth_ipc_export *in, reth_ipc_view_t **out);
int reth_ipc_slot_ptrs(reth_ipc_view_t *v, int slot, void **cols_out /*[n_columns]*/, int64_t **indices_out,
                       double **weights_out);
int reth_ipc_close(reth_ipc_view_t *v);
/* Cross-shard batch assembly over NVLink (north_star "only when the learner is not
 * co-located"): after reth_replay_sample_global() into `slot`, gather this shard's rows STRAIGHT
 * into batch slot `dst_slot` of a peer's ring (`dst`, opened with reth_ipc_open; may live on
 * another GPU: the stores cross NVLink from inside the gather kernel), at the global batch rows
 * [first, first+count) the shard serves; its indices and weights go to the same rows.  dst == NULL
 * names this shard's own ring (the learner rank's own slice; dst_slot != slot).  The
 * reference has no counterpart (its learner polls whole batches shard by shard, trainer.py:33-36). */
int reth_replay_gather_to(reth_replay_t *h, int slot, reth_ipc_view_t *dst, int dst_slot, void *stream);

/* ---- Mode G with a prefetch group: n_slots (<= 32) independent GLOBAL stratified draws of global_batch from
 *      the same tree state in one launch pair (share computation + descent), as sample_multi does for one
 *      shard.  This shard's shares are PACKED back to back from batch slot first_slot on: the n_slots slots
 *      then hold ONE batch of sum(count_g) rows (global_batch must equal max_batch), which gather_multi()
 *      and update_slots() over the same slots handle as such (row count on the device).  slot_quota() reads
 *      {first_g, count_g, offset_g, 0} per draw: rows offset_g .. offset_g+count_g of the packed batch are
 *      positions first_g .. of global batch g.  uniforms_dev: device double[n_slots*global_batch] or NULL. */
int reth_replay_sample_global_multi(reth_replay_t *h, int global_batch, const double *all_roots_dev,
                                    int n_shards, int shard, int first_slot, int n_slots,
                                    const double *uniforms_dev, uint64_t seed, double beta, void *stream);
int reth_replay_slot_quota(reth_replay_t *h, int first_slot, int n_slots, int64_t *out /*[4*n_slots]*/, void *stream);

/* ---- shard roots over NVLink (SURVEY.md 8e "read peers' totals via NVLink-mapped memory"; the
 *      reference has no counterpart: its shards never normalise against each other, trainer.py:33-36).
 *      Mode G needs every shard's root {total, min} on every rank before a global draw.  Instead of a
 *      host-driven collective per step, each shard owns a small "roots board" in its HBM; peers map it
 *      (CUDA IPC from another process, peer access inside one process) and roots_exchange() launches
 *      ONE kernel on `stream` that (phase 1) stores this shard's root into every peer's board over
 *      NVLink and (phase 2) waits on its own board until all n_shards roots of this exchange number
 *      have arrived, then snapshots them into the device array all_roots_dev() returns
 *      (double[2*n_shards], the all_roots argument of reth_replay_sample_global).  Collective in the
 *      NCCL sense: every shard of the group must issue the same sequence of exchanges; a peer that
 *      never publishes ends the wait after RETH_BOARD_TIMEOUT_MS (default 2000) and the next
 *      synchronising call returns RETH_ERR_STATE.  phases: 1 publish, 2 wait, 3 both in one launch
 *      (shards driven by ONE host thread on one stream must publish all before any waits). */
#define RETH_MAX_SHARDS 64
typedef struct reth_board_export {
    int32_t device, pid;
    uint64_t local_ptr;                 /* the board's address inside the exporting process */
    unsigned char handle[64];           /* cudaIpcMemHandle_t */
} reth_board_export;
int reth_replay_board_export(reth_replay_t *h, reth_board_export *out);
int reth_replay_board_connect(reth_replay_t *h, const reth_board_export *boards /*[n_shards], rank order*/,
                              int n_shards, int shard);
int reth_replay_roots_exchange(reth_replay_t *h, int phases, void *stream);
int reth_replay_all_roots_dev(reth_replay_t *h, double **all_roots_dev_out);

/* ---- several sampler topics over one row store (start_server(samplers=[...]), server/main_loop.py:135-159:
 *      every sampler sees every append and every update, loaders pick a topic).  The first topic
 *      uses the row shard's own tree; each further topic is a tree-only handle (store_rows = 0, same
 *      capacity, same device) that receives the same appends / updates.  gather_from() fills batch
 *      slot dst_slot of the ROW shard `h` with the rows, indices and weights of the batch the index
 *      shard `src` drew into its slot src_slot. */
int reth_replay_gather_from(reth_replay_t *h, int dst_slot, reth_replay_t *src, int src_slot, void *stream);

/* ---- measured loop (bench / learner-less soak): `steps` iterations of
 *      sample(batch) -> gather || update_slot(new priorities), the tree chain and the
 *      frame gather on two internal streams joined to `stream` at both ends, slots used
 *      round-robin.  prio_dev: device float32[steps*batch] new |TD| priorities;
 *      uniforms_dev: device double[steps*batch] or NULL (device generator).
 *      flags bit0: skip gather (tree only); bit1: serialise everything on one stream;
 *      bit2: time the gathers with CUDA events on their own stream (see gather_time): one bracket
 *      around the first..last gather of the call when they share a stream (bit5 or bit1), else one per launch;
 *      bit3 / bit4: skip the sample / the update launch (kernel micro-timing only);
 *      bit5: all gathers on ONE internal stream (default: consecutive gathers alternate between two
 *      streams so the tail of one overlaps the ramp-up of the next; bit2 timings are only per-kernel
 *      durations with bit5 set);
 *      bit6: Mode G (after reth_replay_board_connect): every iteration starts with roots_exchange(3)
 *      on the tree stream, draws this shard's share of G GLOBAL stratified batches of `batch` (packed,
 *      see sample_global_multi), gathers and writes back that share (prio_dev: the first share-count
 *      values of each iteration's G*batch); every shard of the group must run the same number of steps;
 *      bits 8..15: prefetch group G (0/1 = strict alternation): sample_multi / gather_multi /
 *      update_slots over G slots per iteration (steps must be a multiple of G). */
int reth_replay_step_loop(reth_replay_t *h, int steps, int batch, const float *prio_dev,
                          const double *uniforms_dev, uint64_t seed, double alpha, double beta,
                          int flags, void *stream);
/* The same loop with actor ingest (BASELINE.json config 4: learner loop + actor append streams; reference: appends
 * and draws are serialised by the one append process and the sampler's SUB socket, server/main_loop.py:44-58,
 * server/sampler_loop.py:29-39).  After every group of G learner steps, rows_per_group device-resident rows
 * (ingest_cols_dev[c]: rows_per_group rows of column c; ingest_prio_dev: their float32 |TD|) are appended:
 * their leaves ride in the write-back's repair launch, as its last entries (batch order: behind the write-back), so
 * the chain draw -> repair -> draw never waits for a gather and pays the repair's fixed cost once per group; the row copy runs on the ingest stream behind every gather enqueued so far (those may
 * still read the slots being recycled) and every later gather waits for it.  Same order of tree mutations as
 * step_loop(G) followed by append(), repeated.  Not with flags bit1 (serial) / bit6 (Mode G). */
int reth_replay_step_loop_ingest(reth_replay_t *h, int steps, int batch, const float *prio_dev,
                                 const double *uniforms_dev, uint64_t seed, double alpha, double beta, int flags,
                                 const void *const *ingest_cols_dev, const float *ingest_prio_dev,
                                 int64_t rows_per_group, void *stream);
/* Sum of the event-bracketed gather launch durations recorded since the last call
 * (step_loop flag bit2), and how many launches that covers.  Synchronises the device. */
int reth_replay_gather_time(reth_replay_t *h, double *total_ms, int64_t *launches);

/* Diagnostic only: SM-clock stamps (int64[32]) of the stage boundaries of the last
 * tree-repair launch; needs RETH_DEBUG_TIMING=1 in the environment at create time. */
int reth_replay_debug_stamps(reth_replay_t *h, int64_t *out);

/* ---- lz4 frame codec of the ingest wire format (host code, no handle).
 *      The reference's actors send compress=True bodies (test/apex-dqn/worker.py:60 ->
 *      utils/pack.py:63-84 lz4.frame.compress) and the server undoes them (utils/pack.py:152-156
 *      lz4.frame.decompress).  These replace the two calls of the `lz4` package (absent here):
 *      any conforming LZ4 frame is decoded (linked or independent blocks, optional checksums,
 *      skippable and concatenated frames); the compressor writes independent 64 KB blocks with the
 *      content size stored.  All four return a byte count, or a negative value:
 *      -1 truncated, -2 malformed, -3 checksum mismatch, -4 output too small, -5 dictionary id. */
int64_t reth_lz4f_content_size(const void *src, int64_t n);   /* -1: the frame does not store it */
int64_t reth_lz4f_decompress(const void *src, int64_t n, void *dst, int64_t cap);
int64_t reth_lz4f_compress_bound(int64_t n);
int64_t reth_lz4f_compress(const void *src, int64_t n, void *dst, int64_t cap);

/* ---- row messages -> columns (host code, no handle).  The reference's Client.append serializes every transition
 *      on its own (client/client.py:26-35) and the server / loaders undo that row by row in Python
 *      (client/numpy_loader.py:35-50).  Rows of one append share a struct, so they form an (n, L) byte matrix:
 *      this copies column c of every row (bytes [lo[c], lo[c]+len[c]) of each row) to the C-contiguous array
 *      dst[c] -- typically pinned staging memory the host->device copy then reads -- in one call without the
 *      interpreter (several ingest threads call it concurrently).  Returns the bytes copied, -1 on bad arguments. */
int64_t reth_pack_scatter_rows(const void *mat, int64_t n, int64_t L, int32_t ncols, const int64_t *lo,
                               const int64_t *len, void *const *dst);

#ifdef __cplusplus
}
#endif
#endif /* RETH_REPLAY_H_ */
