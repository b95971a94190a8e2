/*
 * per_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the prioritized-experience-replay arithmetic of
 * sosp2021/Reth (reth_buffer + its in-process twin reth.buffer).  It is the
 * checker the CUDA engine in reth_b200/ is compared against, and the timed
 * "cpu_baseline"/"--impl reference" leg of bench.py.  Nothing in the product
 * path (reth_b200/) may import, link or call this file.
 *
 * Parity status: PINNED.  the .npz files in tests/golden/ were produced by running the
 * reference's own numba-jitted functions (imported from /root/reference by
 * tests/golden/make_golden.py) and tests/test_oracle_golden.py checks this
 * file against them bit for bit.
 *
 * Every function cites the reference file:line it restates (paths relative to
 * the reference checkout).  Arithmetic is IEEE float64 in the same operation
 * order as the reference; build with -ffp-contract=off (see oracle/Makefile).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define PO_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------ tree -- */

/* reth_buffer/reth_buffer/utils/sumtree.py:82-88 -- three float64[capacity]
 * arrays; node i owns _val[i] and has children 2i+1, 2i+2. */
typedef struct po_tree {
    int64_t capacity;
    double *sum;
    double *min;
    double *val;
} po_tree;

PO_API po_tree *po_tree_create(int64_t capacity)
{
    po_tree *t = (po_tree *)calloc(1, sizeof(po_tree));
    if (!t) return NULL;
    t->capacity = capacity;
    t->sum = (double *)calloc((size_t)capacity, sizeof(double));
    t->min = (double *)calloc((size_t)capacity, sizeof(double));
    t->val = (double *)calloc((size_t)capacity, sizeof(double));
    return t;
}

PO_API void po_tree_destroy(po_tree *t)
{
    if (!t) return;
    free(t->sum); free(t->min); free(t->val); free(t);
}

/* sumtree.py:98-101 */
PO_API void po_tree_clear(po_tree *t)
{
    memset(t->sum, 0, (size_t)t->capacity * sizeof(double));
    memset(t->min, 0, (size_t)t->capacity * sizeof(double));
    memset(t->val, 0, (size_t)t->capacity * sizeof(double));
}

PO_API double *po_tree_sum_ptr(po_tree *t) { return t->sum; }
PO_API double *po_tree_min_ptr(po_tree *t) { return t->min; }
PO_API double *po_tree_val_ptr(po_tree *t) { return t->val; }
PO_API int64_t po_tree_capacity(po_tree *t) { return t->capacity; }

/* sumtree.py:5-21 (_numba_maintain_node) */
static inline void po_maintain_node(po_tree *t, int64_t idx)
{
    const int64_t left = idx * 2 + 1, right = idx * 2 + 2;
    double sum_res = t->val[idx];
    double min_res = (t->val[idx] != 0) ? t->val[idx] : 1.0;
    if (left < t->capacity) {
        sum_res += t->sum[left];
        if (t->min[left] != 0) min_res = (t->min[left] < min_res) ? t->min[left] : min_res;
    }
    if (right < t->capacity) {
        sum_res += t->sum[right];
        if (t->min[right] != 0) min_res = (t->min[right] < min_res) ? t->min[right] : min_res;
    }
    t->sum[idx] = sum_res;
    t->min[idx] = min_res;
}

/* sumtree.py:24-31 (_numba_maintain): idx -> root */
static inline void po_maintain(po_tree *t, int64_t idx)
{
    int64_t cur = idx;
    for (;;) {
        po_maintain_node(t, cur);
        if (cur == 0) break;
        cur = (cur - 1) / 2;
    }
}

/* sumtree.py:34-58 (_numba_find_index) */
PO_API int64_t po_find_index(const po_tree *t, double weight)
{
    int64_t cur = 0;
    for (;;) {
        const int64_t left = cur * 2 + 1, right = cur * 2 + 2;
        if (left < t->capacity) {
            if (weight < t->sum[left]) { cur = left; continue; }
            weight -= t->sum[left];
        }
        if (weight < t->val[cur] + 1e-5) return cur;
        weight -= t->val[cur];
        if (right >= t->capacity) return cur;
        cur = right;
    }
}

/* sumtree.py:61-67 (_numba_update): sequential, last writer wins */
PO_API void po_tree_update(po_tree *t, const int64_t *indices, const double *weights, int64_t n)
{
    for (int64_t i = 0; i < n; ++i) {
        t->val[indices[i]] = weights[i];
        po_maintain(t, indices[i]);
    }
}

/* sumtree.py:70-79 (_numba_sample) with the uniforms injected: the reference
 * draws u from numba's private generator; parity is defined for equal u. */
PO_API void po_tree_sample(const po_tree *t, int64_t batch, const double *u,
                           int64_t *out_idx, double *out_val)
{
    const double segment = t->sum[0] / (double)batch;
    for (int64_t i = 0; i < batch; ++i) {
        const double target = ((double)i + u[i]) * segment;
        const int64_t idx = po_find_index(t, target);
        out_idx[i] = idx;
        out_val[i] = t->val[idx];
    }
}

PO_API double po_tree_total(const po_tree *t) { return t->sum[0]; }

/* min(): reth_buffer/reth_buffer/utils/sumtree.py:109-110 (mode 0, service)
 *        reth/reth/buffer/sumtree.py:112-113            (mode 1, in-process) */
PO_API double po_tree_minimum(const po_tree *t, int mode)
{
    if (mode == 0) return (t->sum[0] != 0) ? t->min[0] : 1.0;
    return (t->min[0] != 0) ? t->min[0] : 1.0;
}

/* ------------------------------------------------------------ PER sampler -- */

/* reth_buffer/reth_buffer/sampler/per_sampler.py:16-17 (_normalize_weights):
 * (w + 1e-6) ** alpha evaluated by NumPy in the dtype of the incoming array.
 * float32 in -> float32 add, float32 pow (alpha demoted to float32). */
PO_API void po_normalize_f32(const float *w, int64_t n, double alpha, double *out)
{
    const float a = (float)alpha, eps = (float)1e-6;
    for (int64_t i = 0; i < n; ++i) out[i] = (double)powf(w[i] + eps, a);
}

PO_API void po_normalize_f64(const double *w, int64_t n, double alpha, double *out)
{
    for (int64_t i = 0; i < n; ++i) out[i] = pow(w[i] + 1e-6, alpha);
}

/* per_sampler.py:24-28: weights = (val / sumtree.min()) ** (-beta) */
PO_API void po_is_weights(const double *val, int64_t n, double pmin, double beta, double *out)
{
    for (int64_t i = 0; i < n; ++i) out[i] = pow(val[i] / pmin, -beta);
}

/* ----------------------------------------------------------- FIFO + ring -- */

/* Slot allocation: reth_buffer/reth_buffer/cache_policy/fifo_policy.py:11-18
 * (tail, tail+1, ... mod capacity).  Columnar ring with wrap and the row
 * gather: reth/reth/buffer/buffer.py:57-86 (append_batch) and :92-93 (select). */
#define PO_MAX_COLS 16
typedef struct po_ring {
    int64_t capacity;
    int32_t n_cols;
    int64_t row_bytes[PO_MAX_COLS];
    uint8_t *cols[PO_MAX_COLS];
    int64_t tail;      /* next slot to write (fifo_policy.py:8) */
    int64_t size;
} po_ring;

PO_API po_ring *po_ring_create(int64_t capacity, int32_t n_cols, const int64_t *row_bytes)
{
    if (n_cols > PO_MAX_COLS) return NULL;
    po_ring *r = (po_ring *)calloc(1, sizeof(po_ring));
    if (!r) return NULL;
    r->capacity = capacity; r->n_cols = n_cols;
    for (int c = 0; c < n_cols; ++c) {
        r->row_bytes[c] = row_bytes[c];
        r->cols[c] = (uint8_t *)malloc((size_t)(capacity * row_bytes[c]));
        if (!r->cols[c]) return NULL;
    }
    return r;
}

PO_API void po_ring_destroy(po_ring *r)
{
    if (!r) return;
    for (int c = 0; c < r->n_cols; ++c) free(r->cols[c]);
    free(r);
}

PO_API uint8_t *po_ring_col_ptr(po_ring *r, int c) { return r->cols[c]; }
PO_API int64_t po_ring_tail(po_ring *r) { return r->tail; }
PO_API int64_t po_ring_size(po_ring *r) { return r->size; }
PO_API void po_ring_reset(po_ring *r) { r->tail = 0; r->size = 0; }

/* fifo_policy.py:11-18 */
PO_API void po_fifo_indices(int64_t capacity, int64_t *tail, int64_t n, int64_t *out)
{
    for (int64_t i = 0; i < n; ++i) {
        out[i] = *tail;
        *tail = (*tail + 1) % capacity;
    }
}

/* buffer.py:57-86: two slice copies per column; returns the slots written */
PO_API void po_ring_append(po_ring *r, const uint8_t *const *src, int64_t n, int64_t *out_idx)
{
    const int64_t start = r->tail;
    const int64_t len1 = (n < r->capacity - start) ? n : r->capacity - start;
    const int64_t len2 = n - len1;
    for (int c = 0; c < r->n_cols; ++c) {
        const int64_t rb = r->row_bytes[c];
        memcpy(r->cols[c] + start * rb, src[c], (size_t)(len1 * rb));
        if (len2 > 0) memcpy(r->cols[c], src[c] + len1 * rb, (size_t)(len2 * rb));
    }
    if (out_idx) po_fifo_indices(r->capacity, &r->tail, n, out_idx);
    else r->tail = (r->tail + n) % r->capacity;
    r->size = (r->size + n < r->capacity) ? r->size + n : r->capacity;
}

/* buffer.py:92-93: [buffer[indices] for buffer in self._buffers].
 * slot_mask: bench-only knob; rows are read from slot (idx & slot_mask) so a
 * bounded host ring can stand behind a full-capacity tree (-1 = no masking). */
PO_API void po_ring_select(const po_ring *r, const int64_t *idx, int64_t n,
                           int64_t slot_mask, uint8_t *const *dst)
{
    for (int c = 0; c < r->n_cols; ++c) {
        const int64_t rb = r->row_bytes[c];
        for (int64_t i = 0; i < n; ++i)
            memcpy(dst[c] + i * rb, r->cols[c] + (idx[i] & slot_mask) * rb, (size_t)rb);
    }
}

/* ------------------------------------------------- timed CPU baseline loop -- */

/* One "step" of the measured path, composed exactly as SURVEY.md 8(d) Tier A:
 * PERSampler.sample (per_sampler.py:24-28) -> NumpyBuffer.select (buffer.py:92-93)
 * -> PERSampler.update (per_sampler.py:34-35) with float32 |TD| priorities.
 * u: [steps*batch] uniforms, new_prio: [steps*batch] float32.
 * Returns a checksum so the work cannot be optimised away. */
PO_API double po_step_loop(po_tree *t, const po_ring *r, int64_t slot_mask, int64_t steps,
                           int64_t batch, const double *u, const float *new_prio,
                           double alpha, double beta, int min_mode,
                           int64_t *idx_buf, double *val_buf, double *w_buf,
                           double *p_buf, uint8_t *const *batch_cols)
{
    double acc = 0;
    for (int64_t s = 0; s < steps; ++s) {
        po_tree_sample(t, batch, u + s * batch, idx_buf, val_buf);
        po_is_weights(val_buf, batch, po_tree_minimum(t, min_mode), beta, w_buf);
        if (r) po_ring_select(r, idx_buf, batch, slot_mask, batch_cols);
        po_normalize_f32(new_prio + s * batch, batch, alpha, p_buf);
        po_tree_update(t, idx_buf, p_buf, batch);
        acc += w_buf[0] + (double)idx_buf[batch - 1];
        if (r) acc += batch_cols[0][0];
    }
    return acc;
}
