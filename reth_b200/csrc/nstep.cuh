// nstep.cuh -- n-step transition assembly at append time (sm_100a).
//
// The reference folds n-step returns on the ACTOR, row by row in Python (NStepAdder.push,
// reth/reth/utils/nstep_adder.py:11-28, driven by test/apex-dqn/worker.py:35,53): a deque of the last
// `step` rows; every push adds gamma^j * r to the rewards of the buffered rows and replaces their s1,
// scanning from the newest row and stopping at the first row whose done flag is set; the row that is
// `step` pushes old leaves the deque and is what gets stored.
//
// Here the actor sends its raw 1-step rows and the shard assembles the n-step transition while it
// appends.  What a row looks like when it leaves the adder depends only on the raw rows t .. t+step-1:
//     m      = largest j < step with done[t] .. done[t+j-1] all clear
//     reward = r[t] + g_1 r[t+1] + ... + g_m r[t+m],   g_j = the running product t_gamma *= gamma
//              (float32 column: float32 accumulation, factor rounded to float32 -- numpy on the 0-d
//               float32 array of worker.py:49; float64 column: float64), same operation order
//     s1     = s1[t+m]   (a full frame stack, or its k frame references in the de-duplicated layout)
//     every other column: row t's own
// so one CTA per (emitted row, column part) reads the <= step raw rows it needs -- from the stream's carry
// (the rows that have not left the adder yet) or from the new rows of this call -- and writes the finished
// transition STRAIGHT into its FIFO slot of the ring: no intermediate copy, no host arithmetic.
// nstep_carry_kernel then keeps the newest min(L, step) raw rows as the stream's next carry.
#pragma once
#include "copy.cuh"

namespace reth {

struct NStepCol {
    const uint8_t *carry;   // the stream's carry rows of this column (oldest first)
    uint8_t *carry_next;    // the other carry buffer (written by nstep_carry_kernel)
    const uint8_t *fresh;   // the new rows of this call
    uint8_t *ring;          // ring column (nullptr: tree-only shard)
    int64_t rb;
    int32_t parts;          // work items per row (wide columns are cut into pieces of part_bytes)
    int32_t part_base;
    int32_t part_bytes;
};

struct NStepArgs {
    NStepCol col[kMaxCols];
    int32_t ncols, step, r_col, s1_col, done_col;
    int32_t r_f64;          // reward column is float64 (else float32)
    int32_t carry_count;    // raw rows in the carry
    int32_t parts_per_row;
    int64_t n;              // new raw rows
    int64_t E;              // rows that leave the adder in this call = max(0, carry_count + n - step)
    int64_t tail, capacity; // emitted row e goes to slot (tail + e) % capacity
    double gamma;
};

__device__ __forceinline__ const uint8_t *nstep_row(const NStepArgs &a, const NStepCol &c, int64_t t)
{
    return t < a.carry_count ? c.carry + t * c.rb : c.fresh + (t - a.carry_count) * c.rb;
}

__device__ __forceinline__ bool nstep_done(const NStepArgs &a, int64_t t)
{
    const NStepCol &c = a.col[a.done_col];
    const uint8_t *p = nstep_row(a, c, t);
    bool any = false;
    for (int64_t q = 0; q < c.rb; ++q) any |= p[q] != 0;      // truthiness of the stored flag (nstep_adder.py:18)
    return any;
}

__global__ void __launch_bounds__(kCopyThreads) nstep_emit_kernel(const NStepArgs a)
{
    const int64_t items = a.E * a.parts_per_row;
    for (int64_t item = blockIdx.x; item < items; item += gridDim.x) {
        const int64_t e = item / a.parts_per_row;
        const int k = (int)(item - e * a.parts_per_row);
        int c = 0;
#pragma unroll 1
        for (int q = 0; q < a.ncols; ++q)
            if (k >= a.col[q].part_base) c = q;
        const NStepCol &col = a.col[c];
        if (!col.ring) continue;
        const int part = k - col.part_base;
        // rows e+1 .. e+m touch row e: m = pushes before a done row stops the scan (uniform over the CTA)
        int m = 0;
        if (c == a.r_col || c == a.s1_col)
            while (m + 1 < a.step && !nstep_done(a, e + m)) ++m;
        uint8_t *dst = col.ring + ((a.tail + e) % a.capacity) * col.rb;
        if (c == a.r_col) {
            if (threadIdx.x == 0) {
                double g = a.gamma;
                if (a.r_f64) {
                    double acc = *(const double *)nstep_row(a, col, e);
                    for (int j = 1; j <= m; ++j) {
                        acc = acc + g * *(const double *)nstep_row(a, col, e + j);     // :21
                        g *= a.gamma;                                                 // :22
                    }
                    *(double *)dst = acc;
                } else {
                    float acc = *(const float *)nstep_row(a, col, e);
                    for (int j = 1; j <= m; ++j) {
                        acc = __fadd_rn(acc, __fmul_rn((float)g, *(const float *)nstep_row(a, col, e + j)));
                        g *= a.gamma;
                    }
                    *(float *)dst = acc;
                }
            }
            continue;
        }
        const uint8_t *src = nstep_row(a, col, c == a.s1_col ? e + m : e);            // :24
        const int64_t off = (int64_t)part * col.part_bytes;
        const int64_t len = min((int64_t)col.part_bytes, col.rb - off);
        span_copy(dst + off, src + off, len, threadIdx.x, kCopyThreads);
    }
}

// carry_next[j] = raw row (L - keep + j) of the sequence carry ++ fresh, keep = min(L, step)
__global__ void __launch_bounds__(kCopyThreads) nstep_carry_kernel(const NStepArgs a, int keep)
{
    const int64_t L = a.carry_count + a.n;
    const int64_t items = (int64_t)keep * a.parts_per_row;
    for (int64_t item = blockIdx.x; item < items; item += gridDim.x) {
        const int64_t j = item / a.parts_per_row;
        const int k = (int)(item - j * a.parts_per_row);
        int c = 0;
#pragma unroll 1
        for (int q = 0; q < a.ncols; ++q)
            if (k >= a.col[q].part_base) c = q;
        const NStepCol &col = a.col[c];
        const int part = k - col.part_base;
        const int64_t off = (int64_t)part * col.part_bytes;
        const int64_t len = min((int64_t)col.part_bytes, col.rb - off);
        span_copy(col.carry_next + j * col.rb + off, nstep_row(a, col, L - keep + j) + off, len, threadIdx.x, kCopyThreads);
    }
}

}  // namespace reth
