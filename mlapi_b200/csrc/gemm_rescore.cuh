// fp32 re-scoring of the rows the tensor-core epilogue flagged.
// One translation unit with gemm_tcgen05.cu (included there, in this order); not a stand-alone header.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "generic_row.cuh"

namespace lrs {
namespace {

// ---- fp32 re-scoring of the rows flagged by the epilogue ------------------------------------------
// A flagged row has a runner-up inside the accumulation error band of its maximum, or a class whose
// probability that error could move by more than the parity bar.  Only the classes that carry
// weight matter for either, so the pass re-computes the logits of those classes in fp32 FMA
// arithmetic (error ~1e-6) and patches the row: one warp per flagged row, traffic = the row of X,
// the row of P and a few rows of W instead of all of W.  The epilogue leaves a compact list of the flagged
// rows, so the warps of this pass share them evenly whatever their density.
//
// Four dot products at once (rows c[0..3] of W, `ldw` floats apart, against the same row of X): the pieces of X
// are loaded once and sixteen independent loads are in flight per lane -- taken one class at a time, a row's ~20
// selected classes were ~20 dependent trips to L2.  `xal`: the row of X is 16-byte aligned and d % 4 == 0
// (128-bit loads of both); otherwise 4-byte loads over exactly d elements.
__device__ __forceinline__ void warp_dot4_x4(const float* __restrict__ x, const float* __restrict__ W, const int (&c)[4],
                                             int d, int ldw, bool xal, int lane, float (&z)[4]) {
    float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (xal) {
        const float4* x4 = reinterpret_cast<const float4*>(x);
        const float4* w4[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) w4[u] = reinterpret_cast<const float4*>(W + (int64_t)c[u] * ldw);
#pragma unroll 4
        for (int k = lane; k < d / 4; k += 32) {
            const float4 xv = __ldg(x4 + k);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float4 wv = __ldg(w4[u] + k);
                acc[u] = fmaf(xv.x, wv.x, acc[u]);
                acc[u] = fmaf(xv.y, wv.y, acc[u]);
                acc[u] = fmaf(xv.z, wv.z, acc[u]);
                acc[u] = fmaf(xv.w, wv.w, acc[u]);
            }
        }
    } else {
        const float* w1[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) w1[u] = W + (int64_t)c[u] * ldw;
#pragma unroll 4
        for (int k = lane; k < d; k += 32) {
            const float xv = __ldg(x + k);
#pragma unroll
            for (int u = 0; u < 4; ++u) acc[u] = fmaf(xv, __ldg(w1[u] + k), acc[u]);
        }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) acc[u] += __shfl_xor_sync(0xffffffffu, acc[u], off);
        z[u] = acc[u];
    }
}

constexpr int kFixWarps = 8;
constexpr int kFixQueue = 64;          // selected classes waiting for their dot products (per warp)

// What a warp adds up over the classes it re-scores for a row.
struct FixPart {
    float old_mass = 0.0f, new_mass = 0.0f, best_z = -INFINITY, best_q = 0.0f;
    int best = 0;
};

// The selected classes of [c_lo, c_hi) -- those with prow[c] >= thr -- get an fp32 logit: queued in ascending order,
// worked off four at a time.  prow[c] becomes exp(z - L), the class's weight in units of the old row sum.
__device__ __forceinline__ void fix_row_classes(const float* __restrict__ x, int d, int ldw, bool xal, const float* __restrict__ W,
                                                const float* __restrict__ b, float* prow, int* queue, int c_lo, int c_hi,
                                                float thr, float L, float* decision_row, int lane, FixPart& part) {
    int nq = 0;
    auto drain = [&](bool all) {
        int done = 0;
        while (nq - done >= 4 || (all && done < nq)) {
            const int take = min(4, nq - done);
            int cls[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) cls[u] = queue[done + (u < take ? u : 0)];
            float z4[4];
            warp_dot4_x4(x, W, cls, d, ldw, xal, lane, z4);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (u >= take) break;
                const int cj = cls[u];
                const float z = z4[u] + __ldg(b + cj);
                const float q = expf(z - L);                    // in units of the old row sum
                part.old_mass += prow[cj];
                part.new_mass += q;
                if (z > part.best_z) { part.best_z = z; part.best = cj; part.best_q = q; }   // ascending classes: first max wins
                __syncwarp();
                if (lane == 0) {
                    prow[cj] = q;
                    if (decision_row) decision_row[cj] = z;
                }
            }
            done += take;
        }
        __syncwarp();
        // what is left (fewer than four) moves to the front
        const int left = nq - done;
        const int keep = lane < left ? queue[done + lane] : 0;
        __syncwarp();
        if (lane < left) queue[lane] = keep;
        nq = left;
        __syncwarp();
    };
    for (int c0 = c_lo; c0 < c_hi; c0 += 32) {
        const int c = c0 + lane;
        const bool sel = c < c_hi && prow[c] >= thr;
        const unsigned mask = __ballot_sync(0xffffffffu, sel);
        if (mask) {
            if (nq + __popc(mask) > kFixQueue) drain(false);          // leaves at most 3
            if (sel) queue[nq + __popc(mask & ((1u << lane) - 1u))] = c;
            nq += __popc(mask);
            __syncwarp();
            if (nq >= 16) drain(false);
        }
    }
    drain(true);
}

// Which classes get an fp32 logit: the rivals of the maximum (near-ties), and every class whose own probability the
// error bound `ez` of this row could move by more than ~4e-6 (p (1 + mass) ez with p >= 2e-6 / ez).  What stays
// behind enters the row only through the renormalisation, as the probability-weighted MEAN of its logit errors:
// random parts average out over the many small classes, the systematic part of the truncation error is ~1/8 of the
// maximum `ez` is built on (measured mean / max, profiles/r02_gemm_accumulation_error.txt; 1/4 is assumed); the
// threshold is lowered until mass_left * ez / 4 * p_max <= 4e-6.  (Every lane of the calling warp gets the same value.)
__device__ __forceinline__ float fix_row_threshold(const float* prow, int C, float pmax, float ez, int lane) {
    float thr = fminf(fmaxf(2e-6f / ez, 1e-4f), 0.9f * pmax);
    for (int round = 0; round < 3; ++round) {
        float left = 0.0f;
        for (int c = lane; c < C; c += 32) left += prow[c] < thr ? prow[c] : 0.0f;
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) left += __shfl_xor_sync(0xffffffffu, left, off);
        if (left * ez * 0.25f * pmax <= 4e-6f) break;
        thr *= 0.25f;
    }
    return thr;
}

// Grid-stride over the list of flagged rows.  Many rows: one warp per row.  Few (at most eight per CTA): one CTA per
// row, its eight warps each taking an eighth of the classes -- a row with a sizeable top probability over a long
// flat tail has hundreds of classes re-scored (four dot products of d per ~2 us trip to L2 for one warp), and a
// request-sized batch waited for its slowest such row: 2048 rows of cfg 3's shape took 131 or 270 us with the
// batch's luck.  Dynamic shared memory: [kFixWarps][c_pad] floats (the rows of P the warps work on; the CTA form
// uses the first) + [kFixWarps][kFixQueue] ints.
__global__ void __launch_bounds__(kFixWarps * 32)
gemm_fixup_sel_kernel(const float* __restrict__ X, int64_t ldx, int d, int ldw, int xal, int C, int c_pad, int64_t out_pitch, int out_col0,
                      const float* __restrict__ W, const float* __restrict__ b,
                      const unsigned* __restrict__ list, const unsigned* __restrict__ n_list,
                      const float* __restrict__ lse, const float* __restrict__ ezr, float* proba,
                      float* decision, int32_t* label, float* maxprob, unsigned long long* rescored,
                      unsigned long long* status) {
    extern __shared__ float fix_smem[];
    __shared__ float s_f[4][kFixWarps];
    __shared__ int s_best[kFixWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int* queue = reinterpret_cast<int*>(fix_smem + (size_t)kFixWarps * c_pad) + warp * kFixQueue;
    const unsigned n = *n_list;
    if (n <= 8u * gridDim.x) {
        // ---- one CTA per row ----
        float* prow = fix_smem;
        const int per = ((C + kFixWarps - 1) / kFixWarps + 31) / 32 * 32;      // classes per warp, whole 32-lane sweeps
        for (unsigned li = blockIdx.x; li < n; li += gridDim.x) {
            const int64_t row = list[li];
            const float* x = X + row * ldx;
            const float L = lse[row];
            if (!finite_f(L)) {                          // (see the warp form below)
                if (warp == 0) {
                    score_row_warp_pitched(x, row, lane, d, ldw, C, W, b, proba, decision, label, maxprob, status, out_pitch, out_col0);
                    if (lane == 0) atomicAdd(rescored, 1ull);
                }
                continue;
            }
            float* pr = proba + row * out_pitch + out_col0;
            float pmax = 0.0f;
            for (int c = threadIdx.x; c < C; c += kFixWarps * 32) {
                const float v = __ldcg(pr + c);
                prow[c] = v;
                pmax = fmaxf(pmax, v);
            }
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) pmax = fmaxf(pmax, __shfl_xor_sync(0xffffffffu, pmax, off));
            if (lane == 0) s_f[0][warp] = pmax;
            __syncthreads();
#pragma unroll
            for (int w = 0; w < kFixWarps; ++w) pmax = fmaxf(pmax, s_f[0][w]);
            const float thr = fix_row_threshold(prow, C, pmax, ezr[row], lane);      // (every warp: the same scan)
            __syncthreads();                             // nobody overwrites prow before every warp has its threshold
            FixPart part;
            const int c_lo = warp * per, c_hi = min(C, (warp + 1) * per);
            if (c_lo < c_hi)
                fix_row_classes(x, d, ldw, xal != 0, W, b, prow, queue, c_lo, c_hi, thr, L,
                                decision ? decision + row * out_pitch + out_col0 : nullptr, lane, part);
            if (lane == 0) {
                s_f[1][warp] = part.old_mass; s_f[2][warp] = part.new_mass; s_f[3][warp] = part.best_z;
                s_f[0][warp] = part.best_q; s_best[warp] = part.best;
            }
            __syncthreads();
            float old_mass = 0.0f, new_mass = 0.0f, best_z = -CUDART_INF_F, best_q = 0.0f;
            int best = 0;
#pragma unroll
            for (int w = 0; w < kFixWarps; ++w) {        // ascending class ranges: the first maximum wins
                old_mass += s_f[1][w];
                new_mass += s_f[2][w];
                if (s_f[3][w] > best_z) { best_z = s_f[3][w]; best = s_best[w]; best_q = s_f[0][w]; }
            }
            const float T = 1.0f - old_mass + new_mass;             // new row sum in units of the old one
            const float invT = __frcp_rn(T);
            const float sc = fabsf(T - 1.0f) > 1e-7f ? invT : 1.0f;
            for (int c = threadIdx.x; c < C; c += kFixWarps * 32) __stcs(pr + c, prow[c] * sc);
            if (threadIdx.x == 0) {
                if (label) label[row] = best;
                if (maxprob) maxprob[row] = best_q * invT;
                atomicAdd(rescored, 1ull);
            }
            __syncthreads();
        }
        return;
    }
    // ---- one warp per row ----
    float* prow = fix_smem + (size_t)warp * c_pad;
    for (unsigned li = blockIdx.x * kFixWarps + warp; li < n; li += gridDim.x * kFixWarps) {
        const int64_t row = list[li];
        const float* x = X + row * ldx;
        const float L = lse[row];
        if (!finite_f(L)) {
            // the tensor pass saw NaN/inf here (bad input, or an fp16 overflow of a finite one):
            // score the whole row in fp32, which also settles whether the input really is non-finite
            score_row_warp_pitched(x, row, lane, d, ldw, C, W, b, proba, decision, label, maxprob, status, out_pitch, out_col0);
            if (lane == 0) atomicAdd(rescored, 1ull);
            continue;
        }
        float* pr = proba + row * out_pitch + out_col0;
        float pmax = 0.0f;
        for (int c = lane; c < C; c += 32) {             // independent loads: one trip to memory for the row
            const float v = __ldcg(pr + c);
            prow[c] = v;
            pmax = fmaxf(pmax, v);
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) pmax = fmaxf(pmax, __shfl_xor_sync(0xffffffffu, pmax, off));
        __syncwarp();
        const float thr = fix_row_threshold(prow, C, pmax, ezr[row], lane);
        FixPart part;
        fix_row_classes(x, d, ldw, xal != 0, W, b, prow, queue, 0, C, thr, L,
                        decision ? decision + row * out_pitch + out_col0 : nullptr, lane, part);
        const float T = 1.0f - part.old_mass + part.new_mass;   // new row sum in units of the old one
        const float invT = __frcp_rn(T);
        const float sc = fabsf(T - 1.0f) > 1e-7f ? invT : 1.0f;
        for (int c = lane; c < C; c += 32) __stcs(pr + c, prow[c] * sc);
        if (lane == 0) {
            if (label) label[row] = part.best;
            if (maxprob) maxprob[row] = part.best_q * invT;
            atomicAdd(rescored, 1ull);
        }
        __syncwarp();
    }
}

// Same purpose when the caller asked for no probability matrix (labels / max probability only, the
// /predict pattern): there is no row of P to pick classes from, so a whole CTA re-scores every class
// of the flagged row, eight warps splitting the classes, four classes per step.
__global__ void __launch_bounds__(256)
gemm_fixup_full_kernel(const float* __restrict__ X, int64_t ldx, int d, int ldw, int xal, int C, int64_t out_pitch, int out_col0,
                       const float* __restrict__ W, const float* __restrict__ b,
                       const unsigned* __restrict__ list, const unsigned* __restrict__ n_list,
                       float* decision, int32_t* label, float* maxprob,
                       unsigned long long* rescored, unsigned long long* status) {
    __shared__ float s_m[8], s_s[8];
    __shared__ int s_a[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned n = *n_list;
    for (unsigned li = blockIdx.x; li < n; li += gridDim.x) {
        const int64_t row = list[li];
        const float* x = X + row * ldx;
        float m = -CUDART_INF_F, ssum = 0.0f;
        int arg = 0;
        const int per = (C + 7) / 8;
        const int c_hi = min(C, (warp + 1) * per);
        for (int c = warp * per; c < c_hi; c += 4) {
            int cls[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) cls[u] = min(c + u, c_hi - 1);
            float z4[4];
            warp_dot4_x4(x, W, cls, d, ldw, xal != 0, lane, z4);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (c + u >= c_hi) break;
                const float z = z4[u] + __ldg(b + c + u);
                if (decision && lane == 0) decision[row * out_pitch + out_col0 + c + u] = z;
                if (z > m) { ssum = ssum * expf(m - z) + 1.0f; m = z; arg = c + u; }
                else ssum += expf(z - m);
            }
        }
        if (lane == 0) { s_m[warp] = m; s_s[warp] = ssum; s_a[warp] = arg; }
        __syncthreads();
        if (threadIdx.x == 0) {
            float gm = -CUDART_INF_F;
            int ga = 0;
            for (int w = 0; w < 8; ++w) if (s_m[w] > gm) { gm = s_m[w]; ga = s_a[w]; }
            float gs = 0.0f;
            for (int w = 0; w < 8; ++w) if (s_s[w] > 0.0f) gs += s_s[w] * expf(s_m[w] - gm);
            if (label) label[row] = ga;
            if (maxprob) maxprob[row] = __frcp_rn(gs);
            atomicAdd(rescored, 1ull);
            if (!(finite_f(gm) && finite_f(gs))) atomicAdd(status, 1ull);
        }
        __syncthreads();
    }
}

}  // namespace
}  // namespace lrs
