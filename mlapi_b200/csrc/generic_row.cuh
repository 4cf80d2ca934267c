// One warp scores one row in plain fp32 FMA arithmetic, any (d, C), any alignment: the body of the
// generic path, shared with the tensor-core path's near-tie re-scoring pass.
// Reference arithmetic: SK/linear_model/_base.py:390-391 (scores), :415-418 (labels), :437-440
// (sigmoid arm), SK/utils/extmath.py:1012-1024 (softmax arm).
#pragma once
#include "common.cuh"

namespace lrs {

#ifdef __CUDACC__
constexpr int kGenericCC = 4;   // score columns per sweep over the row

// `pitch` floats between rows of proba / decision, `col0` the first column this model's classes occupy (a class
// group of a larger model; pitch = C, col0 = 0 otherwise); `ldw` floats between rows of W (>= d)
__device__ __forceinline__ void score_row_warp_pitched(const float* __restrict__ x, int64_t row, int lane, int d, int ldw, int C,
                                                       const float* __restrict__ W, const float* __restrict__ b,
                                                       float* __restrict__ proba, float* __restrict__ decision,
                                                       int32_t* __restrict__ label, float* __restrict__ maxprob,
                                                       unsigned long long* status, int64_t pitch, int col0) {
    constexpr int CC = kGenericCC;
    float run_max = -CUDART_INF_F, run_sum = 0.0f;
    int arg = 0;
    bool ok = true;
    for (int c0 = 0; c0 < C; c0 += CC) {
        float acc[CC];
#pragma unroll
        for (int i = 0; i < CC; ++i) acc[i] = 0.0f;
        for (int k = lane; k < d; k += 32) {
            const float xv = __ldg(x + k);
#pragma unroll
            for (int i = 0; i < CC; ++i)
                if (c0 + i < C) acc[i] = fmaf(xv, __ldg(W + (int64_t)(c0 + i) * ldw + k), acc[i]);
        }
#pragma unroll
        for (int i = 0; i < CC; ++i)
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], off);
#pragma unroll
        for (int i = 0; i < CC; ++i) {
            const int c = c0 + i;
            if (c < C) {
                const float z = acc[i] + __ldg(b + c);
                ok = ok && finite_f(z);
                if (C == 1) {
                    if (lane == 0) {
                        const float2 p = sigmoid_pair(z);
                        if (proba) reinterpret_cast<float2*>(proba)[row] = p;
                        if (decision) decision[row] = z;
                        if (label) label[row] = z > 0.0f ? 1 : 0;
                        if (maxprob) maxprob[row] = fmaxf(p.x, p.y);
                    }
                } else {
                    if (lane == 0) {
                        if (decision) decision[row * pitch + col0 + c] = z;
                        if (proba) proba[row * pitch + col0 + c] = z;     // normalised below
                    }
                    if (z > run_max) {                          // strict: first maximum wins
                        run_sum = run_sum * expf(run_max - z) + 1.0f;
                        run_max = z;
                        arg = c;
                    } else {
                        run_sum += expf(z - run_max);
                    }
                }
            }
        }
    }
    if (!ok && lane == 0 && status) atomicAdd(status, 1ull);
    if (C > 1) {
        const float inv = __frcp_rn(run_sum);
        if (lane == 0) {
            if (label) label[row] = arg;
            if (maxprob) maxprob[row] = inv;
        }
        if (proba) {
            __syncwarp();
            for (int c = lane; c < C; c += 32) {
                const float z = proba[row * pitch + col0 + c];
                proba[row * pitch + col0 + c] = expf(z - run_max) * inv;
            }
        }
    }
}
__device__ __forceinline__ void score_row_warp(const float* __restrict__ x, int64_t row, int lane, int d, int C,
                                               const float* __restrict__ W, const float* __restrict__ b,
                                               float* __restrict__ proba, float* __restrict__ decision,
                                               int32_t* __restrict__ label, float* __restrict__ maxprob,
                                               unsigned long long* status) {
    score_row_warp_pitched(x, row, lane, d, d, C, W, b, proba, decision, label, maxprob, status, C, 0);
}
#endif

}  // namespace lrs
