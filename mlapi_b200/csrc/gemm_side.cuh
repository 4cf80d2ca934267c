// Side kernels of the tensor-core path: per-call batch statistics, per-model balance / scale / split of W.
// One translation unit with gemm_tcgen05.cu (included there, in this order); not a stand-alone header.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "generic_row.cuh"

namespace lrs {
namespace {

// Per-call look at the batch (every arithmetic): up to 2048 rows spread over it are read once, each feature
// multiplied by the model's per-feature power of two (w_colscale_kernel), and per row the largest magnitude and
// the 2-norm are kept.  From those the last CTA to arrive derives
//   out[0], out[1]  the power-of-two scale for X and its inverse: the row maximum that all but ~0.2 % of the sampled
//                   rows stay under lands in [16, 64].  fp16 then holds hi + lo halves at full precision down to
//                   ~2^-9 of that maximum and has ~2^10 of headroom above it; a row holding something larger still
//                   overflows to inf, is seen as non-finite by the epilogue and re-scored in fp32 -- so one outlier
//                   row costs itself a second pass, not every other row its low half;
//   out[6]          the logit error fp16's subnormal floor (quantum q = 2^-24 in scaled units) can cause:
//                   q / x_scale * max_c ||W'_c||_2 from the X side + q * max_c(1 / class scale) * ||x'||_2 from the W
//                   side (rounding errors of the d terms add like a random walk; q is 3.5 sigma of one);
//   out[4]          the arithmetic of this call: 3xFP16 (bits 0) while that error stays under 2e-6, else 3xTF32
//                   (bits 1: fp32's exponent range, half the tensor rate) -- `force` 1 / 2 pins FP16 / TF32;
//   out[5]          the size of a typical partial sum, ||x'||_2 * max_c ||W'_c||_2 / sqrt(d): what the truncating
//                   accumulation error is proportional to when it exceeds the logit itself.
// out[3] is the arrival ticket, zeroed by the caller.  rowstat: [2][kMaxSample] floats of scratch.
constexpr int kXScaleBlocks = 128;
constexpr int kMaxSample = 2048;
constexpr int kStatBins = 2048;          // float bits >> 20: 8 exponent bits + 3 mantissa bits (12 % steps)
template <bool VEC>       // VEC: rows are 16-byte aligned and d % 4 == 0 (128-bit loads); else 4-byte loads
__global__ void __launch_bounds__(256)
x_scale_kernel(const float* __restrict__ X, int64_t n_rows, int64_t ldx, int d, const float* __restrict__ col_scale,
               float w_norm_max, float inv_cscale_max, int force, float* __restrict__ out, float* __restrict__ rowstat) {
    __shared__ unsigned hist[2][kStatBins];
    __shared__ unsigned part[2][256];
    __shared__ float q_s[2];
    __shared__ bool last;
    // up to 2048 sampled rows, fewer for long rows (about 4 MB read in all, at least 64 rows)
    int64_t n_s = (int64_t)(1 << 20) / d;
    n_s = n_s < 64 ? 64 : (n_s > kMaxSample ? kMaxSample : n_s);
    n_s = n_rows < n_s ? n_rows : n_s;
    const int64_t step = n_rows / n_s;
    const int lane = threadIdx.x & 31;
    for (int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); i < n_s; i += (int64_t)gridDim.x * 8) {
        float mx = 0.0f, ss = 0.0f;
        if constexpr (VEC) {
            const float4* x = reinterpret_cast<const float4*>(X + (i * step) * ldx);
            const float4* c4 = reinterpret_cast<const float4*>(col_scale);
#pragma unroll 4
            for (int k = lane; k < d / 4; k += 32) {
                const float4 v = __ldg(x + k), c = __ldg(c4 + k);
                const float a[4] = {fabsf(v.x * c.x), fabsf(v.y * c.y), fabsf(v.z * c.z), fabsf(v.w * c.w)};
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (a[j] <= 1.0e18f) {            // NaN / inf (and what would overflow the sum of squares) are left out
                        mx = fmaxf(mx, a[j]);
                        ss = fmaf(a[j], a[j], ss);
                    }
            }
        } else {
            const float* x = X + (i * step) * ldx;
            for (int k = lane; k < d; k += 32) {
                const float a = fabsf(__ldg(x + k) * __ldg(col_scale + k));
                if (a <= 1.0e18f) {
                    mx = fmaxf(mx, a);
                    ss = fmaf(a, a, ss);
                }
            }
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
            ss += __shfl_xor_sync(0xffffffffu, ss, off);
        }
        if (lane == 0) {
            rowstat[i] = mx;
            rowstat[kMaxSample + i] = sqrtf(ss);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        last = atomicAdd(reinterpret_cast<unsigned*>(out) + 3, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    // The value that all but n_s / 512 of the sampled rows stay under, for both statistics: a histogram over
    // the top 11 bits of the (non-negative) floats, scanned from the top; the answer is the bin's upper edge.
    for (int i = threadIdx.x; i < 2 * kStatBins; i += 256) (&hist[0][0])[i] = 0u;
    __syncthreads();
    for (int i = threadIdx.x; i < (int)n_s; i += 256) {
        atomicAdd(&hist[0][__float_as_uint(__ldcg(rowstat + i)) >> 20], 1u);
        atomicAdd(&hist[1][__float_as_uint(__ldcg(rowstat + kMaxSample + i)) >> 20], 1u);
    }
    __syncthreads();
    constexpr int PER = kStatBins / 256;
    for (int w = 0; w < 2; ++w) {
        unsigned sum = 0;
        for (int j = 0; j < PER; ++j) sum += hist[w][threadIdx.x * PER + j];
        part[w][threadIdx.x] = sum;
    }
    __syncthreads();
    if (threadIdx.x < 2) {
        const int w = threadIdx.x;
        const unsigned drop = (unsigned)(n_s / 512);
        unsigned above = 0;
        int t = 255;
        while (t > 0 && above + part[w][t] <= drop) above += part[w][t--];
        int bin = t * PER + PER - 1;
        while (bin > t * PER && above + hist[w][bin] <= drop) above += hist[w][bin--];
        // all-zero statistics end in bin 0: the edge 2^-126 * ... is harmless (scale clamps below)
        q_s[w] = fminf(__uint_as_float((unsigned)(bin + 1) << 20), 3.0e38f);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const float qmax = q_s[0], qnorm = q_s[1];
        int e = 0;
        frexpf(qmax, &e);                                 // qmax = f * 2^e, f in [0.5, 1)
        e = 6 - e;                                        // qmax * 2^e in [32, 64)
        e = e > 60 ? 60 : (e < -60 ? -60 : e);            // col_scale is clamped likewise: products stay normal
        const float sc = ldexpf(1.0f, e);
        const float q = 5.9604645e-8f;                    // 2^-24
        float floor_err = q / sc * w_norm_max + q * inv_cscale_max * qnorm;
        if (!(floor_err <= 3.0e38f)) floor_err = 3.0e38f;
        const unsigned mode = force == 1 ? 0u : (force == 2 ? 1u : (floor_err > 2.0e-6f ? 1u : 0u));
        out[0] = sc;
        out[1] = 1.0f / sc;
        reinterpret_cast<unsigned*>(out)[4] = mode;
        out[5] = qnorm * w_norm_max * rsqrtf((float)d);
        out[6] = mode == 0u ? floor_err : 0.0f;
    }
}

// Per-feature balance of the FP16 operands, from W alone (one thread per feature): column k of W is multiplied
// by 2^e_k with e_k = -floor(log2 max_c |W[c][k]|), so that every column of the balanced W' peaks in [1, 2),
// and the converter multiplies feature k of X by 2^-e_k.  Powers of two: the products are unchanged to the
// bit.  A model fitted on unstandardised features has weights inversely proportional to the feature scales
// (REF/Logistic Regression.ipynb:28-34 fits raw centimetres), so this also brings the features of X to one
// scale -- without it a feature 2^10 below the largest one loses its fp16 `lo` half to subnormals.
// x_mul[k] = 2^-e_k (0 for an all-zero column: the feature cannot matter), w_mul[k] = 2^e_k.
__global__ void __launch_bounds__(256)
w_colscale_kernel(const double* __restrict__ coef, int C, int d, float* __restrict__ x_mul, double* __restrict__ w_mul) {
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= d) return;
    double mx = 0.0;
    for (int c = 0; c < C; ++c) mx = fmax(mx, fabs(coef[(size_t)c * d + k]));
    int e = 0;
    if (mx > 0.0 && mx <= 1.7976931348623157e308) {
        frexp(mx, &e);                    // mx = f * 2^e, f in [0.5, 1)
        e = 1 - e;                        // mx * 2^e in [1, 2)
        e = e > 60 ? 60 : (e < -60 ? -60 : e);
    }
    x_mul[k] = mx == 0.0 ? 0.0f : (float)ldexp(1.0, -e);
    w_mul[k] = ldexp(1.0, e);
}

// FP16 operands of W, made on the device: class c (one CTA) of the column-balanced W' is scaled by the power of
// two that puts its largest magnitude in [2^14, 2^15) -- the top of fp16's range, the maximum being known
// exactly: entries down to 2^-17 of it keep a normal `lo` half -- and split into fp16 hi + lo halves from the
// float64 value; 1/scale goes to inv[c], the 2-norm of the balanced (unscaled) row to norm[c].
__global__ void __launch_bounds__(256)
w_split16_kernel(const double* __restrict__ coef, const double* __restrict__ w_mul, int d, int d16,
                 __half* __restrict__ hi, __half* __restrict__ lo, float* __restrict__ inv, float* __restrict__ norm) {
    __shared__ double red[8], red2[8];
    const int c = blockIdx.x;
    const double* w = coef + (size_t)c * d;
    double mx = 0.0, ss = 0.0;
    for (int k = threadIdx.x; k < d; k += 256) {
        const double v = w[k] * w_mul[k];
        mx = fmax(mx, fabs(v));      // fmax drops NaN, keeps inf
        ss += v * v;
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, off));
        ss += __shfl_xor_sync(0xffffffffu, ss, off);
    }
    if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5] = mx; red2[threadIdx.x >> 5] = ss; }
    __syncthreads();
    mx = red[0];
    ss = red2[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) { mx = fmax(mx, red[i]); ss += red2[i]; }
    double sc = 1.0;
    if (mx > 0.0 && mx <= 1.7976931348623157e308) {
        int e = 0;
        frexp(mx, &e);
        sc = ldexp(1.0, 15 - e);
    }
    if (threadIdx.x == 0) { inv[c] = (float)(1.0 / sc); norm[c] = (float)sqrt(ss); }
    for (int k = threadIdx.x; k < d; k += 256) {
        const double v = w[k] * w_mul[k] * sc;
        const __half h = __float2half_rn((float)v);
        hi[(size_t)c * d16 + k] = h;
        lo[(size_t)c * d16 + k] = __float2half_rn((float)(v - (double)__half2float(h)));
    }
}

// TF32 operands of W (hi = round-to-nearest TF32, lo = TF32 of the residual), from the float64 coefficients.
__global__ void __launch_bounds__(256)
w_split_tf32_kernel(const double* __restrict__ coef, int d, int d4, float* __restrict__ hi, float* __restrict__ lo) {
    const double* w = coef + (size_t)blockIdx.x * d;
    for (int k = threadIdx.x; k < d; k += 256) {
        const float h = to_tf32((float)w[k]);
        hi[(size_t)blockIdx.x * d4 + k] = h;
        lo[(size_t)blockIdx.x * d4 + k] = to_tf32((float)(w[k] - (double)h));
    }
}

}  // namespace
}  // namespace lrs
