// Wide binary path (d in the thousands and up, one score column; BASELINE cfg 2: d = 100 000).
//
// Replaces sklearn's decision_function + _predict_proba_lr + predict for a (1,d) coef_
// (SK/linear_model/_base.py:390-391, :415-418, :437-440).
//
// A row no longer fits a warp's registers, so a CTA streams R rows at once: in each sweep step a
// thread loads one 128-bit piece of W (hot in L1/L2: W is d*4 bytes, 400 KB at d=100k) and the
// matching 128-bit pieces of R different rows of X (R independent HBM requests in flight per
// thread, 2R with the two-step form used from d = 2048 up), so W's L2 traffic is 1/R of X's HBM
// traffic.  Partial sums go through a warp-shuffle and
// one shared-memory step; the sigmoid epilogue is fused.
//
// HBM roofline: algorithmic bytes per row = 4*d read + 8 written (proba) -- 400 008 B at d=100k.
#include "common.cuh"

namespace lrs {
namespace {

constexpr int kThreads = 256;

template <int R, int UN>
__global__ void __launch_bounds__(kThreads, UN == 2 ? 2 : 3)   // UN = 1: <= 85 registers, 3 CTAs per SM, 96 KB of loads in flight
wide_kernel(const float* __restrict__ X, int64_t n_rows, int64_t ldx, int d,
            const float* __restrict__ W, const float* __restrict__ b,
            float* __restrict__ proba, float* __restrict__ decision, int32_t* __restrict__ label,
            float* __restrict__ maxprob, unsigned long long* status) {
    __shared__ float part[kThreads / 32][R];

    for (int64_t blk = blockIdx.x; blk * R < n_rows; blk += gridDim.x) {
        const int64_t row0 = blk * R;
        const float* xr[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            int64_t row = row0 + r;
            if (row >= n_rows) row = n_rows - 1;      // clamp: re-reads a valid row, result dropped
            xr[r] = X + row * ldx;
        }
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = 0.0f;

        const int d4 = d >> 2;
        int k = threadIdx.x;
        if constexpr (UN == 2) {
            // two sweep steps of all R rows requested before the first FMA
            for (; k + kThreads < d4; k += 2 * kThreads) {
                float4 x0[R], x1[R];
#pragma unroll
                for (int r = 0; r < R; ++r) x0[r] = ldg_stream4(xr[r] + 4 * k);
#pragma unroll
                for (int r = 0; r < R; ++r) x1[r] = ldg_stream4(xr[r] + 4 * (k + kThreads));
                const float4 w0 = __ldg(reinterpret_cast<const float4*>(W) + k);
                const float4 w1 = __ldg(reinterpret_cast<const float4*>(W) + k + kThreads);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    acc[r] = fmaf(x0[r].x, w0.x, acc[r]);
                    acc[r] = fmaf(x0[r].y, w0.y, acc[r]);
                    acc[r] = fmaf(x0[r].z, w0.z, acc[r]);
                    acc[r] = fmaf(x0[r].w, w0.w, acc[r]);
                }
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    acc[r] = fmaf(x1[r].x, w1.x, acc[r]);
                    acc[r] = fmaf(x1[r].y, w1.y, acc[r]);
                    acc[r] = fmaf(x1[r].z, w1.z, acc[r]);
                    acc[r] = fmaf(x1[r].w, w1.w, acc[r]);
                }
            }
        }
        for (; k < d4; k += kThreads) {
            float4 x[R];
#pragma unroll
            for (int r = 0; r < R; ++r) x[r] = ldg_stream4(xr[r] + 4 * k);
            const float4 w = __ldg(reinterpret_cast<const float4*>(W) + k);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                acc[r] = fmaf(x[r].x, w.x, acc[r]);
                acc[r] = fmaf(x[r].y, w.y, acc[r]);
                acc[r] = fmaf(x[r].z, w.z, acc[r]);
                acc[r] = fmaf(x[r].w, w.w, acc[r]);
            }
        }
        // tail when d is not a multiple of 4
        {
            const int k = (d4 << 2) + threadIdx.x;
            if (k < d) {
                const float w = __ldg(W + k);
#pragma unroll
                for (int r = 0; r < R; ++r) acc[r] = fmaf(ldg_stream1(xr[r] + k), w, acc[r]);
            }
        }

#pragma unroll
        for (int r = 0; r < R; ++r) {
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) acc[r] += __shfl_xor_sync(0xffffffffu, acc[r], off);
        }
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        if (lane == 0) {
#pragma unroll
            for (int r = 0; r < R; ++r) part[warp][r] = acc[r];
        }
        __syncthreads();
        if (threadIdx.x < R) {
            const int64_t row = row0 + threadIdx.x;
            if (row < n_rows) {
                float z = b[0];
#pragma unroll
                for (int w = 0; w < kThreads / 32; ++w) z += part[w][threadIdx.x];
                if (!finite_f(z)) atomicAdd(status, 1ull);
                const float2 p = sigmoid_pair(z);
                if (proba) __stcs(reinterpret_cast<float2*>(proba) + row, p);
                if (decision) __stcs(decision + row, z);
                if (label) __stcs(label + row, z > 0.0f ? 1 : 0);
                if (maxprob) __stcs(maxprob + row, fmaxf(p.x, p.y));
            }
        }
        __syncthreads();   // part[] is reused by the next row block
    }
}

// Rows that are not 16-byte aligned (odd feature count, odd pitch, unaligned base): the same sweep with
// 4-byte loads.  A warp's load is still one contiguous 128-byte span, so the instruction rate is no
// concern (0.2 loads per clock per SM at HBM rate); what matters is bytes in flight, hence four sweep
// steps of all R rows are requested before the first FMA (32 loads per thread).
template <int R>
__global__ void __launch_bounds__(kThreads, 2)
wide_scalar_kernel(const float* __restrict__ X, int64_t n_rows, int64_t ldx, int d,
                   const float* __restrict__ W, const float* __restrict__ b,
                   float* __restrict__ proba, float* __restrict__ decision, int32_t* __restrict__ label,
                   float* __restrict__ maxprob, unsigned long long* status) {
    constexpr int UN = 4;
    __shared__ float part[kThreads / 32][R];
    for (int64_t blk = blockIdx.x; blk * R < n_rows; blk += gridDim.x) {
        const int64_t row0 = blk * R;
        const float* xr[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            int64_t row = row0 + r;
            if (row >= n_rows) row = n_rows - 1;      // clamp: re-reads a valid row, result dropped
            xr[r] = X + row * ldx;
        }
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = 0.0f;
        for (int k = threadIdx.x; k < d; k += kThreads * UN) {
            float x[UN][R], w[UN];
#pragma unroll
            for (int u = 0; u < UN; ++u) {
                const int kk = k + u * kThreads;
                const int kc = kk < d ? kk : 0;       // past the end: column 0 against a zero weight
#pragma unroll
                for (int r = 0; r < R; ++r) x[u][r] = ldg_stream1(xr[r] + kc);
                w[u] = kk < d ? __ldg(W + kk) : 0.0f;
            }
#pragma unroll
            for (int u = 0; u < UN; ++u)
#pragma unroll
                for (int r = 0; r < R; ++r) acc[r] = fmaf(x[u][r], w[u], acc[r]);
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) acc[r] += __shfl_xor_sync(0xffffffffu, acc[r], off);
        }
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        if (lane == 0) {
#pragma unroll
            for (int r = 0; r < R; ++r) part[warp][r] = acc[r];
        }
        __syncthreads();
        if (threadIdx.x < R) {
            const int64_t row = row0 + threadIdx.x;
            if (row < n_rows) {
                float z = b[0];
#pragma unroll
                for (int w = 0; w < kThreads / 32; ++w) z += part[w][threadIdx.x];
                if (!finite_f(z)) atomicAdd(status, 1ull);
                const float2 p = sigmoid_pair(z);
                if (proba) __stcs(reinterpret_cast<float2*>(proba) + row, p);
                if (decision) __stcs(decision + row, z);
                if (label) __stcs(label + row, z > 0.0f ? 1 : 0);
                if (maxprob) __stcs(maxprob + row, fmaxf(p.x, p.y));
            }
        }
        __syncthreads();   // part[] is reused by the next row block
    }
}

template <int R>
int launch_scalar_r(const Model& m, const ScoreArgs& a) {
    const int64_t blocks = (a.n_rows + R - 1) / R;
    const int64_t cap = (int64_t)m.num_sms * 16;
    const int grid = (int)(blocks < cap ? blocks : cap);
    wide_scalar_kernel<R><<<grid, kThreads, 0, a.stream>>>(a.X, a.n_rows, a.ldx, m.n_features, m.W, m.b,
                                                           a.proba, a.decision, a.label, a.maxprob, m.status);
    count_launch();
    LRS_CUDA(cudaGetLastError());
    return LR_OK;
}

template <int R>
int launch_r(const Model& m, const ScoreArgs& a) {
    const int64_t blocks = (a.n_rows + R - 1) / R;
    const int64_t cap = (int64_t)m.num_sms * 16;     // grid-stride beyond this; multiple of the SM count
    const int grid = (int)(blocks < cap ? blocks : cap);
    // Rows of 2048 floats and more: two sweep steps in flight per thread at 2 CTAs per SM (128 KB of loads
    // per SM) -- measured 6.2 -> 7.1 TB/s at d = 4096..20000, cfg 2 (d = 100k) 4.96 -> 4.49 ms; shorter rows
    // have a single step per row and keep the 3-CTA form.  LRS_WIDE_UNROLL=1|2 pins one (tuning).
    static const int forced = [] { const char* e = getenv("LRS_WIDE_UNROLL"); return e ? atoi(e) : 0; }();
    const int un = forced ? forced : (m.n_features / 4 >= 2 * kThreads ? 2 : 1);
    if (un == 2)
        wide_kernel<R, 2><<<grid, kThreads, 0, a.stream>>>(a.X, a.n_rows, a.ldx, m.n_features, m.W, m.b,
                                                           a.proba, a.decision, a.label, a.maxprob, m.status);
    else
        wide_kernel<R, 1><<<grid, kThreads, 0, a.stream>>>(a.X, a.n_rows, a.ldx, m.n_features, m.W, m.b,
                                                           a.proba, a.decision, a.label, a.maxprob, m.status);
    count_launch();
    LRS_CUDA(cudaGetLastError());
    return LR_OK;
}

}  // namespace

bool wide_supports(const Model& m, const ScoreArgs&) { return m.n_scores == 1 && m.n_features >= 4; }

int wide_launch(const Model& m, const ScoreArgs& a) {
    if (!wide_supports(m, a)) return fail(LR_ERR_INVALID, "wide path needs a binary model with at least 4 features");
    // fewer rows per CTA when the batch is too small to give every SM two CTAs
    const int64_t want = 2 * (int64_t)m.num_sms;
    const bool aligned = a.ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(a.X) & 15u) == 0;
    if (!aligned) {
        if (a.n_rows >= 8 * want) return launch_scalar_r<8>(m, a);
        if (a.n_rows >= 2 * want) return launch_scalar_r<2>(m, a);
        return launch_scalar_r<1>(m, a);
    }
    if (a.n_rows >= 8 * want) return launch_r<8>(m, a);
    if (a.n_rows >= 4 * want) return launch_r<4>(m, a);
    if (a.n_rows >= 2 * want) return launch_r<2>(m, a);
    return launch_r<1>(m, a);
}

}  // namespace lrs
