// Shared host/device helpers for liblrscore (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include <atomic>
#include <string>
#include <vector>

#include "../../include/lrscore.h"

namespace lrs {

// ---- error plumbing ------------------------------------------------------------------------
void set_error(const std::string& msg);
int fail(int code, const std::string& msg);

#define LRS_CUDA(expr)                                                                         \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return ::lrs::fail(LR_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
    } while (0)

extern std::atomic<uint64_t> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

// ---- model -----------------------------------------------------------------------------------
struct HostStage;   // api.cu
struct GemmPlan;    // gemm_tcgen05.cu

struct Model {
    int device = 0;
    int n_scores = 0;     // coef_.shape[0]
    int n_features = 0;
    int n_out = 0;        // max(n_scores, 2): columns of a probability row
    int num_sms = 148;
    int forced_path = LR_PATH_AUTO;

    float* W = nullptr;   // n_scores x n_features fp32 (device)
    float* b = nullptr;   // n_scores
    // narrow path: W zero-padded to [ns_pad][d_pad], intercept padded with -inf
    float* Wn = nullptr;
    float* bn = nullptr;
    int narrow_ns = 0, narrow_dpad = 0, narrow_lanes = 0;
    bool narrow_bulk = false;             // opt-in (LRS_NARROW_BULK=1): cp.async.bulk ring, measured 4% slower than plain loads
    int64_t narrow_bulk_min_rows = 65536;
    // status words: [0] rows that scored to a non-finite value, [1] rows re-scored in fp32 (tensor-core path,
    // running), [2] rows the tensor-core path scored in 3xTF32 because 3xFP16 would have lost precision (running)
    unsigned long long* status = nullptr;
    std::atomic<unsigned long long> rescored_taken{0};   // host shadows for lr_model_take_rescored / _take_fallback
    std::atomic<unsigned long long> fallback_taken{0};

    GemmPlan* gemm = nullptr;
    // more than 2048 classes: class groups of equal size, each an ordinary tensor-path model over a slice of W / b
    // (which they do not own); see gemm_groups_launch
    std::vector<Model*> groups;
    int group_classes = 0;
    HostStage* stage = nullptr;
    int64_t host_chunk_rows = 0;
};

struct ScoreArgs {
    const float* X;
    int64_t n_rows;
    int64_t ldx;
    float* proba;
    float* decision;
    int32_t* label;
    float* maxprob;
    cudaStream_t stream;
    // a model that is one class group of a larger one (more than 2048 classes, gemm_groups_launch): where its columns
    // sit in the caller's proba / decision rows, and where the log-sum-exp of its own classes goes
    int64_t out_pitch = 0;      // 0: n_scores
    int out_col0 = 0;
    float* lse_out = nullptr;
};

// per-family entry points (each returns LR_OK / error code, sets last error)
bool narrow_supports(const Model& m);
int narrow_prepare(Model& m, const float* W_host, const float* b_host);
int narrow_launch(const Model& m, const ScoreArgs& a);

bool wide_supports(const Model& m, const ScoreArgs& a);
int wide_launch(const Model& m, const ScoreArgs& a);

int generic_launch(const Model& m, const ScoreArgs& a);

bool gemm_supports(const Model& m, const ScoreArgs& a);
int gemm_prepare(Model& m, const double* coef, const double* intercept);
int gemm_load(Model& m, const double* coef, const double* intercept);
void gemm_destroy(Model& m);
void gemm_feedback(Model& m, int64_t rows, uint64_t rescored_total);
int gemm_active_layout(const Model& m);
int gemm_launch(const Model& m, const ScoreArgs& a);
bool gemm_short_rows(const Model& m, const ScoreArgs& a);
int gemm_groups_prepare(Model& m, const double* coef, const double* intercept);    // > 2048 classes
int gemm_groups_load(Model& m, const double* coef, const double* intercept);
int gemm_groups_launch(const Model& m, const ScoreArgs& a);
void gemm_groups_destroy(Model& m);   // 13..32 classes, dense rows of <= 64 floats: the short-row tensor form

// ---- device helpers ----------------------------------------------------------------------------
#ifdef __CUDACC__

__device__ __forceinline__ float4 ldg_stream4(const float* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ float ldg_stream1(const float* p) {
    float v;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}

// Binary arm of predict_proba (SK/linear_model/_base.py:437-440): returns (1-p, p) with
// p = expit(z), evaluated from e = exp(-|z|) so that both columns keep full relative accuracy and
// saturate to exactly 0/1 like scipy's expit does in float64.
__device__ __forceinline__ float2 sigmoid_pair(float z) {
    float e = expf(-fabsf(z));
    float big = __frcp_rn(1.0f + e);   // probability of the predicted side, >= 0.5
    float small = e * big;
    return z > 0.0f ? make_float2(small, big) : make_float2(big, small);
}

// ---- mbarrier + bulk-copy (TMA engine) primitives -----------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// (A suspend-time hint on try_wait -- up to 20 us -- was tried to stop idle warps from polling: ncu counts 41M polls
// per 400k rows in the GEMM kernel, a fifth of its instructions.  The poll count did not change on this driver.)
// Bounded wait: a pipeline bug traps (-> cudaErrorLaunchFailure) after a few seconds instead of hanging
// the GPU.  The retry loop lives inside the asm block on purpose: a C++ loop with a per-lane exit
// would make the surrounding control flow divergent in the compiler's eyes, and every value computed
// after it (stage indices, shared-memory descriptors) would lose its place in uniform registers.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .u32 n;\n\t"
        "mov.u32 n, 0;\n"
        "LRS_WAIT:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra LRS_DONE;\n\t"
        "add.u32 n, n, 1;\n\t"
        "setp.lt.u32 p, n, 50000000;\n\t"
        "@p bra LRS_WAIT;\n\t"
        "trap;\n"
        "LRS_DONE:\n\t"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D bulk copy global -> shared, completion counted in bytes on `bar`.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// "Transpose reduce" over the LANES lanes that share a row (LANES = 2 * OFF): each lane enters with
// partial sums of U rows (x NS score columns) and leaves owning the complete sums of U / min(U, LANES)
// distinct rows in acc[0..]; see narrow.cu for the mapping.
template <int OFF, int ROWS, int NS, int U>
__device__ __forceinline__ void lane_reduce(float (&acc)[U][NS], int s) {
    if constexpr (OFF >= 1) {
        if constexpr (ROWS > 1) {
            constexpr int H = ROWS / 2;
            const bool upper = (s & OFF) != 0;
#pragma unroll
            for (int r = 0; r < H; ++r) {
#pragma unroll
                for (int c = 0; c < NS; ++c) {
                    const float keep = upper ? acc[r + H][c] : acc[r][c];
                    const float send = upper ? acc[r][c] : acc[r + H][c];
                    acc[r][c] = keep + __shfl_xor_sync(0xffffffffu, send, OFF);
                }
            }
            lane_reduce<OFF / 2, H, NS, U>(acc, s);
        } else {
#pragma unroll
            for (int c = 0; c < NS; ++c) acc[0][c] += __shfl_xor_sync(0xffffffffu, acc[0][c], OFF);
            lane_reduce<OFF / 2, 1, NS, U>(acc, s);
        }
    }
}

__device__ __forceinline__ bool finite_f(float z) { return fabsf(z) <= 3.402823466e38f; }

#endif  // __CUDACC__

}  // namespace lrs
