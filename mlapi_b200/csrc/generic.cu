// Generic path: any (n_scores, n_features), any alignment or row pitch.  One warp per row, score
// columns in groups of four, scalar coalesced loads; softmax by an online (running max / running
// sum) pass followed by an in-place normalisation of the row the warp has just written.  It is the
// shape-complete safety net under the tuned paths (narrow / wide / gemm), not a roofline kernel:
// X is re-read from L1/L2 once per group of four score columns.
//
// Same reference arithmetic as the other paths: SK/linear_model/_base.py:390-391 (scores),
// :415-418 (labels), :437-440 (sigmoid arm), SK/utils/extmath.py:1012-1024 (softmax arm).
#include "generic_row.cuh"

namespace lrs {
namespace {

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads)
generic_kernel(const float* __restrict__ X, int64_t n_rows, int64_t ldx, int d, int C,
               const float* __restrict__ W, const float* __restrict__ b,
               float* __restrict__ proba, float* __restrict__ decision,
               int32_t* __restrict__ label, float* __restrict__ maxprob,
               unsigned long long* status) {
    const int lane = threadIdx.x & 31;
    const int64_t warps_total = (int64_t)gridDim.x * (kThreads / 32);
    for (int64_t row = (int64_t)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5); row < n_rows;
         row += warps_total)
        score_row_warp(X + row * ldx, row, lane, d, C, W, b, proba, decision, label, maxprob, status);
}

}  // namespace

int generic_launch(const Model& m, const ScoreArgs& a) {
    const int64_t need = (a.n_rows + (kThreads / 32) - 1) / (kThreads / 32);
    const int64_t cap = (int64_t)m.num_sms * 8;
    const int grid = (int)(need < cap ? need : cap);
    generic_kernel<<<grid, kThreads, 0, a.stream>>>(a.X, a.n_rows, a.ldx, m.n_features, m.n_scores,
                                                    m.W, m.b, a.proba, a.decision, a.label,
                                                    a.maxprob, m.status);
    count_launch();
    LRS_CUDA(cudaGetLastError());
    return LR_OK;
}

}  // namespace lrs
