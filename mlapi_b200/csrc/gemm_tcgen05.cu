// GEMM path: multinomial models with many classes (BASELINE cfg 3: d=1024, C=1000) -- the only
// case on this path that is a real dense contraction, so it runs on the 5th-gen tensor cores.
//
// Replaces, for a row tile, sklearn's decision_function (X @ coef_.T + intercept_,
// SK/linear_model/_base.py:390-391), the row softmax of predict_proba
// (SK/utils/extmath.py:1012-1024: subtract row max, exp, divide by row sum) and the argmax of
// predict (SK/linear_model/_base.py:415-418), with the softmax fused into the epilogue so the
// n x C score matrix is written exactly once (as probabilities).
//
// Precision.  tcgen05 has no fp32 MMA; a single reduced-precision pass leaves ~1e-3 error on a d=1024
// logit, far outside the 1e-5 parity bar.  Each product is therefore split,
//     x = x_hi + x_lo,  w = w_hi + w_lo,     x.w ~= x_hi.w_hi + x_lo.w_hi + x_hi.w_lo   (fp32 accumulation in TMEM)
// in fp16 halves by default (3xFP16, kind::f16 at the full tensor rate) or TF32 halves (3xTF32, half the
// rate, fp32's exponent range).  fp16 has 5 exponent bits, so range is managed explicitly: W is balanced
// per feature and scaled per class when the model is loaded (w_colscale_kernel, w_split16_kernel), X is
// scaled per call from a robust row maximum of a sample of the batch (x_scale_kernel), rows that overflow
// are re-scored in fp32, and a per-call bound on what the fp16 subnormal floor can still cost decides on the
// device whether the call runs in 3xFP16 or 3xTF32 (both kernels are enqueued; the one not chosen returns at
// once).  The halves of W are made once per model; those of X on the fly in shared memory by four converter
// warps, so X is read from HBM exactly once.  The dropped x_lo.w_lo term is ~2^-22 relative.  Algorithmic
// flops (2*d*C per row) are counted once; the tensor pipe executes three times that.
//
// Structure (one CTA = 128 rows x up to 512 classes; a cluster of CTAs covers all classes of the same rows
// and exchanges softmax statistics through distributed shared memory; in the 512-class layout two CTAs
// form a cta_group::2 pair that shares every W tile):
//   warp 0      TMA producer, W tiles (hi, lo) into an mbarrier ring
//   warp 1      MMA issuer: one lane issues tcgen05.mma (M=128 or 256 across the pair, N up to 256) into TMEM
//   warps 2-5   converters: raw fp32 X tile -> scaled fp16 hi | lo operand tiles, in place (128B swizzle)
//   warps 6-13  epilogue: two TMEM passes (exp written back in between), one DSMEM exchange per tile,
//               outputs, re-scoring list
//   warp 14     TMA producer, X tiles
// Rows whose answer the accumulation error could change (near-ties, a top probability near one half)
// are put on a list by the epilogue and re-scored in fp32 by a second kernel.
// Models of 13..32 classes on dense rows of at most 64 floats take short_rows_kernel instead.
//
// Tensor roofline: 2*d*C flops per row against the measured 16-bit tensor peak (TF32: half of it).
//
// This file: the host side (operand copies of W, layouts, tensor maps, launches, class groups).  The device code is one
// translation unit with it, in parts:
//   gemm_tc.cuh          tile constants, shared-memory plan, GemmParams, PTX wrappers (mbarrier / TMA / tcgen05 / TMEM)
//   gemm_epilogue.cuh    register epilogues for models of at most 128 classes
//   gemm_kernel.cuh      gemm_softmax_kernel
//   gemm_rescore.cuh     fp32 re-scoring of flagged rows
//   gemm_short_rows.cuh  short_rows_kernel and its re-scoring kernels
//   gemm_side.cuh        batch statistics (x_scale_kernel), balance / scale / split of W
#include <cuda.h>
#include <cuda_fp16.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "generic_row.cuh"
#include "gemm_tc.cuh"
#include "gemm_epilogue.cuh"
#include "gemm_kernel.cuh"
#include "gemm_rescore.cuh"
#include "gemm_short_rows.cuh"
#include "gemm_side.cuh"

namespace lrs {

struct GemmPlan {
    double* coef64 = nullptr;   // [C][d] float64 coef_ as handed in: what every operand copy below is made from
    float* w_hi = nullptr;      // [c_pad][d4] TF32 (round-to-nearest) part of coef_
    float* w_lo = nullptr;      // [c_pad][d4] TF32 of the residual
    float* bias = nullptr;      // [c_pad], -inf beyond n_classes
    // FP16 operands: coef_ balanced per feature (w_colscale_kernel), scaled per class by a power of two to the
    // top of fp16's range and split into two halves (w_split16_kernel)
    __half* w16_hi = nullptr;   // [c_pad][d16]
    __half* w16_lo = nullptr;
    float* col_inv_scale = nullptr;   // [c_pad] 1 / class scale
    float* col_scale = nullptr;       // [d64] per-feature power of two applied to X, zero beyond n_features
    double* w_mul = nullptr;          // [d] its inverse, applied to W
    float* w_norm = nullptr;          // [c_pad] 2-norm of each balanced class row
    float w_norm_max = 0.0f;          // largest of those                          } inputs of the per-call
    float inv_cscale_max = 0.0f;      // largest 1 / class scale                    } precision guard
    float* w32_pad = nullptr;   // [C][d4] fp32 coef_ at a 16-byte pitch (only when n_features % 4 != 0)
    int d4 = 0;                 // n_features rounded up to 4
    int d16 = 0;                // row pitch of the fp16 copies (n_features rounded up to 8: TMA wants 16-byte pitches)
    int d64 = 0;                // n_features rounded up to whole FP16 k-chunks
    // Arithmetic: 0 = per call (3xFP16 unless the batch's dynamic range says 3xTF32), 1 = 3xFP16 always
    // (LRS_GEMM_F16=2, measurements only: no guard), 2 = 3xTF32 always (LRS_GEMM_F16=0)
    int force = 0;
    // Two CTA layouts over the same W buffers:
    //   [0] fast:     512 classes per CTA (CTA pairs), one TMEM accumulator that fills TMEM   (C in 257..2048)
    //   [1] accurate: 256 classes per CTA, correction products accumulate apart, 3x smaller logit
    //                 error, so far fewer rows need the fp32 re-scoring pass                   (C <= 1024)
    // `active` starts on the fast layout where both exist and moves to the accurate one when the re-scoring
    // pass reports that the data is confident (lr_score_host feeds that back).
    // (A third layout -- 256 classes per CTA, unsplit, two accumulators of 256 columns so that the epilogue of a
    // row tile runs under the MMAs of the next -- was built and measured in round 2: 8 % fewer cycles on cfg 3,
    // but every X tile is converted by four CTAs instead of two and W crosses L2 twice as often; under the 1 kW
    // cap the SM clock fell from ~1.9 to ~1.6 GHz and the step took 27.4-30.6 ms against 24.9-29.3; d=768, C=256
    // ran at 2.32 TB/s against 2.72 on the accurate layout.  Removed.)
    struct Layout {
        int nb = 0, ncc = 0, nrt = 1;
        int mma_n = BN;             // see GemmParams::mma_n
        bool pair = false;          // cta_group::2: cluster = 2 * ncc CTAs
        bool split = false;
        int max_clusters[64] = {0};
        CUtensorMap map_whi, map_wlo, map_w16hi, map_w16lo;    // W boxes are BN / nrt rows
    };
    static constexpr int kLayouts = 2;
    Layout layout[kLayouts];
    bool has[kLayouts] = {false, false};
    std::atomic<int> active{0};
    bool adaptive = true;
    int c_pad = 0;
    uint64_t rows_seen = 0, rescored_seen = 0;     // since the layout was last chosen (lr_score_host holds the staging lock)
    // Device-path feedback: after every launch the status block is copied (asynchronously) to this pinned mirror;
    // the next launch reads whatever has arrived -- rows scored [3] and rows re-scored [1] as of the same point of
    // the stream -- and applies the rule of gemm_feedback without a synchronisation.
    unsigned long long* mirror = nullptr;
    unsigned long long mirror_rows0 = 0, mirror_resc0 = 0;
    bool configured[64] = {false};
    bool fix_configured[64] = {false};
    bool fixup = true;          // re-score near-tie rows in fp32 (LRS_GEMM_FIXUP=0 turns it off)
};

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// 2-D fp32 row-major [rows][cols] with row pitch `ld` elements; box = box_rows x 32 columns, 128B swizzle.
// 2-D row-major [rows][cols] tensor with row pitch `ld` elements of `elem_bytes`; box = box_rows x box_cols.
int make_map_ex(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                uint32_t box_cols, int elem_bytes, bool swizzle128) {
    EncodeTiledFn fn = encode_tiled();
    if (!fn) return fail(LR_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * (uint64_t)elem_bytes};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                    const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(LR_ERR_CUDA, "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    return LR_OK;
}
// fp32, box_rows x 32 columns, 128-byte swizzle (TF32 mode operands)
int make_map(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
    return make_map_ex(map, base, rows, cols, ld, box_rows, (uint32_t)BK, 4, true);
}

}  // namespace

// Shapes the tensor-core path serves: more classes than the narrow path holds, up to 2048.  TMA wants a
// 16-byte aligned base and pitch: batches without are fetched by the kernel's own loader, see gemm_launch.
bool gemm_supports(const Model& m, const ScoreArgs& a) {
    return m.gemm != nullptr && a.n_rows < (int64_t)1 << 31;
}

// (Re)builds every operand copy of W from the float64 coefficients: called at creation and by
// lr_model_update (a refit hands in a new iterate per evaluation; nothing is allocated here).
int gemm_load(Model& m, const double* coef, const double* intercept) {
    GemmPlan* g = m.gemm;
    if (!g) return LR_OK;
    const int C = m.n_scores, d = m.n_features, d4 = g->d4;
    std::vector<float> bias(g->c_pad, -INFINITY);
    for (int c = 0; c < C; ++c) bias[c] = (float)intercept[c];
    LRS_CUDA(cudaMemcpy(g->bias, bias.data(), bias.size() * sizeof(float), cudaMemcpyHostToDevice));
    LRS_CUDA(cudaMemcpy(g->coef64, coef, (size_t)C * d * sizeof(double), cudaMemcpyHostToDevice));
    if (g->w32_pad) {
        // fp32 coefficients at the padded pitch, for the re-scoring pass (its W rows are read with 128-bit loads)
        std::vector<float> w32((size_t)C * d4, 0.0f);
        for (int c = 0; c < C; ++c)
            for (int k = 0; k < d; ++k) w32[(size_t)c * d4 + k] = (float)coef[(size_t)c * d + k];
        LRS_CUDA(cudaMemcpy(g->w32_pad, w32.data(), w32.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    // Scale + hi/lo splits on the device from the float64 coefficients: on the host the software fp16
    // conversions took 40 ms per million coefficients, which a refit pays per iteration.
    w_split_tf32_kernel<<<C, 256>>>(g->coef64, d, d4, g->w_hi, g->w_lo);
    w_colscale_kernel<<<(d + 255) / 256, 256>>>(g->coef64, C, d, g->col_scale, g->w_mul);
    w_split16_kernel<<<C, 256>>>(g->coef64, g->w_mul, d, g->d16, g->w16_hi, g->w16_lo, g->col_inv_scale, g->w_norm);
    count_launch(3);
    std::vector<float> inv(C), norm(C);
    LRS_CUDA(cudaMemcpy(inv.data(), g->col_inv_scale, C * sizeof(float), cudaMemcpyDeviceToHost));   // synchronises
    LRS_CUDA(cudaMemcpy(norm.data(), g->w_norm, C * sizeof(float), cudaMemcpyDeviceToHost));
    LRS_CUDA(cudaGetLastError());
    float nmax = 0.0f, imax = 0.0f;
    for (int c = 0; c < C; ++c) {
        if (norm[c] > nmax) nmax = norm[c];          // NaN never compares greater
        if (inv[c] > imax) imax = inv[c];
    }
    g->w_norm_max = nmax;
    g->inv_cscale_max = imax;
    return LR_OK;
}

int gemm_prepare(Model& m, const double* coef, const double* intercept) {
    if (m.n_scores < 13 || m.n_scores > kMaxCluster * 2 * BN) return LR_OK;
    if (!encode_tiled()) return LR_OK;            // no TMA descriptor API: generic path serves the model
    {
        // The per-call scratch of this path (X scale, flags, log-sum-exp) comes from the device's
        // default stream-ordered pool.  Left at its default threshold of 0 the pool hands its memory back to the
        // driver at every synchronisation, i.e. once per lr_score_host call: let it keep up to 1 GiB.
        cudaMemPool_t pool = nullptr;
        if (cudaDeviceGetDefaultMemPool(&pool, m.device) == cudaSuccess) {
            uint64_t cur = 0;
            if (cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &cur) == cudaSuccess && cur < (1ull << 30)) {
                uint64_t want = 1ull << 30;
                cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &want);
            }
        }
        cudaGetLastError();
    }
    GemmPlan* g = new GemmPlan();
    m.gemm = g;                                     // from here on lr_model_destroy frees what was allocated
    const int C = m.n_scores, d = m.n_features;
    const int d4 = (d + 3) / 4 * 4;                 // row pitch of the fp32 copies: TMA wants 16-byte pitches
    g->d4 = d4;
    g->d16 = (d + 7) / 8 * 8;
    g->d64 = (d + BK16 - 1) / BK16 * BK16;
    auto& fast = g->layout[0];
    auto& acc = g->layout[1];
    // Short rows against 129..256 classes: a tile's MMAs are over in no time and the epilogue is most of the work
    // (d=200, 160 classes: 3.4 ms per 5.4M rows, 2.1 without the epilogue), so it should at least overlap the next
    // tile's MMAs.  Rows this short leave a single accumulator as accurate as split ones are at d=1024, so the layout
    // drops the split and keeps TWO 256-column accumulator buffers instead, unpaired: d=200, 160 classes 3.49 ->
    // 3.07 ms, d=384, 256 classes 2.53 -> 2.35, d=128, 200 classes 4.63 -> 4.41.  (With more than one class slice --
    // tried up to 1024 classes -- the same trade loses: four CTAs meeting at every tile's exchange cost more than
    // the overlap returns, d=128, 1000 classes 3.59 -> 3.79 ms, d=64, 300 classes 4.28 -> 5.60.)
    const bool overlap256 = d4 <= 384 && C > 128 && C <= BN && !getenv("LRS_GEMM_NO_OVERLAP256");
    if (C > BN && !overlap256) {                    // fast layout: 512 classes per CTA
        fast.nb = 2;
        fast.ncc = (C + 2 * BN - 1) / (2 * BN);
        if (fast.ncc == 3) fast.ncc = 4;
        g->has[0] = true;
    }
    if (C <= kMaxCluster * BN) {                    // accurate layout: 256 classes per CTA, split accumulators
        acc.nb = 1;
        acc.ncc = (C + BN - 1) / BN;
        if (acc.ncc == 3) acc.ncc = 4;
        acc.split = !overlap256;
        g->has[1] = true;
    }
    // two row tiles per cluster share every W tile through TMA multicast (cluster <= 8 CTAs)
    bool mc = true;
    if (const char* e = getenv("LRS_GEMM_MC")) mc = atoi(e) != 0;
    bool pair = true;
    if (const char* e = getenv("LRS_GEMM_PAIR")) pair = atoi(e) != 0;
    // Measured on cfg 3 (tools/gpu_gemm_perf.sh): pairs help while the cluster stays at <= 4 CTAs
    // (fast layout, 26.8-28.0 ms against 28.4-28.7); clusters of 8 leave SMs idle (GPCs hold 16/18/20
    // SMs), so the accurate layout (4 class slices) runs unpaired and without multicast (37.3 ms against
    // 40.3 with multicast and 43.2 paired).  LRS_GEMM_PAIR / LRS_GEMM_MC = 1 force them where legal.
    const bool force_pair = getenv("LRS_GEMM_PAIR") && pair, force_mc = getenv("LRS_GEMM_MC") && mc;
    for (int i = 0; i < 2; ++i) {
        auto& L = g->layout[i];
        // (the accurate layout with two class slices: 3.39 ms unpaired against 4.40 paired on d=1024, 300 classes,
        // 3.93 / 4.19 on 500, 4.77 / 5.25 on d=512, 400 -- a cluster of 4 leaves 16 SMs idle and its X ring is shorter)
        L.pair = pair && !(overlap256 && i == 1) && ((i == 0 ? L.ncc * 2 <= 4 : L.ncc == 1) || (force_pair && L.ncc * 2 <= 8));
        L.nrt = (!L.pair && force_mc && L.ncc * 2 <= 8) ? 2 : 1;
    }
    // A model whose classes fit one W tile multiplies by exactly that many columns (in whole 32-column
    // epilogue chunks) instead of a zero-padded 256: the tensor time of a 32-class model drops 8x.  That
    // form is unpaired (a pair would split a W tile that is already small).
    if (acc.ncc == 1 && C <= BN - 32 && !force_pair) {
        acc.pair = false;
        acc.nrt = 1;
        acc.mma_n = (C + 31) / 32 * 32;
        if (const char* e = getenv("LRS_GEMM_MMA_N")) acc.mma_n = atoi(e) >= acc.mma_n ? BN : acc.mma_n;   // debug: any value >= N restores 256
    }
    // More than one class block: the classes are spread evenly over the blocks of all slices, in whole 32-column
    // chunks, instead of filling 256-wide blocks and padding the last ones with zeros (300 classes: 2 x 160 instead
    // of 2 x 256; 1100: 8 x 160 instead of 8 x 256).  cfg 3 (1000 classes: 4 x 256) is unchanged.
    if (!getenv("LRS_GEMM_MMA_N")) {
        for (int i = 0; i < GemmPlan::kLayouts; ++i) {
            auto& L = g->layout[i];
            if (!g->has[i] || L.ncc * L.nb == 1) continue;
            const int blocks = L.ncc * L.nb;
            int n = ((C + blocks - 1) / blocks + 31) / 32 * 32;
            if (L.nrt > 1) n = (n + 63) / 64 * 64;      // (opt-in W multicast: every CTA fetches mma_n / nrt rows)
            L.mma_n = n < BN ? n : BN;
        }
    }
    g->active = g->has[0] ? 0 : 1;                  // C <= 256: the accurate layout costs nothing extra
    if (const char* e = getenv("LRS_GEMM_SPLIT")) {  // pin a layout: 1 = accurate, 0 = fast
        const int want = atoi(e) != 0 ? 1 : 0;
        if (g->has[want]) g->active = want;
        g->adaptive = false;
    }
    if (const char* e = getenv("LRS_GEMM_LAYOUT")) { // pin a layout by index
        const int want = atoi(e);
        if (want >= 0 && want < GemmPlan::kLayouts && g->has[want]) g->active = want;
        g->adaptive = false;
    }
    g->c_pad = 0;
    for (int i = 0; i < GemmPlan::kLayouts; ++i)
        if (g->has[i] && g->layout[i].ncc * g->layout[i].nb * BN > g->c_pad) g->c_pad = g->layout[i].ncc * g->layout[i].nb * BN;
    if (const char* e = getenv("LRS_GEMM_FIXUP")) g->fixup = atoi(e) != 0;
    if (const char* e = getenv("LRS_GEMM_F16")) g->force = atoi(e) == 0 ? 2 : (atoi(e) == 2 ? 1 : 0);

    const size_t b32 = (size_t)g->c_pad * d4 * sizeof(float), b16 = (size_t)g->c_pad * g->d16 * sizeof(__half);
    LRS_CUDA(cudaMalloc(&g->coef64, (size_t)C * d * sizeof(double)));
    LRS_CUDA(cudaMalloc(&g->bias, g->c_pad * sizeof(float)));
    LRS_CUDA(cudaMalloc(&g->w_hi, b32));
    LRS_CUDA(cudaMalloc(&g->w_lo, b32));
    LRS_CUDA(cudaMalloc(&g->w16_hi, b16));
    LRS_CUDA(cudaMalloc(&g->w16_lo, b16));
    LRS_CUDA(cudaMalloc(&g->col_inv_scale, g->c_pad * sizeof(float)));
    LRS_CUDA(cudaMalloc(&g->w_norm, g->c_pad * sizeof(float)));
    LRS_CUDA(cudaMalloc(&g->col_scale, g->d64 * sizeof(float)));
    LRS_CUDA(cudaMalloc(&g->w_mul, d * sizeof(double)));
    LRS_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&g->mirror), 4 * sizeof(unsigned long long), cudaHostAllocDefault));
    std::memset(g->mirror, 0, 4 * sizeof(unsigned long long));
    if (d4 != d) LRS_CUDA(cudaMalloc(&g->w32_pad, (size_t)C * d4 * sizeof(float)));
    // padding (classes beyond C, columns beyond d) stays zero across reloads
    LRS_CUDA(cudaMemset(g->w_hi, 0, b32));
    LRS_CUDA(cudaMemset(g->w_lo, 0, b32));
    LRS_CUDA(cudaMemset(g->w16_hi, 0, b16));
    LRS_CUDA(cudaMemset(g->w16_lo, 0, b16));
    LRS_CUDA(cudaMemset(g->col_scale, 0, g->d64 * sizeof(float)));
    {
        std::vector<float> ones(g->c_pad, 1.0f);
        LRS_CUDA(cudaMemcpy(g->col_inv_scale, ones.data(), ones.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    int rc;
    for (int i = 0; i < GemmPlan::kLayouts; ++i) {
        if (!g->has[i]) continue;
        auto& L = g->layout[i];
        const int wbox = L.pair ? L.mma_n / 2 : L.mma_n / L.nrt;
        if ((rc = make_map(&L.map_whi, g->w_hi, g->c_pad, d, d4, wbox)) != LR_OK) return rc;
        if ((rc = make_map(&L.map_wlo, g->w_lo, g->c_pad, d, d4, wbox)) != LR_OK) return rc;
        if ((rc = make_map_ex(&L.map_w16hi, g->w16_hi, g->c_pad, d, g->d16, wbox, BK16, 2, true)) != LR_OK) return rc;
        if ((rc = make_map_ex(&L.map_w16lo, g->w16_lo, g->c_pad, d, g->d16, wbox, BK16, 2, true)) != LR_OK) return rc;
    }
    return gemm_load(m, coef, intercept);
}

// Called by lr_score_host after a synchronised batch: `rescored_total` is the running count kept in
// the model's status block.  If the fast layout sends more than a fifth of the rows to the fp32
// re-scoring pass, later batches use the accurate layout (which takes a third longer on the tensor
// cores; re-scoring a row costs ~1/4000 of that, so the layouts break even near 30 %).
void gemm_feedback(Model& m, int64_t rows, uint64_t rescored_total) {
    GemmPlan* g = m.gemm;
    if (!g || !g->adaptive || g->active == 1 || !g->has[1]) return;
    g->rows_seen += (uint64_t)rows;
    const uint64_t resc = rescored_total - g->rescored_seen;
    if (g->rows_seen >= 4096 && resc * 5 > g->rows_seen) {
        g->active = 1;
    } else if (g->rows_seen > (1ull << 24)) {          // judge on recent data only
        g->rows_seen = 0;
        g->rescored_seen = rescored_total;
    }
}

int gemm_active_layout(const Model& m) { return m.gemm ? m.gemm->active.load() : -1; }

void gemm_destroy(Model& m) {
    if (!m.gemm) return;
    cudaFree(m.gemm->w_hi);
    cudaFree(m.gemm->w_lo);
    cudaFree(m.gemm->bias);
    cudaFree(m.gemm->w16_hi);
    cudaFree(m.gemm->w16_lo);
    cudaFree(m.gemm->w32_pad);
    cudaFree(m.gemm->col_inv_scale);
    cudaFree(m.gemm->coef64);
    cudaFree(m.gemm->col_scale);
    cudaFree(m.gemm->w_mul);
    cudaFree(m.gemm->w_norm);
    if (m.gemm->mirror) cudaFreeHost(m.gemm->mirror);
    delete m.gemm;
    m.gemm = nullptr;
}

namespace {

// Stream-ordered scratch that is handed back on every way out of the launch code.
struct AsyncScratch {
    void* p = nullptr;
    cudaStream_t s = nullptr;
    ~AsyncScratch() { if (p) cudaFreeAsync(p, s); }
};

template <bool F16>
cudaError_t launch_gemm_kernel(const cudaLaunchConfig_t& cfg, bool pair, const CUtensorMap& map_x,
                               const GemmPlan::Layout& L, const GemmParams& p) {
    // 65..128 classes in one class slice: the instantiation whose epilogue keeps two chunks per warp in registers
    const bool epi2 = !pair && p.ncc == 1 && p.nb == 1 && p.mma_n > 64 && p.mma_n <= 128;
    if constexpr (F16) {
        if (epi2) return cudaLaunchKernelEx(&cfg, gemm_softmax_kernel<true, false, true>, map_x, L.map_w16hi, L.map_w16lo, p);
        return pair ? cudaLaunchKernelEx(&cfg, gemm_softmax_kernel<true, true>, map_x, L.map_w16hi, L.map_w16lo, p)
                    : cudaLaunchKernelEx(&cfg, gemm_softmax_kernel<true, false>, map_x, L.map_w16hi, L.map_w16lo, p);
    } else {
        if (epi2) return cudaLaunchKernelEx(&cfg, gemm_softmax_kernel<false, false, true>, map_x, L.map_whi, L.map_wlo, p);
        return pair ? cudaLaunchKernelEx(&cfg, gemm_softmax_kernel<false, true>, map_x, L.map_whi, L.map_wlo, p)
                    : cudaLaunchKernelEx(&cfg, gemm_softmax_kernel<false, false>, map_x, L.map_whi, L.map_wlo, p);
    }
}

// `d_eff`: row length as the W copies see it (n_features rounded up to 4); `w_fix`: fp32 W at pitch d_eff for the
// re-scoring pass; `sw_load`: TMA cannot address X (GemmParams::sw_load).
int gemm_launch_aligned(const Model& m, const ScoreArgs& a, int d_eff, const float* w_fix, bool sw_load) {
    GemmPlan* g = m.gemm;
    // Labels + max probability without a P matrix (REF/main.py:21-22's pair of calls, in bulk): the rows whose top
    // probability the error bound could move are re-scored, and with no row of P to pick the classes that matter
    // from, that pass recomputes every class of such a row -- 4 MB of W per row of a 1000-class model, 21.9 ms per
    // 1M rows of cfg 3's shape where the whole P matrix takes 6.4.  So the batch is scored in blocks of rows into a
    // scratch P of at most 512 MB (stream-ordered; blocks of 64k rows left a quarter of the last wave idle: 9.8 ms)
    // and re-scored class by class like any other.  No memory for the scratch: the old way.
    // (A class group of a larger model comes with a pitch for the caller's rows: the scratch has its own, so a group
    // that also has decision values to write at the caller's pitch stays on the old way.)
    if (!a.proba && a.maxprob && g->fixup && m.n_scores > 32 && (!a.out_pitch || !a.decision) && !getenv("LRS_GEMM_NO_SCRATCH_P")) {
        const int C = m.n_scores;
        const char* mb = getenv("LRS_GEMM_SCRATCH_P_MB");           // (tests shrink it to see several blocks)
        int64_t block = ((int64_t)(mb && atoi(mb) > 0 ? atoi(mb) : 512) << 20) / ((int64_t)C * 4) / 1024 * 1024;
        block = block < 4096 ? 4096 : block;
        if (block > a.n_rows) block = a.n_rows;
        AsyncScratch sp;
        sp.s = a.stream;
        if (cudaMallocAsync(&sp.p, (size_t)block * C * sizeof(float), a.stream) != cudaSuccess) {
            cudaGetLastError();
            sp.p = nullptr;
        }
        int rc = LR_OK;
        for (int64_t r0 = 0; sp.p && r0 < a.n_rows && rc == LR_OK; r0 += block) {
            ScoreArgs s = a;
            s.X = a.X + r0 * a.ldx;
            s.n_rows = a.n_rows - r0 < block ? a.n_rows - r0 : block;
            s.proba = static_cast<float*>(sp.p);
            s.out_pitch = 0;
            s.out_col0 = 0;
            if (a.lse_out) s.lse_out = a.lse_out + r0;
            if (a.decision) s.decision = a.decision + r0 * C;
            if (a.label) s.label = a.label + r0;
            s.maxprob = a.maxprob + r0;
            rc = gemm_launch_aligned(m, s, d_eff, w_fix, sw_load);
        }
        if (sp.p) return rc;
    }
    const int d_x = m.n_features;
    const bool xal = !sw_load && d_x % 4 == 0;      // 128-bit loads of X rows in the side kernels
    if (g->adaptive && g->active.load() == 0 && g->has[1]) {
        // what the mirror shows of earlier launches on this model (it may lag by a launch or two)
        const unsigned long long rows = *reinterpret_cast<volatile unsigned long long*>(g->mirror + 3);
        const unsigned long long resc = *reinterpret_cast<volatile unsigned long long*>(g->mirror + 1);
        if (rows - g->mirror_rows0 >= 4096 && (resc - g->mirror_resc0) * 5 > rows - g->mirror_rows0) {
            g->active = 1;
        } else if (rows - g->mirror_rows0 > (1ull << 24)) {          // judge on recent data only
            g->mirror_rows0 = rows;
            g->mirror_resc0 = resc;
        }
    }
    GemmPlan::Layout& L = g->layout[g->active.load()];
    // X as the two arithmetics read it: fp32 boxes of 64 columns, unswizzled (3xFP16: converted in place) or of
    // 32 columns, 128-byte swizzle (3xTF32)
    // (columns past the real row length and rows past the batch read as zeros; with sw_load the kernel takes no X map,
    // a W map stands in)
    CUtensorMap map_x16 = L.map_w16hi, map_x32 = L.map_whi;
    int rc;
    if (!sw_load && g->force != 2 &&
        (rc = make_map_ex(&map_x16, a.X, (uint64_t)a.n_rows, (uint64_t)d_x, (uint64_t)a.ldx, BM, BK16, 4, false)) != LR_OK)
        return rc;
    if (!sw_load && g->force != 1 && (rc = make_map(&map_x32, a.X, (uint64_t)a.n_rows, (uint64_t)d_x, (uint64_t)a.ldx, BM)) != LR_OK)
        return rc;
    const size_t smem = (L.pair ? SmemLayoutT<true>::total : SmemLayoutT<false>::total) + 1024;
    const int dev = m.device & 63;
    if (!g->configured[dev]) {
        const int s0 = SmemLayoutT<false>::total + 1024, s1 = SmemLayoutT<true>::total + 1024;
        LRS_CUDA(cudaFuncSetAttribute(gemm_softmax_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, s0));
        LRS_CUDA(cudaFuncSetAttribute(gemm_softmax_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, s0));
        LRS_CUDA(cudaFuncSetAttribute(gemm_softmax_kernel<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, s0));
        LRS_CUDA(cudaFuncSetAttribute(gemm_softmax_kernel<true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, s0));
        LRS_CUDA(cudaFuncSetAttribute(gemm_softmax_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, s1));
        LRS_CUDA(cudaFuncSetAttribute(gemm_softmax_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, s1));
        g->configured[dev] = true;
    }
    // Per-call scratch, stream-ordered (cudaMallocAsync) so that concurrent lr_score calls on one model stay
    // independent: [256 B: what x_scale_kernel decides | its per-row statistics | per row the log-sum-exp the
    // tensor path saw | one flag byte per row (near-ties, risky probabilities) for the fp32 re-scoring pass].
    const bool fixup = g->fixup && (a.label || a.maxprob || a.proba || a.decision);
    unsigned* flags = nullptr;
    unsigned* flag_list = nullptr;
    float* lse = nullptr;
    float* ezr = nullptr;
    AsyncScratch scratch;
    scratch.s = a.stream;
    const size_t stat_bytes = 2 * kMaxSample * sizeof(float);
    const size_t lse_bytes = fixup ? ((size_t)a.n_rows * sizeof(float) + 255) / 256 * 256 : 0;
    LRS_CUDA(cudaMallocAsync(&scratch.p, 256 + stat_bytes + 4 * lse_bytes, a.stream));
    float* x_scale = reinterpret_cast<float*>(scratch.p);
    float* rowstat = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch.p) + 256);
    if (fixup) {
        lse = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch.p) + 256 + stat_bytes);
        flags = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(scratch.p) + 256 + stat_bytes + lse_bytes);
        flag_list = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(scratch.p) + 256 + stat_bytes + 2 * lse_bytes);
        ezr = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch.p) + 256 + stat_bytes + 3 * lse_bytes);
        LRS_CUDA(cudaMemsetAsync(flags, 0, (size_t)a.n_rows * sizeof(unsigned), a.stream));
    }
    if (a.lse_out) lse = a.lse_out;                 // a class group hands its log-sum-exp to the caller
    LRS_CUDA(cudaMemsetAsync(x_scale, 0, 64, a.stream));          // ticket, decisions, [8] = length of the re-scoring list
    if (xal)
        x_scale_kernel<true><<<kXScaleBlocks, 256, 0, a.stream>>>(a.X, a.n_rows, a.ldx, d_x, g->col_scale, g->w_norm_max,
                                                                  g->inv_cscale_max, g->force, x_scale, rowstat);
    else
        x_scale_kernel<false><<<kXScaleBlocks, 256, 0, a.stream>>>(a.X, a.n_rows, a.ldx, d_x, g->col_scale, g->w_norm_max,
                                                                   g->inv_cscale_max, g->force, x_scale, rowstat);
    count_launch();
    GemmParams p;
    // (8-byte copies where every row starts on an 8-byte boundary)
    p.sw_load = !sw_load ? 0 : ((a.ldx % 2 == 0 && (reinterpret_cast<uintptr_t>(a.X) & 7u) == 0) ? 2 : 1);
    p.x_raw = a.X;
    p.ldx = a.ldx;
    p.d_x = d_x;
    p.flags = flags;
    p.flag_list = flag_list;
    p.n_flagged = reinterpret_cast<unsigned*>(x_scale) + 8;
    p.ezr = ezr;
    p.lse = lse;
    p.n_rows = a.n_rows;
    p.n_features = d_eff;
    p.n_classes = m.n_scores;
    p.nb = L.nb;
    p.split = L.split ? 1 : 0;
    p.ncc = L.ncc;
    p.nrt = L.nrt;
    p.mma_n = L.mma_n;
    // Shared-memory split of the unpaired kernel (192 KB): W tiles are mma_n rows x 128 B.  With the
    // fixed 4 x 32 KB W ring only two k-chunks of W (hi + lo tile each) were in flight, and a W slot is
    // held from its L2 fetch until the MMAs that read it retire: ~1.15 us per chunk however deep the X
    // ring was.  Small tiles buy a 12- or 6-deep W ring and a third X stage.
    // (Converting the fp32 tile in a ring of its own instead of in place was tried: 4137 against 4690 GB/s at
    // d=256, C=32, 3214 against 3457 at d=4096, C=64 -- the rings it leaves room for are shallower.)
    if (L.pair) {
        p.b_tile_bytes = SmemLayoutT<true>::b_tile; p.a_stages = SmemLayoutT<true>::a_stages; p.b_stages = SmemLayoutT<true>::b_stages;
    } else {
        // W stages of exactly mma_n rows (multiples of 4 KB, so every tile keeps the 1 KB alignment the swizzle
        // wants); a third X stage whenever four W stages still fit beside it
        p.b_tile_bytes = L.mma_n * 128;
        p.a_stages = 3 * 32 * 1024 + 4 * p.b_tile_bytes <= 192 * 1024 ? 3 : 2;
        p.b_stages = (192 * 1024 - p.a_stages * 32 * 1024) / p.b_tile_bytes;
        if (p.b_stages > kMaxBStages) p.b_stages = kMaxBStages;
    }
    if (const char* e = getenv("LRS_GEMM_RINGS")) {          // tuning: "a_stages,b_stages"
        int as = 0, bs = 0;
        if (sscanf(e, "%d,%d", &as, &bs) == 2 && as >= 1 && as <= kMaxAStages && bs >= 2 && bs <= kMaxBStages &&
            as * 32 * 1024 + bs * p.b_tile_bytes <= 192 * 1024) {
            p.a_stages = as; p.b_stages = bs;
        }
    }
    // two accumulator buffers where TMEM holds them: split layout (main + correction) up to N = 128, 128 columns
    // apart in each half; single accumulators up to N = 256, 256 apart
    p.acc_bufs = (L.nb == 1 && !L.pair && L.mma_n <= (L.split ? 128 : 256)) ? 2 : 1;
    p.acc_stride = (!L.split && L.mma_n > 128) ? 256 : 128;
    // Long rows of a model with at most 128 classes: several main accumulators, each over a contiguous piece of K of
    // at most 800 columns (see GemmParams::ksplit).  Beside two accumulator buffers there is room for 3 (7 when N <= 32).
    // Rows that want more give up the second buffer -- a tile of such rows takes HBM several times longer to
    // deliver than the MMAs and the (now serial) epilogue need together -- for up to 7 (15): d=4096, 64 classes
    // re-scored 13.5 % of its rows with 3 pieces (1.27 ms per 262k rows) and 0.006 % with 7 (1.08 ms).
    p.ksplit = 1;
    p.seg_w = L.mma_n <= 32 ? 32 : (L.mma_n <= 64 ? 64 : 128);
    p.seg_slot = 128;
    if (L.split && L.mma_n <= 128 && !L.pair && L.nb == 1) {
        const int most2 = 2 * (128 / p.seg_w) - 1, most1 = 2 * (256 / p.seg_w) - 1;
        // (a row is flagged for its top probability once the error bound passes 3.2e-5: pieces of up to 800 columns
        // keep it under that for logits up to ~5; shorter pieces only add TMEM reads -- d=784, 16 classes lost 8 % to
        // a second piece it had no use for)
        const int want = (d_eff + 799) / 800;
        // 65..128 classes: one main accumulator beside a second buffer, three without.  Their epilogue is long
        // enough to miss the overlap on rows of ~1000 floats (d=1024, 128 classes: 1.34 ms with 3.1 % re-scored
        // against 1.49 with none), so they switch later: d=1536, 96 classes 1.82 -> 1.34 ms (33 % -> 2.3 % re-scored),
        // d=2048, 100 classes 1.75 -> 1.32, d=4096, 128 classes 4.5 -> 1.4 (70 % -> 7 %).
        if (want > most2 && (p.seg_w < 128 || d_eff > 1280) && !getenv("LRS_GEMM_TWO_BUFFERS")) {   // (64 classes: d > 2400; 32: d > 5600)
            p.acc_bufs = 1;
            p.seg_slot = 256;
            p.ksplit = want > most1 ? most1 : want;
        } else {
            p.ksplit = want > most2 ? most2 : want;
        }
    }
    if (const char* e = getenv("LRS_GEMM_KSPLIT")) p.ksplit = (atoi(e) >= 1 && atoi(e) <= p.ksplit) ? atoi(e) : p.ksplit;
    if (const char* e = getenv("LRS_GEMM_ACC_BUFS")) p.acc_bufs = (atoi(e) == 2 && p.acc_bufs == 2) ? 2 : 1;   // debug: 1 turns the overlap off
    p.bias = g->bias;
    p.col_inv_scale = g->col_inv_scale;
    p.col_scale = g->col_scale;
    p.x_scale = x_scale;
    p.count_nonfinite = fixup ? 0 : 1;
    p.need_prob = (a.proba || a.maxprob) ? 1 : 0;
    p.debug_skip = 0;
    if (const char* e = getenv("LRS_GEMM_DEBUG_SKIP")) p.debug_skip = atoi(e);
    p.proba = a.proba;
    p.decision = a.decision;
    p.label = a.label;
    p.maxprob = a.maxprob;
    p.status = m.status;
    p.out_pitch = a.out_pitch ? a.out_pitch : m.n_scores;
    p.out_col0 = a.out_col0;
    const bool al = (p.out_pitch % 4 == 0) && (p.out_col0 % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.proba) & 15u) == 0) &&
                    ((reinterpret_cast<uintptr_t>(a.decision) & 15u) == 0);
    const bool al32 = al && (p.out_pitch % 8 == 0) && (p.out_col0 % 8 == 0) && ((reinterpret_cast<uintptr_t>(a.proba) & 31u) == 0);
    p.vec_store = al32 ? 2 : (al ? 1 : 0);
    // Rows that cannot take 32-byte vector stores leave through 16 KB of store tiles at the end of the ring area, paid
    // for with a W stage where the rings fill it (the register epilogue of <= 64 classes has its tiles elsewhere).
    // Measured (d=512, 2M rows): 101 classes through the tiles 1.31 ms, 100 classes with 16-byte stores 1.40 ms.
    p.stage_off = 0;
    if (p.vec_store < 2 && (a.proba || a.decision) && (L.pair || L.ncc > 1 || L.nb * L.mma_n > 64)) {
        const int ring_end = 192 * 1024 - 8 * kStoreTileBytes;
        int bs = p.b_stages;
        while (bs > 2 && p.a_stages * 32 * 1024 + bs * p.b_tile_bytes > ring_end) --bs;
        if (p.a_stages * 32 * 1024 + bs * p.b_tile_bytes <= ring_end) {
            p.b_stages = bs;
            p.stage_off = ring_end;
        }
    }

    const int64_t n_tiles = (a.n_rows + BM - 1) / BM;
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = a.stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    const int gsz = L.pair ? 2 : L.nrt;             // row tiles per cluster iteration
    const int csize = L.ncc * gsz;
    attr[0].val.clusterDim.x = (unsigned)csize;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // Persistent grid = as many clusters as can be co-resident.  A cluster must sit inside one GPC and
    // GPCs hold 16/18/20 SMs, so clusters of 4 do not tile all 148 SMs: asking for floor(148/ncc)
    // clusters would leave the last one waiting for a whole wave (measured: 2x slower).
    if (L.max_clusters[dev] == 0) {
        cfg.gridDim = dim3((unsigned)(m.num_sms / csize * csize));
        int nc = 0;
        cudaError_t qe = L.pair ? cudaOccupancyMaxActiveClusters(&nc, gemm_softmax_kernel<true, true>, &cfg)
                                : cudaOccupancyMaxActiveClusters(&nc, gemm_softmax_kernel<true, false>, &cfg);
        if (qe != cudaSuccess || nc < 1) {
            cudaGetLastError();
            nc = m.num_sms / csize / 2 > 0 ? m.num_sms / csize / 2 : 1;
        }
        L.max_clusters[dev] = nc;
    }
    int64_t clusters = L.max_clusters[dev];
    const int64_t n_groups = (n_tiles + gsz - 1) / gsz;
    if (clusters > n_groups) clusters = n_groups;
    cfg.gridDim = dim3((unsigned)(clusters * csize));
    // Truncating accumulation: the largest |dz| / max(1, |z|) over 2M logits at d = 1024 (tools/gemm_kappa.py:
    // 1.68e-5 / 6.5e-6 for 3xFP16 on one / split accumulators, 3.1e-5 / 1.08e-5 for 3xTF32), plus a fifth.  It
    // grows with the number of accumulation steps, i.e. linearly with the row length (measured at d = 256, 1024,
    // 4096, 16384: 6.3e-6, 1.68e-5, 6.7e-5, 2.4e-4), a little slower than that below 1024.
    const int d_acc = d_eff / p.ksplit;              // row length behind one main accumulator
    const float len = (float)(d_acc > 384 ? d_acc : 384) / 1024.0f;
    // Both arithmetics are enqueued; the kernel x_scale_kernel did not choose returns at once (a few us).
    if (g->force != 2) {
        p.num_k_chunks = (d_eff + BK16 - 1) / BK16;
        p.kappa = (L.split ? 8e-6f : 2.0e-5f) * len;
        p.count_fallback = 0;
        LRS_CUDA(launch_gemm_kernel<true>(cfg, L.pair, map_x16, L, p));
        count_launch();
    }
    if (g->force != 1) {
        p.num_k_chunks = (d_eff + BK - 1) / BK;
        p.kappa = (L.split ? 1.3e-5f : 3.7e-5f) * len;
        p.count_fallback = g->force == 0 ? 1 : 0;
        LRS_CUDA(launch_gemm_kernel<false>(cfg, L.pair, map_x32, L, p));
        count_launch();
    }
    if (fixup) {
        if (a.proba) {
            const int c_pad = (m.n_scores + 31) / 32 * 32;
            const size_t fsm = (size_t)kFixWarps * c_pad * sizeof(float) + (size_t)kFixWarps * kFixQueue * sizeof(int);
            if (fsm > 48 * 1024 && !g->fix_configured[dev]) {
                LRS_CUDA(cudaFuncSetAttribute(gemm_fixup_sel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024));
                g->fix_configured[dev] = true;
            }
            gemm_fixup_sel_kernel<<<m.num_sms * 3, kFixWarps * 32, fsm, a.stream>>>(
                a.X, a.ldx, d_x, d_eff, xal ? 1 : 0, m.n_scores, c_pad, p.out_pitch, p.out_col0, w_fix, m.b, flag_list, p.n_flagged, lse, ezr,
                a.proba, a.decision,
                a.label, a.maxprob, m.status + 1, m.status);
        } else {
            gemm_fixup_full_kernel<<<m.num_sms * 4, 256, 0, a.stream>>>(a.X, a.ldx, d_x, d_eff, xal ? 1 : 0, m.n_scores, p.out_pitch,
                                                                       p.out_col0, w_fix, m.b,
                                                                       flag_list, p.n_flagged, a.decision, a.label,
                                                                       a.maxprob, m.status + 1, m.status);
        }
        count_launch();
        LRS_CUDA(cudaGetLastError());
    }
    if (g->adaptive && g->active.load() == 0 && g->has[1])
        LRS_CUDA(cudaMemcpyAsync(g->mirror, m.status, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, a.stream));
    return LR_OK;
}

// 13..32 classes on dense rows of at most 64 floats: short_rows_kernel (see there).
bool short_rows_serve(const Model& m, const ScoreArgs& a) {
    const GemmPlan* g = m.gemm;
    return g && g->force != 2 && m.n_scores >= 13 && m.n_scores <= 32 && m.n_features <= 64 && a.ldx == m.n_features &&
           (reinterpret_cast<uintptr_t>(a.X) & 15u) == 0 && a.n_rows >= 4096 && g->has[1] && g->layout[1].mma_n == 32 &&
           !g->layout[1].pair && !getenv("LRS_NO_SHORT_ROWS");
}

int short_launch(const Model& m, const ScoreArgs& a) {
    GemmPlan* g = m.gemm;
    GemmPlan::Layout& L = g->layout[1];
    const int d = m.n_features, C = m.n_scores;
    const bool fixup = g->fixup && (a.label || a.maxprob || a.proba || a.decision);
    AsyncScratch scratch;
    scratch.s = a.stream;
    const size_t stat_bytes = 2 * kMaxSample * sizeof(float);
    const size_t row_bytes = fixup ? ((size_t)a.n_rows * sizeof(unsigned) + 255) / 256 * 256 : 0;
    LRS_CUDA(cudaMallocAsync(&scratch.p, 256 + stat_bytes + 2 * row_bytes, a.stream));
    float* x_scale = reinterpret_cast<float*>(scratch.p);
    float* rowstat = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch.p) + 256);
    unsigned* flags = nullptr;
    unsigned* flag_list = nullptr;
    if (fixup) {
        flags = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(scratch.p) + 256 + stat_bytes);
        flag_list = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(scratch.p) + 256 + stat_bytes + row_bytes);
        LRS_CUDA(cudaMemsetAsync(flags, 0, (size_t)a.n_rows * sizeof(unsigned), a.stream));
    }
    LRS_CUDA(cudaMemsetAsync(x_scale, 0, 64, a.stream));
    if (d % 4 == 0)
        x_scale_kernel<true><<<kXScaleBlocks, 256, 0, a.stream>>>(a.X, a.n_rows, a.ldx, d, g->col_scale, g->w_norm_max,
                                                                  g->inv_cscale_max, g->force, x_scale, rowstat);
    else
        x_scale_kernel<false><<<kXScaleBlocks, 256, 0, a.stream>>>(a.X, a.n_rows, a.ldx, d, g->col_scale, g->w_norm_max,
                                                                   g->inv_cscale_max, g->force, x_scale, rowstat);
    count_launch();
    ShortParams p;
    p.X = a.X;
    p.n_rows = a.n_rows;
    p.d = d;
    p.n_classes = C;
    p.stage_bytes = (BM * d * 4 + 127) / 128 * 128;
    // rows of at most 32 floats leave room for two operand tiles per converter set: a set then converts its next
    // tile while the MMAs of its previous one are still on their way (measured: the hand-over, not the work, set the pace)
    p.op_stages = 2;             // four (rows of <= 32 floats leave the room) measured the same
    // measured (tools/shape_sweep.py): d=16, C=26 2.8 -> 3.9 TB/s with 1 + 3 sets, d=54, C=20 3.0 -> 3.9 with 2 + 2
    p.conv_sets = d < 32 ? 1 : 2;                    // (d = 32, 20 classes: 3.70 TB/s with one set, 3.80 with two; d = 16, 26 classes: 3.74 against 3.04)
    if (const char* e = getenv("LRS_SHORT_CONV_SETS")) p.conv_sets = atoi(e) == 1 ? 1 : 2;
    const int room = 227 * 1024 - 1024 - SrSmem::op - p.op_stages * kSrOpBytes;
    p.raw_stages = room / p.stage_bytes < kSrMaxRaw ? room / p.stage_bytes : kSrMaxRaw;
    p.bias = g->bias;
    p.col_inv_scale = g->col_inv_scale;
    p.x_scale = x_scale;
    p.col_scale = g->col_scale;
    // 3 x (dk / 16) accumulation steps only: the truncation error measured on these shapes (tools/gemm_kappa.py
    // --short) is 0.7-1.9e-6 |z|; the longer-row kernel's floor (7.5e-6) flagged 2-5 % of the rows of a
    // 26..32-class model for nothing
    p.kappa = float(getenv("LRS_SHORT_KAPPA") ? atof(getenv("LRS_SHORT_KAPPA")) : 1.0e-6 + 2.5e-6 * ((d + 15) / 16) / 4.0);
    p.count_nonfinite = fixup ? 0 : 1;
    p.need_prob = (a.proba || a.maxprob) ? 1 : 0;
    p.vec_store = 0;
    p.proba = a.proba;
    p.decision = a.decision;
    p.label = a.label;
    p.maxprob = a.maxprob;
    p.status = m.status;
    p.flags = flags;
    p.flag_list = flag_list;
    p.n_flagged = reinterpret_cast<unsigned*>(x_scale) + 8;
    p.ezr = nullptr;
    const size_t smem = (size_t)SrSmem::op + (size_t)p.op_stages * kSrOpBytes + (size_t)p.raw_stages * p.stage_bytes + 1024;
    const int64_t n_tiles = (a.n_rows + BM - 1) / BM;
    const int grid = (int)(n_tiles < m.num_sms ? n_tiles : m.num_sms);
    const bool vec = d % 4 == 0;
    auto launch = [&](auto kern) -> int {
        LRS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        kern<<<grid, kSrThreads, smem, a.stream>>>(L.map_w16hi, L.map_w16lo, p);
        return LR_OK;
    };
    int rc;
    switch ((d + 15) / 16) {
        case 1: rc = vec ? launch(short_rows_kernel<16, true>) : launch(short_rows_kernel<16, false>); break;
        case 2: rc = vec ? launch(short_rows_kernel<32, true>) : launch(short_rows_kernel<32, false>); break;
        case 3: rc = vec ? launch(short_rows_kernel<48, true>) : launch(short_rows_kernel<48, false>); break;
        default: rc = vec ? launch(short_rows_kernel<64, true>) : launch(short_rows_kernel<64, false>); break;
    }
    if (rc != LR_OK) return rc;
    count_launch();
    if (g->force == 0) {
        short_fallback_kernel<<<m.num_sms * 8, 256, 0, a.stream>>>(a.X, a.n_rows, d, C, m.W, m.b, x_scale, a.proba,
                                                                   a.decision, a.label, a.maxprob, m.status);
        count_launch();
    }
    if (fixup) {
        short_fixup_kernel<<<m.num_sms * 4, 256, 0, a.stream>>>(a.X, d, C, m.W, m.b, flag_list, p.n_flagged, a.proba,
                                                                a.decision, a.label, a.maxprob, m.status + 1, m.status);
        count_launch();
    }
    LRS_CUDA(cudaGetLastError());
    return LR_OK;
}

}  // namespace

bool gemm_short_rows(const Model& m, const ScoreArgs& a) { return short_rows_serve(m, a); }

int gemm_launch(const Model& m, const ScoreArgs& a) {
    if (!gemm_supports(m, a)) return fail(LR_ERR_INVALID, "gemm path needs 13..2048 classes and fewer than 2^31 rows");
    GemmPlan* g = m.gemm;
    const int d = m.n_features;
    if (short_rows_serve(m, a)) return short_launch(m, a);
    // TMA wants a 16-byte aligned base and row pitch (any row length: columns past it are out of bounds and read as
    // zeros, which is what the zero-padded W copies multiply them with).  Everything else -- odd pitches, the dense
    // rows of an odd feature count, a base off by a few floats -- is fetched by the kernel's X producer warps with
    // 4-byte asynchronous copies (GemmParams::sw_load).  (Until late in round 2 such batches were first copied to an
    // aligned scratch: one extra write and read of X, 9.9 ms against 6.4 on d=1022 / 1024 with 1000 classes.)
    const bool tma_ok = a.ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(a.X) & 15u) == 0;
    const int d4 = g->d4;
    return gemm_launch_aligned(m, a, d4, d4 == d ? m.W : g->w32_pad, !tma_ok);
}

// ---- more than 2048 classes: class groups --------------------------------------------------------------------
// A model of C > 2048 classes is cut into G = ceil(C / 2048) groups of equal size (a multiple of 8 classes; the last
// takes what is left), each an ordinary tensor-path model over its slice of W / b.  A call scores the batch once per
// group -- each writes the softmax over its OWN classes into its columns of the caller's rows, and per row the
// log-sum-exp, best class and best probability of the group -- then one kernel per row combines the groups: row
// log-sum-exp L = logsumexp_g(lse_g), every group's columns scaled by exp(lse_g - L), the best class taken across
// groups (lowest group on a tie; two groups whose best logits lie within the accumulation error band of each other
// have those two logits recomputed in fp32).  X is read G times (it stays in L2 for batches of a few hundred MB);
// P is written once and rescaled once.  Replaces the SIMT fallback these models ran on (one warp per row: ~0.03 TB/s).
namespace {

constexpr int kGroupClasses = 2048;

__global__ void __launch_bounds__(256)
groups_combine_kernel(const float* __restrict__ X, int64_t n_rows, int64_t ldx, int d, int C, int G, int S,
                      const float* __restrict__ W, const float* __restrict__ b, float band,
                      float* __restrict__ lse_g /* [G][n]: in: group log-sum-exp, out: scale of the group's columns */,
                      const int32_t* __restrict__ label_g, const float* __restrict__ maxp_g,
                      float* proba, int32_t* label, float* maxprob, unsigned long long* status) {
    const int lane = threadIdx.x & 31;
    for (int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); row < n_rows; row += (int64_t)gridDim.x * 8) {
        // lanes hold groups (G <= 32)
        const float l = lane < G ? lse_g[(int64_t)lane * n_rows + row] : -CUDART_INF_F;
        const float mp = lane < G ? maxp_g[(int64_t)lane * n_rows + row] : 1.0f;
        const int lb = lane < G ? label_g[(int64_t)lane * n_rows + row] : 0;
        float zbest = lane < G ? l + logf(mp) : -CUDART_INF_F;         // the group's largest logit
        float m = l;
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
        float e = lane < G ? expf(l - m) : 0.0f;
        float ssum = e;
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) ssum += __shfl_xor_sync(0xffffffffu, ssum, off);
        const float L = m + logf(ssum);
        const float scale = e / ssum;                                  // exp(lse_g - L)
        if (lane < G) lse_g[(int64_t)lane * n_rows + row] = scale;
        // best group (lowest on a tie), runner-up
        float z1 = zbest;
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) z1 = fmaxf(z1, __shfl_xor_sync(0xffffffffu, z1, off));
        unsigned win = __ballot_sync(0xffffffffu, zbest == z1);
        int g1 = __ffs(win) - 1;
        float z2 = lane == g1 ? -CUDART_INF_F : zbest;
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) z2 = fmaxf(z2, __shfl_xor_sync(0xffffffffu, z2, off));
        int best = __shfl_sync(0xffffffffu, lb, g1) + g1 * S;
        float pbest = __shfl_sync(0xffffffffu, mp * scale, g1);
        if (G > 1 && z1 - z2 <= band * fmaxf(1.0f, fabsf(z1))) {
            // two groups' champions within the error band: their logits again, in fp32
            const unsigned win2 = __ballot_sync(0xffffffffu, lane != g1 && zbest == z2);
            const int g2 = __ffs(win2) - 1;
            const int c1 = best, c2 = __shfl_sync(0xffffffffu, lb, g2) + g2 * S;
            const float* x = X + row * ldx;
            float a1 = 0.0f, a2 = 0.0f;
            for (int k = lane; k < d; k += 32) {
                const float xv = __ldg(x + k);
                a1 = fmaf(xv, __ldg(W + (int64_t)c1 * d + k), a1);
                a2 = fmaf(xv, __ldg(W + (int64_t)c2 * d + k), a2);
            }
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
                a1 += __shfl_xor_sync(0xffffffffu, a1, off);
                a2 += __shfl_xor_sync(0xffffffffu, a2, off);
            }
            a1 += __ldg(b + c1);
            a2 += __ldg(b + c2);
            if (a2 > a1 || (a2 == a1 && c2 < c1)) { best = c2; pbest = expf(a2 - L); }
            else pbest = expf(a1 - L);
        }
        if (lane == 0) {
            if (label) label[row] = best;
            if (maxprob) maxprob[row] = pbest;
            if (!finite_f(L)) atomicAdd(status, 1ull);
        }
    }
}

// every group's columns of a row times that group's scale (groups_combine_kernel left it in lse_g)
__global__ void __launch_bounds__(256)
groups_scale_kernel(float* __restrict__ proba, int64_t n_rows, int C, int S, const float* __restrict__ scale_g) {
    const int64_t total4 = n_rows * (int64_t)(C / 4);
    if (C % 4 == 0 && (reinterpret_cast<uintptr_t>(proba) & 15u) == 0) {
        for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total4; t += (int64_t)gridDim.x * 256) {
            const int64_t row = t / (C / 4);
            const int c = (int)(t - row * (C / 4)) * 4;               // S % 8 == 0: four columns never straddle groups
            const float sc = __ldg(scale_g + (int64_t)(c / S) * n_rows + row);
            float4 v = reinterpret_cast<float4*>(proba)[t];
            v.x *= sc; v.y *= sc; v.z *= sc; v.w *= sc;
            reinterpret_cast<float4*>(proba)[t] = v;
        }
    } else {
        const int64_t total = n_rows * (int64_t)C;
        for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
            const int64_t row = t / C;
            const int c = (int)(t - row * C);
            proba[t] *= __ldg(scale_g + (int64_t)(c / S) * n_rows + row);
        }
    }
}

}  // namespace

int gemm_groups_prepare(Model& m, const double* coef, const double* intercept) {
    if (m.n_scores <= kGroupClasses || !encode_tiled()) return LR_OK;
    const int C = m.n_scores, d = m.n_features;
    const int G = (C + kGroupClasses - 1) / kGroupClasses;
    if (G > 32) return LR_OK;                       // 65 536 classes and more: the generic path
    const int S = ((C + G - 1) / G + 7) / 8 * 8;    // group size: columns of a group start 32-byte aligned
    m.group_classes = S;
    for (int g = 0; g < G; ++g) {
        const int c0 = g * S, cnt = (c0 + S <= C ? S : C - c0);
        if (cnt < 13) return fail(LR_ERR_INVALID, "internal: class group too small");
        Model* sub = new Model();
        sub->device = m.device;
        sub->n_scores = cnt;
        sub->n_features = d;
        sub->n_out = cnt;
        sub->num_sms = m.num_sms;
        sub->W = m.W + (size_t)c0 * d;              // slices of the parent's device copies (not owned)
        sub->b = m.b + c0;
        sub->status = m.status + 4;                 // the groups' own counters; non-finite rows are counted once, by
                                                    // groups_combine_kernel, into the model's
        m.groups.push_back(sub);
        const int rc = gemm_prepare(*sub, coef + (size_t)c0 * d, intercept + c0);
        if (rc != LR_OK) return rc;
        if (!sub->gemm) return fail(LR_ERR_CUDA, "internal: class group has no tensor-path plan");
    }
    return LR_OK;
}

int gemm_groups_load(Model& m, const double* coef, const double* intercept) {
    for (size_t g = 0; g < m.groups.size(); ++g) {
        const size_t c0 = g * (size_t)m.group_classes;
        const int rc = gemm_load(*m.groups[g], coef + c0 * m.n_features, intercept + c0);
        if (rc != LR_OK) return rc;
    }
    return LR_OK;
}

void gemm_groups_destroy(Model& m) {
    for (Model* sub : m.groups) {
        gemm_destroy(*sub);
        delete sub;
    }
    m.groups.clear();
}

int gemm_groups_launch(const Model& m, const ScoreArgs& a) {
    const int G = (int)m.groups.size(), C = m.n_scores, S = m.group_classes;
    if (G == 0) return fail(LR_ERR_INVALID, "model has no class groups");
    if (a.n_rows >= (int64_t)1 << 31) return fail(LR_ERR_INVALID, "gemm path needs fewer than 2^31 rows");
    AsyncScratch scratch;
    scratch.s = a.stream;
    const size_t per = ((size_t)a.n_rows * sizeof(float) + 255) / 256 * 256;
    LRS_CUDA(cudaMallocAsync(&scratch.p, 3 * (size_t)G * per, a.stream));
    char* base = static_cast<char*>(scratch.p);
    float* lse_g = reinterpret_cast<float*>(base);
    int32_t* label_g = reinterpret_cast<int32_t*>(base + (size_t)G * per);
    float* maxp_g = reinterpret_cast<float*>(base + 2 * (size_t)G * per);
    // (the per-group arrays are [G][n_rows], dense; `per` is only the allocation granule)
    for (int g = 0; g < G; ++g) {
        ScoreArgs s = a;
        s.out_pitch = C;
        s.out_col0 = g * S;
        s.lse_out = lse_g + (size_t)g * a.n_rows;
        s.label = label_g + (size_t)g * a.n_rows;
        s.maxprob = maxp_g + (size_t)g * a.n_rows;
        const int rc = gemm_launch(*m.groups[g], s);
        if (rc != LR_OK) return rc;
    }
    // what separates two groups' champions from a tie the tensor path cannot call: the band of its one-accumulator
    // layout at this row length (gemm_launch_aligned)
    const int d = m.n_features;
    const float band = 2.0f * 2.0e-5f * (float)(d > 384 ? d : 384) / 1024.0f;
    const int64_t want = (a.n_rows + 7) / 8;
    const int grid = (int)(want < (int64_t)m.num_sms * 8 ? want : (int64_t)m.num_sms * 8);
    groups_combine_kernel<<<grid, 256, 0, a.stream>>>(a.X, a.n_rows, a.ldx, d, C, G, S, m.W, m.b, band, lse_g, label_g, maxp_g,
                                                      a.proba, a.label, a.maxprob, m.status);
    count_launch();
    if (a.proba) {
        groups_scale_kernel<<<m.num_sms * 8, 256, 0, a.stream>>>(a.proba, a.n_rows, C, S, lse_g);
        count_launch();
    }
    LRS_CUDA(cudaGetLastError());
    return LR_OK;
}

}  // namespace lrs
