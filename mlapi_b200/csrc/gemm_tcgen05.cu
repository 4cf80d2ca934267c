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
// Models of 13..32 classes on dense rows of at most 64 floats take short_rows_kernel instead (below).
//
// Tensor roofline: 2*d*C flops per row against the measured 16-bit tensor peak (TF32: half of it).
#include <cuda.h>
#include <cuda_fp16.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "generic_row.cuh"

namespace lrs {

namespace {

constexpr int BM = 128;          // rows per tile (TMEM lanes)
constexpr int BK = 32;           // TF32 mode: fp32 elements per k-chunk = 128 bytes = one swizzle-128B row
constexpr int BK16 = 64;         // FP16 mode: fp16 elements per k-chunk, again one 128-byte row
constexpr int BN = 256;          // classes per MMA / per W stage
constexpr int A_TILE_BYTES = BM * BK * 4;     // 16 KB: one operand tile of X (hi or lo), either mode
constexpr int kConvWarps = 4;                 // converter warps (all on one k-chunk)
constexpr int kConvThreads = kConvWarps * 32;
constexpr int kLoadWarps = 1;                 // X producer: drives TMA, or copies the rows TMA cannot address itself (a second
                                              // loader warp helps odd pitches, 2.0 -> 2.5 TB/s, but 512 threads leave the
                                              // epilogue short of registers: 56-byte spill frames instead of 24)
constexpr int kThreads = (2 + kConvWarps + 8 + kLoadWarps) * 32;   // warps: TMA (W), MMA, converters, 8 epilogue, X producers
constexpr int kEpiThreads = 256;
constexpr int kMaxCluster = 4;          // class slices per cluster -> up to 2048 classes
constexpr int kMaxAStages = 6;          // X ring depth the barrier block has room for
constexpr int kMaxBStages = 16;         // W ring depth, likewise

// Shared-memory plan.  PAIR = the CTA is half of a cta_group::2 pair: it holds half of every W tile
// (the MMA reads the other half from the peer's shared memory), which halves the shared-memory read
// traffic of each MMA and frees room for a four-deep X ring.
template <bool PAIR>
struct SmemLayoutT {
    static constexpr int a_stages = PAIR ? 4 : 2;
    static constexpr int b_stages = 4;
    static constexpr int b_tile = (PAIR ? BN / 2 : BN) * 128;        // 16 KB / 32 KB
    // offsets from the 1024-byte aligned base
    static constexpr int a_hi = 0;                                   // [a_stages][16 KB]
    static constexpr int a_lo = a_hi + a_stages * A_TILE_BYTES;      // [a_stages][16 KB]
    static constexpr int b = a_lo + a_stages * A_TILE_BYTES;         // [b_stages][b_tile]
    static constexpr int bias = b + b_stages * b_tile;               // [512] floats
    static constexpr int cscale = bias + 512 * 4;                    // [512] floats: 1 / (x scale * class scale), FP16 mode
    static constexpr int xchg = cscale + 512 * 4;                    // [2 tile parities][kMaxCluster][BM] float4 (max, class, sum, -)
    static_assert(2 * kMaxCluster * BM * 16 >= 8 * 32 * 16 * 4, "xchg doubles as the register epilogue's store tiles");
    static constexpr int halves = xchg + 2 * kMaxCluster * BM * 16;  // [BM] float4: column half 1 -> column half 0, CTA-local
    static constexpr int glob = halves + BM * 16;                    // [BM] float4: row statistics of the whole cluster
    static constexpr int chunk_ref = glob + BM * 16;                 // [8][kEpiThreads] floats: the maximum each 32-column
                                                                     // chunk's exponentials (kept in TMEM) are relative to
    static constexpr int bars = chunk_ref + 8 * kEpiThreads * 4;     // mbarriers
    static constexpr int n_bars = 3 * kMaxAStages + 2 * kMaxBStages + 4 + 4;
    static constexpr int tmem_slot = bars + n_bars * 8;
    static constexpr int total = tmem_slot + 16;
};
static_assert(SmemLayoutT<false>::total + 1024 <= 227 * 1024, "shared memory budget");
static_assert(SmemLayoutT<true>::total + 1024 <= 227 * 1024, "shared memory budget");

struct GemmParams {
    int64_t n_rows;
    int n_features;
    int n_classes;
    int nb;                 // 256-class blocks per CTA (1 or 2)
    int split;              // 1: nb == 1 and the correction products accumulate in TMEM columns 256..511
    int stage_off;          // byte offset of the epilogue's 8 x 2 KB store tiles inside the ring area (0: none; see store_block_by_rows)
    int sw_load;            // 1, 2: TMA cannot address X (row pitch or base not a multiple of 16 bytes -- and the innermost box
                            // coordinate has to be one too, so no tensor map over every fourth row helps): the X producer
                            // warp fills the raw tiles with 4-byte (1) or 8-byte (2) cp.async copies from x_raw instead
    const float* x_raw;     // X and its pitch (floats) for that
    int64_t ldx;
    int d_x;                // real row length of X (n_features is the length the W copies are padded to)
    int ncc;                // class slices per cluster: CTAs that share a row tile
    int nrt;                // row tiles per cluster (1 or 2): CTAs that share a class slice
                            // cluster = ncc * nrt CTAs, rank = slice + ncc * row_tile
    int num_k_chunks;
    const float* bias;      // padded to ncc*nb*256, -inf beyond n_classes
    const float* col_inv_scale;   // FP16 mode: 1 / class scale, padded like bias
    const float* x_scale;         // per-call block made by x_scale_kernel: [0] scale applied to X, [1] its inverse,
                                  // [4] arithmetic chosen for this call (bits: 0 = 3xFP16, 1 = 3xTF32), [5] size of a
                                  // typical partial sum of this batch, [6] logit error the fp16 subnormal floor can cause
    const float* col_scale;       // FP16 mode: per-feature power of two applied to X (its inverse is folded into W), padded
                                  // with zeros to whole k-chunks
    float kappa;                  // relative logit error of the truncating TMEM accumulation (layout, arithmetic, row length)
    int acc_stride;               // TMEM columns between the two accumulator buffers (acc_bufs == 2)
    int seg_w;                    // columns reserved per main accumulator when ksplit > 1: 32, 64 or 128
    int seg_slot;                 // TMEM columns a buffer owns in each half: 128 (two buffers) or 256 (one), see seg_offset
    int ksplit;                   // 1..15 main accumulators per tile, each over a contiguous piece of the k-chunks (split
                                  // layout, N <= 128 only: TMEM has room for 3 beside a second buffer when N <= 64, 7 without;
                                  // 7 / 15 when N <= 32; 1 / 3 when N <= 128): the truncation error of a logit grows with
                                  // the accumulation steps behind it, so long rows are summed in pieces
    int count_fallback;           // 3xTF32 kernel launched as the fallback of a 3xFP16 model: add n_rows to status[2]
    int a_stages;           // unpaired kernel: X ring depth (32 KB stages) ...
    int b_tile_bytes;       // ... bytes reserved per W stage ...
    int b_stages;           // ... and W ring depth: a_stages * 32 KB + b_stages * b_tile_bytes <= 192 KB
    int acc_bufs;           // 2: N <= 128, two accumulator buffers 128 TMEM columns apart -- the epilogue of a tile
                            // overlaps the MMAs of the next; 1 otherwise
    int mma_n;              // N of one MMA = rows of one W tile: 256, or the class count rounded up to 16 when that is less (unpaired)
    int debug_skip;         // profiling only: bit0 = converters skip their work, bit1 = epilogue skips passes
    int count_nonfinite;    // 1: the epilogue counts non-finite rows itself (no re-scoring pass to do it)
    float* proba;
    float* decision;
    int32_t* label;
    float* maxprob;
    unsigned long long* status;
    unsigned* flags;        // [n_rows] one word per row, non-zero once the row is on the re-scoring list (several threads
                            // may find a reason to flag the same row)
    int need_prob;          // 0: the caller takes labels / decision values only -- a row is then flagged for near-ties alone,
                            // not for what the error bound could do to its probabilities
    unsigned* flag_list;    // [n_rows] the rows to re-score in fp32 (near-ties, risky probabilities, overflowed rows) ...
    unsigned* n_flagged;    // ... and how many there are
    float* ezr;             // [n_rows] per flagged row: the logit error bound it was flagged under
    float* lse;             // [n_rows] log-sum-exp of the row as the tensor path saw it (for the re-scoring pass)
    int64_t out_pitch;      // floats between rows of proba / decision (n_classes, or the full class count when this
    int out_col0;           // model is one class group of a larger one: then its columns start at out_col0)
    int vec_store;          // 1: rows of proba/decision are 16-byte aligned (C % 4 == 0, base aligned);
                            // 2: 32-byte aligned as well (C % 8 == 0): 256-bit stores, whole sectors per lane
};

// ---- PTX wrappers ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t remote_bar_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote_bar_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .u32 n;\n\t"
        "mov.u32 n, 0;\n"
        "LRS_CWAIT:\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra LRS_CDONE;\n\t"
        "add.u32 n, n, 1;\n\t"
        "setp.lt.u32 p, n, 50000000;\n\t"
        "@p bra LRS_CWAIT;\n\t"
        "trap;\n"
        "LRS_CDONE:\n\t"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
// 4-byte asynchronous copy global -> shared (src_bytes = 0: four zero bytes), and "all my copies so far have landed"
// as an arrival on an mbarrier whose count already includes it
__device__ __forceinline__ void cp_async4(uint32_t dst, const float* src, uint32_t src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_arrive_noinc(uint64_t* bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
// Same copy delivered to the same CTA-relative offset (data and mbarrier signal) in every CTA of `mask`.
__device__ __forceinline__ void tma_load_2d_mc(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar,
                                               uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
        " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
        : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// Arrive on the barrier at the same offset in every CTA of `mask` once the MMAs issued so far are done.
__device__ __forceinline__ void tc_commit_mc(uint64_t* bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(bar)),
                 "h"(mask)
                 : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
template <bool F16>
__device__ __forceinline__ void tc_mma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate);
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
// The same load without the wait: the registers hold the data only after tmem_ld_wait(), which takes them as
// in/out operands so that no use can be scheduled ahead of it.
__device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, float (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7]),
          "=f"(r[8]), "=f"(r[9]), "=f"(r[10]), "=f"(r[11]), "=f"(r[12]), "=f"(r[13]), "=f"(r[14]), "=f"(r[15]),
          "=f"(r[16]), "=f"(r[17]), "=f"(r[18]), "=f"(r[19]), "=f"(r[20]), "=f"(r[21]), "=f"(r[22]), "=f"(r[23]),
          "=f"(r[24]), "=f"(r[25]), "=f"(r[26]), "=f"(r[27]), "=f"(r[28]), "=f"(r[29]), "=f"(r[30]), "=f"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait(float (&r)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+f"(r[0]), "+f"(r[1]), "+f"(r[2]), "+f"(r[3]), "+f"(r[4]), "+f"(r[5]), "+f"(r[6]), "+f"(r[7]),
                   "+f"(r[8]), "+f"(r[9]), "+f"(r[10]), "+f"(r[11]), "+f"(r[12]), "+f"(r[13]), "+f"(r[14]), "+f"(r[15]),
                   "+f"(r[16]), "+f"(r[17]), "+f"(r[18]), "+f"(r[19]), "+f"(r[20]), "+f"(r[21]), "+f"(r[22]), "+f"(r[23]),
                   "+f"(r[24]), "+f"(r[25]), "+f"(r[26]), "+f"(r[27]), "+f"(r[28]), "+f"(r[29]), "+f"(r[30]), "+f"(r[31])
                 :
                 : "memory");
}
// v[0..7] += eight columns (the correction accumulator of the split layout, taken in small pieces: a second
// 32-register buffer next to the pipelined loads would spill)
__device__ __forceinline__ void tmem_add8(uint32_t taddr, float* v) {
    uint32_t r[8];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] += __uint_as_float(r[i]);
}
// sixteen at a time: half the round trips of tmem_add8 where the registers allow it
__device__ __forceinline__ void tmem_add16(uint32_t taddr, float* v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] += __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
        "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
        "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
        "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15])),
        "r"(__float_as_uint(v[16])), "r"(__float_as_uint(v[17])), "r"(__float_as_uint(v[18])), "r"(__float_as_uint(v[19])),
        "r"(__float_as_uint(v[20])), "r"(__float_as_uint(v[21])), "r"(__float_as_uint(v[22])), "r"(__float_as_uint(v[23])),
        "r"(__float_as_uint(v[24])), "r"(__float_as_uint(v[25])), "r"(__float_as_uint(v[26])), "r"(__float_as_uint(v[27])),
        "r"(__float_as_uint(v[28])), "r"(__float_as_uint(v[29])), "r"(__float_as_uint(v[30])), "r"(__float_as_uint(v[31]))
        : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
// 256-bit streaming store (sm_100): one full 32-byte sector per lane and instruction
__device__ __forceinline__ void st_cs_v8(float* p, const float* v) {
    asm volatile("st.global.cs.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]),
                 "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7])
                 : "memory");
}
// A warp's 32 rows x 32 columns (lane = row, v = its columns) to global memory through a tile in shared memory,
// so that a store instruction writes the columns of two rows side by side (2 x 64 contiguous bytes) instead of one
// column of 32 rows (32 sectors per instruction, which is what scalar per-row stores cost when the row pitch rules
// vectors out: 1.4-1.8x on the whole step, measured on 101 / 150 / 255 / 1001 classes).  Two sweeps of 16 columns;
// `stage` = 32 x 16 floats per warp (2 KB), column index xor-ed with the row so that the reads are conflict-free
// and the writes two-way.  `dst` = column 0 of the warp's first row.
constexpr int kStoreTileBytes = 32 * 16 * 4;
__device__ __forceinline__ void store_block_by_rows(float* __restrict__ stage, const float (&v)[32], int lane,
                                                    float* __restrict__ dst, int64_t pitch, int rows_valid, int n_cols) {
    const int col = lane & 15, rsel = lane >> 4;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        if (half * 16 < n_cols) {
#pragma unroll
            for (int i = 0; i < 16; ++i) stage[lane * 16 + (i ^ (lane & 15))] = v[half * 16 + i];
            __syncwarp();
            if (half * 16 + col < n_cols) {
                float* o = dst + half * 16 + col;
#pragma unroll 4
                for (int r = rsel; r < rows_valid; r += 2) __stcs(o + r * pitch, stage[r * 16 + (col ^ (r & 15))]);
            }
            __syncwarp();
        }
    }
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// Two scaled values -> packed fp16 (hi, hi) and packed fp16 residuals (lo, lo), with the packed
// fp32x2 / f16x2 conversions of sm_100: 7 instructions per pair.
__device__ __forceinline__ void split_pair(float a, float b, float sa, float sb, uint32_t& hi, uint32_t& lo) {
    asm("{\n\t"
        ".reg .b64 v, s, r;\n\t"
        ".reg .f32 ha, hb;\n\t"
        ".reg .b16 h0, h1;\n\t"
        "mov.b64 v, {%2, %3};\n\t"
        "mov.b64 s, {%4, %5};\n\t"
        "mul.rn.f32x2 v, v, s;\n\t"
        "mov.b64 {ha, hb}, v;\n\t"
        "cvt.rn.f16x2.f32 %0, hb, ha;\n\t"            // low half <- a, high half <- b
        "mov.b32 {h0, h1}, %0;\n\t"
        "cvt.f32.f16 ha, h0;\n\t"
        "cvt.f32.f16 hb, h1;\n\t"
        "mov.b64 r, {ha, hb};\n\t"
        "sub.rn.f32x2 v, v, r;\n\t"
        "mov.b64 {ha, hb}, v;\n\t"
        "cvt.rn.f16x2.f32 %1, hb, ha;\n\t"
        "}"
        : "=r"(hi), "=r"(lo)
        : "f"(a), "f"(b), "f"(sa), "f"(sb));
}
__device__ __forceinline__ float to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

// K-major, 128-byte-swizzled operand tile (rows of 128 B, 8-row atoms of 1024 B).
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
    uint64_t desc = 0;
    desc |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);          // start address
    desc |= (uint64_t)1 << 16;                                // leading byte offset (unused for SW128 K-major)
    desc |= (uint64_t)(1024 >> 4) << 32;                      // stride byte offset: 8 rows x 128 B
    desc |= (uint64_t)1 << 46;                                // descriptor version (sm_100)
    desc |= (uint64_t)2 << 61;                                // SWIZZLE_128B
    return desc;
}
constexpr uint32_t kIdesc = (1u << 4)      // D = f32
                            | (2u << 7)    // A = tf32
                            | (2u << 10)   // B = tf32
                            | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);   // K-major A and B

constexpr uint32_t kIdesc16 = (1u << 4)    // D = f32, A = B = f16 (format 0), K-major
                              | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
template <>
__device__ __forceinline__ void tc_mma<false>(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) { tc_mma_tf32(d, a, b, idesc, acc); }
template <>
__device__ __forceinline__ void tc_mma<true>(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) { tc_mma_f16(d, a, b, idesc, acc); }

constexpr float kLog2e = 1.4426950408889634f;

// TMEM columns of main accumulator `seg` relative to the tile's first one (GemmParams::ksplit; split layout, the
// correction accumulator BN columns up).  `slot` = columns a buffer owns in each half of TMEM: 128 with two
// accumulator buffers, 256 with one.  `w` = 128, 64 or 32 columns per accumulator (N <= 128 / 64 / 32): slot / w of
// them in the main half, slot / w - 1 more beside the correction accumulator (two buffers: 1, 3 or 7 in all; one
// buffer: 3, 7 or 15).
__device__ __forceinline__ uint32_t seg_offset(int seg, int w, int slot) {
    const int per_slot = slot / w;
    return seg < per_slot ? (uint32_t)(seg * w) : (uint32_t)BN + (uint32_t)((seg - per_slot + 1) * w);
}

// Put `row` on the re-scoring list, once; `ez` = the logit error the row was judged against (every thread that
// flags a row has the same value for it).
__device__ __forceinline__ void flag_row(unsigned* flags, unsigned* list, unsigned* count, float* ezr, int64_t row, float ez) {
    if (atomicExch(flags + row, 1u) == 0u) {
        if (ezr) ezr[row] = ez;
        list[atomicAdd(count, 1u)] = (unsigned)row;
    }
}

// ---- cta_group::2 (CTA pair) forms ---------------------------------------------------------------
// The pair shares every MMA: the even CTA of the pair issues it, each SM multiplies its own 128 rows
// of X with a W tile whose two halves sit in the two CTAs' shared memories.
__device__ __forceinline__ void tma_load_2d_2sm(void* dst, const CUtensorMap* map, int c0, int c1,
                                                uint32_t bar_cluster_addr) {
    // data lands in THIS CTA's shared memory, the byte count is credited to the barrier at
    // `bar_cluster_addr` (a shared::cluster address: the issuing CTA's copy of the barrier)
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tc_commit_2sm(uint64_t* bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(bar)),
                 "h"(mask)
                 : "memory");
}
template <bool F16>
__device__ __forceinline__ void tc_mma_2sm(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc1,
                                           uint32_t accumulate) {
    const uint32_t idesc = idesc1 + ((uint32_t)(BM >> 4) << 24);   // M = 256 across the pair
    if constexpr (F16)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                     "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
                     : "memory");
    else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                     "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
                     : "memory");
}
// `idesc` (unpaired form only): instruction descriptor with the model's N, see GemmParams::mma_n
template <bool F16, bool PAIR>
__device__ __forceinline__ void mma_issue(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    if constexpr (PAIR) tc_mma_2sm<F16>(d, a, b, idesc, acc);
    else tc_mma<F16>(d, a, b, idesc, acc);
}

// The epilogue of a model with at most 64 classes: whole rows stay in registers, one lane per row.
// One read of TMEM, everything else in registers: no write-back, no second pass over TMEM (the general form spent
// 2.4 us per 128-row tile on a 32-class model, which is what bounded it).  A warp holds one 32-column chunk of its
// 32 rows (TMEM lets only the two warps of a lane quadrant read those rows): up to 32 classes the four warps of
// column half 0 do it alone, with 33..64 the warps of half 1 take columns 32..63 of the same rows and the two warps
// of a quadrant meet twice in shared memory (row maximum, then sum), 64 threads per named barrier.  Same arithmetic
// and the same re-scoring flags as the general form.  (65..128 classes: epilogue_in_registers2.)
// `xch` = 8 x 128 floats of scratch, `stg` = the warp's store tile (store_block_by_rows).
template <bool F16>
__device__ __forceinline__ void epilogue_in_registers(const GemmParams& p, uint32_t t_lane, int h, int q, int lane, int n_real,
                                                      int64_t row, bool live, int rows_valid, int64_t blk_w,
                                                      const float* cscale_s, const float* bias_s, float* xch, float* stg,
                                                      uint64_t* acc_empty_bar, float psum_scale, float floor_err) {
    const bool two = n_real > 32;                     // the warps of column half 1 take part
    if (h == 1 && !two) {
        tc_fence_before();
        mbar_arrive(acc_empty_bar);
        return;
    }
    const int r_in_tile = q * 32 + lane;
    const int cbase = h * 32;
    const int n_here = n_real - cbase < 32 ? n_real - cbase : 32;   // real classes in this warp's columns
    const uint32_t t_col = t_lane + (uint32_t)cbase;
    float v[32];
    tmem_ld32(t_col, v);
    if (p.split) {
        float u[32];
        tmem_ld32(t_col + (uint32_t)BN, u);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += u[i];
        for (int sg = 1; sg < p.ksplit; ++sg) {
            tmem_ld32(t_col + seg_offset(sg, p.seg_w, p.seg_slot), u);
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] += u[i];
        }
    }
    tc_fence_before();
    mbar_arrive(acc_empty_bar);                       // the accumulators are in registers now
    float gm = -INFINITY;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        if constexpr (F16) v[i] = fmaf(v[i], cscale_s[cbase + i], bias_s[cbase + i]);
        else v[i] += bias_s[cbase + i];               // padding carries -inf
        gm = fmaxf(gm, v[i]);
    }
    int garg = 31;
#pragma unroll
    for (int i = 30; i >= 0; --i) garg = (v[i] == gm) ? i : garg;   // first index wins
    garg += cbase;
    if (two) {
        float* xm = xch;                              // [2 halves][128 rows] maxima, then classes
        int* xa = reinterpret_cast<int*>(xch + 2 * BM);
        xm[h * BM + r_in_tile] = gm;
        xa[h * BM + r_in_tile] = garg;
        asm volatile("bar.sync %0, 64;" ::"r"(3 + q) : "memory");
        const float om = xm[(h ^ 1) * BM + r_in_tile];
        const int oa = xa[(h ^ 1) * BM + r_in_tile];
        if (h == 0 ? om > gm : om >= gm) { gm = om; garg = oa; }   // ties: the lower class, i.e. half 0
    }
    const int64_t blk = blk_w + cbase;
    if (p.decision) {
        if (n_here == 32 && p.vec_store) {
            if (live) {
                float* zr = p.decision + row * p.out_pitch + p.out_col0 + cbase;
#pragma unroll
                for (int i = 0; i < 32; i += 4)
                    __stcs(reinterpret_cast<float4*>(zr + i), make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]));
            }
        } else {
            store_block_by_rows(stg, v, lane, p.decision + blk, p.out_pitch, rows_valid, n_here);
        }
    }
    const float ez = fmaxf(p.kappa * fmaxf(psum_scale, fabsf(gm)) + floor_err, 1e-30f);
    const float tau = gm - 2.0f * ez;
    const float gml = gm * kLog2e;
    float gs = 0.0f;
    int gnear = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        gnear += v[i] > tau ? 1 : 0;
        v[i] = ex2(fmaf(v[i], kLog2e, -gml));
        gs += v[i];
    }
    if (two) {
        float* xs = xch + 4 * BM;                     // [2 halves][128 rows] sums, then near-tie counts
        int* xn = reinterpret_cast<int*>(xch + 6 * BM);
        xs[h * BM + r_in_tile] = gs;
        xn[h * BM + r_in_tile] = gnear;
        asm volatile("bar.sync %0, 64;" ::"r"(3 + q) : "memory");
        gs += xs[(h ^ 1) * BM + r_in_tile];           // a + b == b + a: both warps hold the same sum
        gnear += xn[(h ^ 1) * BM + r_in_tile];
    }
    const float inv = __frcp_rn(gs);
    const float risk_thr = p.need_prob ? 8e-6f / ez : 1.0f;      // (p (1 - p) <= 1/4: never)
    if (p.proba) {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] *= inv;
        if (n_here == 32 && p.vec_store == 2) {
            if (live) {
                float* pr = p.proba + row * p.out_pitch + p.out_col0 + cbase;
#pragma unroll
                for (int i = 0; i < 32; i += 8) st_cs_v8(pr + i, v + i);
            }
        } else {
            store_block_by_rows(stg, v, lane, p.proba + blk, p.out_pitch, rows_valid, n_here);
        }
    }
    if (live && h == 0) {
        if (p.label) __stcs(p.label + row, garg);
        if (p.maxprob) __stcs(p.maxprob + row, inv);      // exp(gm - gm) == 1
        const bool finite_row = finite_f(gm) && finite_f(gs);
        if (!finite_row) {
            if (p.flags && !p.count_nonfinite) flag_row(p.flags, p.flag_list, p.n_flagged, p.ezr, row, ez);
            else atomicAdd(p.status, 1ull);
        }
        if (p.lse) p.lse[row] = gm + logf(gs);
        if (p.flags && finite_row && (gnear >= 2 || inv * (1.0f - inv) > risk_thr))
            flag_row(p.flags, p.flag_list, p.n_flagged, p.ezr, row, ez);
    }
}

// 65..128 classes: the same single read of TMEM with TWO chunks per warp (columns 64 h + {0, 32}), taken one after
// the other so that at most one chunk's raw accumulators and sixteen columns of the correction accumulator are live
// beside the first chunk's logits (both chunks loaded at once spilled 270 bytes per thread and lost to the general
// form).  Everything else as in epilogue_in_registers.
// (A kernel instantiation of its own, EPI2: compiled into the common kernel its 64 live values pushed ptxas into
// spilling loop-carried state of the OTHER epilogue forms -- d=128, 16 classes lost 9 % -- and as a non-inlined
// function it forced the kernel's parameter block into local memory, which cost everybody 5-40 %.)
template <bool F16>
__device__ __forceinline__ void epilogue_in_registers2(const GemmParams& p, uint32_t t_lane, int h, int q, int lane, int n_real,
                                                       int64_t row, bool live, int rows_valid, int64_t blk_w,
                                                       const float* cscale_s, const float* bias_s, float* xch, float* stg,
                                                       uint64_t* acc_empty_bar, float psum_scale, float floor_err) {
    const int r_in_tile = q * 32 + lane;
    const int cbase = h * 64;
    const int n_a = n_real - cbase >= 32 ? 32 : n_real - cbase;                            // >= 1 (n_real > 64)
    const int n_b = n_real - cbase - 32 >= 32 ? 32 : (n_real > cbase + 32 ? n_real - cbase - 32 : 0);
    float a[32], bb[32];
    float m_a = -INFINITY, m_b = -INFINITY;
    int arg_a = 0, arg_b = 0;
    auto chunk_logits = [&](float (&v)[32], int col, float& cm, int& carg) {
        const uint32_t t_col = t_lane + (uint32_t)col;
        tmem_ld32(t_col, v);
        if (p.split) {
#pragma unroll
            for (int i8 = 0; i8 < 32; i8 += 16) tmem_add16(t_col + (uint32_t)(BN + i8), v + i8);
            for (int sg = 1; sg < p.ksplit; ++sg)
#pragma unroll
                for (int i8 = 0; i8 < 32; i8 += 16) tmem_add16(t_col + seg_offset(sg, p.seg_w, p.seg_slot) + (uint32_t)i8, v + i8);
        }
#pragma unroll
        for (int i8 = 0; i8 < 32; i8 += 8) {
#pragma unroll
            for (int i = i8; i < i8 + 8; ++i) {
                if constexpr (F16) v[i] = fmaf(v[i], cscale_s[col + i], bias_s[col + i]);
                else v[i] += bias_s[col + i];             // padding carries -inf
                cm = fmaxf(cm, v[i]);
            }
            asm volatile("" ::: "memory");               // (keeps later groups' scale / bias loads out of the registers)
        }
        int idx = 31;
#pragma unroll
        for (int i = 30; i >= 0; --i) idx = (v[i] == cm) ? i : idx;   // first index wins
        carg = col + idx;
    };
    chunk_logits(a, cbase, m_a, arg_a);
    if (n_b > 0) {
        chunk_logits(bb, cbase + 32, m_b, arg_b);
    } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) bb[i] = -INFINITY;
    }
    tc_fence_before();
    mbar_arrive(acc_empty_bar);                       // the accumulators are in registers now
    float gm = m_a;
    int garg = arg_a;
    if (m_b > m_a) { gm = m_b; garg = arg_b; }        // strict: the lower chunk keeps a tie
    {
        float* xm = xch;                              // [2 halves][128 rows] maxima, then classes
        int* xa = reinterpret_cast<int*>(xch + 2 * BM);
        xm[h * BM + r_in_tile] = gm;
        xa[h * BM + r_in_tile] = garg;
        asm volatile("bar.sync %0, 64;" ::"r"(3 + q) : "memory");
        const float om = xm[(h ^ 1) * BM + r_in_tile];
        const int oa = xa[(h ^ 1) * BM + r_in_tile];
        if (h == 0 ? om > gm : om >= gm) { gm = om; garg = oa; }   // ties: the lower class, i.e. half 0
    }
    const int64_t blk = blk_w + cbase;
    if (p.decision) {
        store_block_by_rows(stg, a, lane, p.decision + blk, p.out_pitch, rows_valid, n_a);
        if (n_b > 0) store_block_by_rows(stg, bb, lane, p.decision + blk + 32, p.out_pitch, rows_valid, n_b);
    }
    const float ez = fmaxf(p.kappa * fmaxf(psum_scale, fabsf(gm)) + floor_err, 1e-30f);
    const float tau = gm - 2.0f * ez;
    const float gml = gm * kLog2e;
    float gs0 = 0.0f, gs1 = 0.0f;
    int gnear = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        gnear += (a[i] > tau ? 1 : 0) + (bb[i] > tau ? 1 : 0);
        a[i] = ex2(fmaf(a[i], kLog2e, -gml));
        bb[i] = ex2(fmaf(bb[i], kLog2e, -gml));
        gs0 += a[i];
        gs1 += bb[i];
    }
    float gs = gs0 + gs1;
    {
        float* xs = xch + 4 * BM;                     // [2 halves][128 rows] sums, then near-tie counts
        int* xn = reinterpret_cast<int*>(xch + 6 * BM);
        xs[h * BM + r_in_tile] = gs;
        xn[h * BM + r_in_tile] = gnear;
        asm volatile("bar.sync %0, 64;" ::"r"(3 + q) : "memory");
        gs += xs[(h ^ 1) * BM + r_in_tile];           // a + b == b + a: both warps hold the same sum
        gnear += xn[(h ^ 1) * BM + r_in_tile];
    }
    const float inv = __frcp_rn(gs);
    const float risk_thr = p.need_prob ? 8e-6f / ez : 1.0f;      // (p (1 - p) <= 1/4: never)
    if (p.proba) {
#pragma unroll
        for (int i = 0; i < 32; ++i) { a[i] *= inv; bb[i] *= inv; }
        if (p.vec_store == 2 && n_a == 32) {
            if (live) {
                float* pr = p.proba + row * p.out_pitch + p.out_col0 + cbase;
#pragma unroll
                for (int i = 0; i < 32; i += 8) st_cs_v8(pr + i, a + i);
            }
        } else {
            store_block_by_rows(stg, a, lane, p.proba + blk, p.out_pitch, rows_valid, n_a);
        }
        if (p.vec_store == 2 && n_b == 32) {
            if (live) {
                float* pr = p.proba + row * p.out_pitch + p.out_col0 + cbase + 32;
#pragma unroll
                for (int i = 0; i < 32; i += 8) st_cs_v8(pr + i, bb + i);
            }
        } else if (n_b > 0) {
            store_block_by_rows(stg, bb, lane, p.proba + blk + 32, p.out_pitch, rows_valid, n_b);
        }
    }
    if (live && h == 0) {
        if (p.label) __stcs(p.label + row, garg);
        if (p.maxprob) __stcs(p.maxprob + row, inv);      // exp(gm - gm) == 1
        const bool finite_row = finite_f(gm) && finite_f(gs);
        if (!finite_row) {
            if (p.flags && !p.count_nonfinite) flag_row(p.flags, p.flag_list, p.n_flagged, p.ezr, row, ez);
            else atomicAdd(p.status, 1ull);
        }
        if (p.lse) p.lse[row] = gm + logf(gs);
        if (p.flags && finite_row && (gnear >= 2 || inv * (1.0f - inv) > risk_thr))
            flag_row(p.flags, p.flag_list, p.n_flagged, p.ezr, row, ez);
    }
}

// F16 = false: 3xTF32 (operands are fp32 containers, 8 K-elements per MMA).
// F16 = true : 3xFP16 (x and w scaled by powers of two into fp16 range and split hi/lo like above,
//              16 K-elements per MMA at twice the TF32 rate; scales are undone in the epilogue).
// PAIR = true: cta_group::2 -- see SmemLayoutT.
template <bool F16, bool PAIR, bool EPI2 = false>
__global__ void __launch_bounds__(kThreads, 1)
gemm_softmax_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_whi,
                    const __grid_constant__ CUtensorMap map_wlo, const GemmParams p) {
    // The arithmetic of a call is chosen on the device (x_scale_kernel: 3xFP16 unless the data's dynamic range
    // would cost precision); both kernels are enqueued and the one not chosen leaves here, every CTA alike.
    if (reinterpret_cast<const uint32_t*>(p.x_scale)[4] != (F16 ? 0u : 1u)) return;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(p.status + 3, (unsigned long long)p.n_rows);          // rows this path has scored (layout feedback)
        if (!F16 && p.count_fallback) atomicAdd(p.status + 2, (unsigned long long)p.n_rows);
    }
    extern __shared__ unsigned char smem_dyn[];
    // 1024-byte alignment by pointer arithmetic on the shared array itself: going through uintptr_t made
    // every access below a generic LD/ST instead of LDS/STS
    unsigned char* smem = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);

    using SmemLayout = SmemLayoutT<PAIR>;
    // Unpaired: a model with few classes has small W tiles and trades their room for a deeper X ring
    // (GemmParams::a_stages / b_tile_bytes; same 192 KB in total).  Paired: the fixed plan.
    const int A_STAGES = p.a_stages;
    const int B_STAGES = p.b_stages;
    const int B_TILE_BYTES = PAIR ? SmemLayout::b_tile : p.b_tile_bytes;
    unsigned char* a_hi = smem + SmemLayout::a_hi;
    unsigned char* a_lo = a_hi + A_STAGES * A_TILE_BYTES;
    unsigned char* b_t = a_lo + A_STAGES * A_TILE_BYTES;
    constexpr int BKx = F16 ? BK16 : BK;                 // K-elements per chunk
    constexpr int RAW_BYTES = BM * BKx * 4;              // raw fp32 X tile: 16 KB (TF32) / 32 KB (FP16)
    constexpr int A_STAGE = 2 * A_TILE_BYTES;            // FP16: one contiguous 32 KB stage (raw -> hi | lo in place)
    float* bias_s = reinterpret_cast<float*>(smem + SmemLayout::bias);
    float* cscale_s = reinterpret_cast<float*>(smem + SmemLayout::cscale);
    float4* xchg4 = reinterpret_cast<float4*>(smem + SmemLayout::xchg);
    float4* halves4 = reinterpret_cast<float4*>(smem + SmemLayout::halves);
    float4* glob4 = reinterpret_cast<float4*>(smem + SmemLayout::glob);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SmemLayout::bars);
    uint64_t* x_full = bars;                       // [A_STAGES] TMA -> converters
    uint64_t* a_ready = x_full + kMaxAStages;      // [A_STAGES] converters -> MMA
    uint64_t* a_empty = a_ready + kMaxAStages;     // [A_STAGES] MMA -> TMA
    uint64_t* b_full = a_empty + kMaxAStages;      // [B_STAGES] TMA -> MMA
    uint64_t* b_empty = b_full + kMaxBStages;      // [B_STAGES] MMA -> TMA
    uint64_t* acc_full = b_empty + kMaxBStages;       // [2 accumulator buffers] MMA -> epilogue
    uint64_t* acc_empty = acc_full + 2;            // [2] epilogue -> MMA
    uint64_t* xchg_bar = acc_empty + 2;            // [2 tile parities] row statistics from every CTA of the cluster
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + SmemLayout::tmem_slot);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    // cluster layout: PAIR: 2 * ncc CTAs, rank = 2 * slice + row tile (pair = two row tiles of one class slice);
    //                 else: ncc * nrt CTAs, rank = slice + ncc * row tile
    const int gsz = PAIR ? 2 : p.nrt;                   // row tiles per cluster iteration
    const int csize = p.ncc * gsz;
    const uint32_t crank = csize > 1 ? cluster_ctarank() : 0;
    const uint32_t rank = PAIR ? (crank >> 1) : crank % (uint32_t)p.ncc;      // class slice of this CTA
    const uint32_t rt = PAIR ? (crank & 1u) : crank / (uint32_t)p.ncc;        // row tile inside the cluster's tile group
    const bool lead = !PAIR || rt == 0;                 // the CTA whose MMA lane issues for the pair
    const uint32_t lead_rank = crank & ~1u;
    const uint16_t pair_mask = (uint16_t)(3u << lead_rank);
    // X is the caller's memory: every CTA fetches its own copy of the tile with a plain TMA load.  (TMA
    // multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another
    // GPU's memory hung the GPU -- and nothing stops a caller from handing in a peer's buffer.)
    const bool w_mc = !PAIR && p.nrt > 1;               // W (the library's own copy, on this GPU) multicast across row tiles
    // classes of this CTA: nb blocks of mma_n (the class count is spread evenly over the blocks of all slices, in
    // whole 32-column chunks: 300 classes are 2 x 160, not 2 x 256); block k's accumulator starts at TMEM column k * BN
    const int n_cols_cta = p.nb * p.mma_n;
    const int col0 = (int)rank * n_cols_cta;       // first class handled by this CTA
    // Multicast group of W: the nrt CTAs of a class slice need the same W tiles, each fetches 1/nrt of the rows.
    uint16_t w_mask = 0;
    for (int j = 0; j < p.nrt; ++j) w_mask |= (uint16_t)(1u << (rank + j * p.ncc));
    const int w_rows = PAIR ? p.mma_n / 2 : p.mma_n / p.nrt;  // rows of each W tile this CTA fetches

    if (threadIdx.x == 0) {
        for (int i = 0; i < kMaxAStages; ++i) {     // (operand ring and raw ring may differ in depth)
            mbar_init(x_full + i, p.sw_load ? kLoadWarps * 32 : 1);
            mbar_init(a_ready + i, PAIR ? 2 * kConvThreads : kConvThreads);    // PAIR: both CTAs' converters report to the issuing CTA
            mbar_init(a_empty + i, 1);
        }
        for (int i = 0; i < B_STAGES; ++i) {
            mbar_init(b_full + i, 1);
            mbar_init(b_empty + i, w_mc ? p.nrt : 1);    // every CTA sharing the W tile has consumed the stage
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(acc_full + i, 1);
            mbar_init(acc_empty + i, PAIR ? 2 * kEpiThreads : kEpiThreads);
        }
        for (int i = 0; i < 4; ++i) mbar_init(xchg_bar + i, 128 * p.ncc);
        mbar_fence_init();
    }
    for (int i = threadIdx.x; i < n_cols_cta; i += kThreads) {
        bias_s[i] = p.bias[col0 + i];
        if constexpr (F16) cscale_s[i] = p.col_inv_scale[col0 + i] * p.x_scale[1];
    }
    if (warp == 1) {
        if constexpr (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                         "r"(512)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                         "r"(512)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (csize > 1) cluster_sync_all();             // peers' barriers are initialised before any remote arrive
    const uint32_t tmem_base = *tmem_slot;

    const int64_t n_tiles = (p.n_rows + BM - 1) / BM;
    // tiles are handed out in groups of nrt consecutive row tiles, one group per cluster iteration
    const int64_t n_groups = (n_tiles + gsz - 1) / gsz;
    const int64_t cluster_id = blockIdx.x / csize;
    const int64_t n_clusters = gridDim.x / csize;

    if (warp == 0) {
        // ===================== TMA producer: W tiles =====================
        // (whole warp walks the loop, one elected lane issues: see the MMA issuer for why)
        // X has a producer warp of its own (below): issued from one program-ordered loop, the next X
        // tile waited behind the W slots of the current chunk, which free up only as the MMAs of the
        // previous chunk retire -- the HBM latency of X plus the conversion then sat on the critical
        // path of every chunk (measured: 5300 cycles per chunk against 3072 of MMA work).
        const bool leader = elect_one();
        int bs = 0;
        uint32_t bph = 0;
        for (int64_t grp = cluster_id; grp < n_groups; grp += n_clusters) {
            for (int kc = 0; kc < p.num_k_chunks; ++kc) {
                for (int part = 0; part < 2; ++part) {            // 0: w_hi, 1: w_lo
                    for (int nbk = 0; nbk < p.nb; ++nbk) {
                        mbar_wait(b_empty + bs, bph ^ 1);
                        if (leader && PAIR) {
                            // each CTA of the pair fetches its half of the tile into its own shared
                            // memory; both transfers are counted on the issuing CTA's barrier
                            if (lead) mbar_expect_tx(b_full + bs, p.mma_n * 128);
                            tma_load_2d_2sm(b_t + bs * B_TILE_BYTES, part == 0 ? &map_whi : &map_wlo, kc * BKx,
                                            col0 + nbk * p.mma_n + (int)rt * w_rows,
                                            map_to_cta(smem_u32(b_full + bs), lead_rank));
                        } else if (leader) {
                            mbar_expect_tx(b_full + bs, p.mma_n * 128);
                            if (w_mc)
                                tma_load_2d_mc(b_t + bs * B_TILE_BYTES + rt * w_rows * 128,
                                               part == 0 ? &map_whi : &map_wlo, kc * BKx,
                                               col0 + nbk * p.mma_n + (int)rt * w_rows, b_full + bs, w_mask);
                            else
                                tma_load_2d(b_t + bs * B_TILE_BYTES, part == 0 ? &map_whi : &map_wlo, kc * BKx,
                                            col0 + nbk * p.mma_n, b_full + bs);
                        }
                        __syncwarp();
                        if (++bs == B_STAGES) { bs = 0; bph ^= 1; }
                    }
                }
            }
        }
    } else if (warp >= 2 + kConvWarps + 8 && p.sw_load) {
        // ===================== X producers, rows TMA cannot address =====================
        // The warp copies the 128 rows of the tile; a warp instruction = consecutive floats of one row
        // (coalesced), 4 bytes per lane -- 8 where every row is 8-byte aligned (sw_load == 2: even pitch, e.g. the
        // dense rows of an even feature count).  Rows past the end and columns past the row length arrive as zeros,
        // like TMA's out-of-bounds fill.  The copies are asynchronous: a stage's instructions are issued back to
        // back and the stage's barrier counts every lane's "all landed".  A tile and k-chunk that lie fully inside
        // the batch take a loop without predicates or pointer selects: one warp issues dependent integer work at
        // one instruction per ~4 cycles, and with ~10 instructions per copy the loader managed 1.8 TB/s.
        const int lw = warp - (2 + kConvWarps + 8);
        constexpr int ROWS = BM / kLoadWarps;
        int as = 0;
        uint32_t aph = 0;
        for (int64_t grp = cluster_id; grp < n_groups; grp += n_clusters) {
            const int64_t tile = grp * gsz + rt;
            const int64_t row_lo = tile * BM + lw * ROWS;
            for (int kc = 0; kc < p.num_k_chunks; ++kc) {
                mbar_wait(a_empty + as, aph ^ 1);
                const int k0 = kc * BKx;
                const uint32_t xdst = smem_u32(F16 ? a_hi + as * A_STAGE : a_hi + as * A_TILE_BYTES) +
                                      (uint32_t)(lw * ROWS * BKx * 4);
                if (row_lo + ROWS <= p.n_rows && k0 + BKx <= p.d_x) {
                    const float* src = p.x_raw + row_lo * p.ldx + k0;
                    if constexpr (F16) {                 // raw tile: [128][64] floats, linear
                        if (p.sw_load == 2) {
                            src += 2 * lane;
                            const uint32_t dst = xdst + (uint32_t)(lane * 8);
#pragma unroll 8
                            for (int r = 0; r < ROWS; ++r, src += p.ldx)
                                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst + (uint32_t)(r * 256)), "l"(src) : "memory");
                        } else {
                            src += lane;
                            const uint32_t dst = xdst + (uint32_t)(lane * 4);
#pragma unroll 8
                            for (int r = 0; r < ROWS; ++r, src += p.ldx) {
                                cp_async4(dst + (uint32_t)(r * 256), src, 4u);
                                cp_async4(dst + (uint32_t)(r * 256 + 128), src + 32, 4u);
                            }
                        }
                    } else {                             // [128][32] floats in the operand's 128-byte swizzle
                        src += lane;
                        const uint32_t dst = xdst + (uint32_t)((lane & 3) * 4);
#pragma unroll 8
                        for (int r = 0; r < ROWS; ++r, src += p.ldx)      // (ROWS % 8 == 0: r & 7 is the tile row's)
                            cp_async4(dst + (uint32_t)(r * 128 + (((lane >> 2) ^ (r & 7)) << 4)), src, 4u);
                    }
                } else {
                    for (int r = 0; r < ROWS; ++r) {
                        const bool rv = row_lo + r < p.n_rows;
                        const float* src = p.x_raw + (rv ? row_lo + r : 0) * p.ldx + k0;
                        if constexpr (F16) {
#pragma unroll
                            for (int c = lane; c < BKx; c += 32) {
                                const bool ok = rv && k0 + c < p.d_x;
                                cp_async4(xdst + (uint32_t)(r * (BKx * 4) + c * 4), ok ? src + c : p.x_raw, ok ? 4u : 0u);
                            }
                        } else {
                            const bool ok = rv && k0 + lane < p.d_x;
                            cp_async4(xdst + (uint32_t)(r * 128 + (((lane >> 2) ^ (r & 7)) << 4) + (lane & 3) * 4),
                                      ok ? src + lane : p.x_raw, ok ? 4u : 0u);
                        }
                    }
                }
                cp_async_arrive_noinc(x_full + as);
                if (++as == A_STAGES) { as = 0; aph ^= 1; }
            }
        }
    } else if (warp == 2 + kConvWarps + 8) {
        // ===================== TMA producer: X tiles =====================
        const bool leader = elect_one();
        int as = 0;
        uint32_t aph = 0;
        for (int64_t grp = cluster_id; grp < n_groups; grp += n_clusters) {
            const int64_t tile = grp * gsz + rt;
            const int row0 = (int)(tile * BM);
            for (int kc = 0; kc < p.num_k_chunks; ++kc) {
                mbar_wait(a_empty + as, aph ^ 1);
                if (leader) {
                    mbar_expect_tx(x_full + as, RAW_BYTES);
                    unsigned char* xdst = F16 ? a_hi + as * A_STAGE : a_hi + as * A_TILE_BYTES;
                    tma_load_2d(xdst, &map_x, kc * BKx, row0, x_full + as);
                }
                __syncwarp();
                if (++as == A_STAGES) { as = 0; aph ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        // The whole warp walks the pipeline (waits included) and only lane 0 issues: with warp-uniform
        // control flow the compiler keeps stage indices, shared-memory descriptors and TMEM addresses
        // in uniform registers, which is what UTCHMMA takes.  Under `if (lane == 0) { loop }` every
        // operand was re-broadcast (R2UR) per instruction, ~14 instructions per MMA, and the issuing
        // lane -- not the tensor pipe -- set the pace (ncu: tensor pipe 61 % active).
        const bool leader = elect_one() && lead;                      // PAIR: only the even CTA issues
        const uint64_t desc0 = make_smem_desc(0);                    // every field but the address
        const uint64_t ahi0 = desc0 | (uint64_t)((smem_u32(a_hi) & 0x3FFFF) >> 4);
        const uint64_t alo0 = desc0 | (uint64_t)((smem_u32(a_lo) & 0x3FFFF) >> 4);
        const uint64_t b0 = desc0 | (uint64_t)((smem_u32(b_t) & 0x3FFFF) >> 4);
        const uint32_t c_off = p.split ? (uint32_t)BN : 0u;          // split: corrections -> columns 256..511
        const uint32_t corr_first = p.split ? 0u : 1u;               // accumulate flag of the first correction MMA
        // N of the instruction = the model's class count rounded up to 16 when one W tile holds them all
        const uint32_t idesc = ((F16 ? kIdesc16 : kIdesc) & ~(0x3Fu << 17)) | ((uint32_t)(p.mma_n >> 3) << 17);
        int as = 0, bs = 0;
        uint32_t aph = 0, bph = 0, it = 0;
        for (int64_t grp = cluster_id; lead && grp < n_groups; grp += n_clusters, ++it) {
            const uint32_t buf = p.acc_bufs == 2 ? (it & 1u) : 0u;                  // accumulator buffer of this tile
            const uint32_t accph = (p.acc_bufs == 2 ? (it >> 1) : it) & 1u;         // its use count, as a parity
            if constexpr (PAIR) mbar_wait_cluster(acc_empty + buf, accph ^ 1); else mbar_wait(acc_empty + buf, accph ^ 1);
            tc_fence_after();
            int prev_seg = -1;
            for (int kc = 0; kc < p.num_k_chunks; ++kc) {
                if constexpr (PAIR) mbar_wait_cluster(a_ready + as, aph); else mbar_wait(a_ready + as, aph);
                // TF32: hi tiles and lo tiles are two arrays; FP16: stage = [hi 16 KB | lo 16 KB] contiguous
                const uint64_t ahi = ahi0 + (uint64_t)(as * ((F16 ? A_STAGE : A_TILE_BYTES) >> 4));
                const uint64_t alo = F16 ? ahi + (uint64_t)(A_TILE_BYTES >> 4)
                                         : alo0 + (uint64_t)(as * (A_TILE_BYTES >> 4));
                const int seg = p.ksplit > 1 ? kc * p.ksplit / p.num_k_chunks : 0;
                const uint32_t first = seg == prev_seg ? 1u : 0u;           // a segment's first chunk overwrites
                const uint32_t cfirst = kc > 0 ? 1u : 0u;                   // the correction accumulator spans all of K
                prev_seg = seg;
                for (int part = 0; part < 2; ++part) {
                    for (int nbk = 0; nbk < p.nb; ++nbk) {
                        mbar_wait(b_full + bs, bph);
                        tc_fence_after();
                        const uint64_t bd = b0 + (uint64_t)(bs * (B_TILE_BYTES >> 4));
                        const uint32_t t0 = tmem_base + buf * (uint32_t)p.acc_stride + (uint32_t)(nbk * BN);
                        const uint32_t d_tmem = t0 + seg_offset(seg, p.seg_w, p.seg_slot);
                        const uint32_t c_tmem = t0 + c_off;
                        if (leader) {
                            if (part == 0) {
                                // x_hi.w_hi into the main accumulator, x_lo.w_hi into the correction one
                                mma_issue<F16, PAIR>(d_tmem, ahi, bd, idesc, first);
#pragma unroll
                                for (int k = 1; k < BK / 8; ++k)      // 32-byte steps inside the swizzled row
                                    mma_issue<F16, PAIR>(d_tmem, ahi + 2 * k, bd + 2 * k, idesc, 1u);
                                mma_issue<F16, PAIR>(c_tmem, alo, bd, idesc, cfirst | corr_first);
#pragma unroll
                                for (int k = 1; k < BK / 8; ++k)
                                    mma_issue<F16, PAIR>(c_tmem, alo + 2 * k, bd + 2 * k, idesc, 1u);
                            } else {
#pragma unroll
                                for (int k = 0; k < BK / 8; ++k) mma_issue<F16, PAIR>(c_tmem, ahi + 2 * k, bd + 2 * k, idesc, 1u);
                            }
                            if constexpr (PAIR) tc_commit_2sm(b_empty + bs, pair_mask);
                            else if (w_mc) tc_commit_mc(b_empty + bs, w_mask);
                            else tc_commit(b_empty + bs);
                        }
                        __syncwarp();
                        if (++bs == B_STAGES) { bs = 0; bph ^= 1; }
                    }
                }
                if (leader) {
                    if constexpr (PAIR) tc_commit_2sm(a_empty + as, pair_mask);
                    else tc_commit(a_empty + as);
                }
                __syncwarp();
                if (++as == A_STAGES) { as = 0; aph ^= 1; }
            }
            if (leader) { if constexpr (PAIR) tc_commit_2sm(acc_full + buf, pair_mask); else tc_commit(acc_full + buf); }
            __syncwarp();
        }
    } else if (warp < 2 + kConvWarps) {
        // ===================== converters: raw fp32 -> (hi, lo) TF32 pair =====================
        const int t = threadIdx.x - 64;             // 0..kConvThreads-1
        const float sx = F16 ? p.x_scale[0] : 1.0f;
        // FP16: the four columns of a k-chunk this thread converts are the same for all of its quads
        // (quad index = t + 128 j, 16 quads per row), so their scales are one 128-bit load per chunk
        const float4* cs4 = reinterpret_cast<const float4*>(p.col_scale) + (t & 15);
        int as = 0;
        uint32_t aph = 0;
        for (int64_t grp = cluster_id; grp < n_groups; grp += n_clusters) {
            for (int kc = 0; kc < p.num_k_chunks; ++kc) {
                float4 cs = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
                if constexpr (F16) {
                    cs = __ldg(cs4 + kc * (BK16 / 4));
                    cs.x *= sx; cs.y *= sx; cs.z *= sx; cs.w *= sx;      // powers of two: exact
                }
                mbar_wait(x_full + as, aph);
                if (p.debug_skip & 1) {
                } else if constexpr (!F16) {
                    float4* hi = reinterpret_cast<float4*>(a_hi + as * A_TILE_BYTES);
                    float4* lo = reinterpret_cast<float4*>(a_lo + as * A_TILE_BYTES);
#pragma unroll
                    for (int j = 0; j < A_TILE_BYTES / 16 / kConvThreads; ++j) {
                        const int i = t + kConvThreads * j;
                        const float4 x = hi[i];
                        float4 h, l;
                        h.x = to_tf32(x.x); h.y = to_tf32(x.y); h.z = to_tf32(x.z); h.w = to_tf32(x.w);
                        l.x = to_tf32(x.x - h.x); l.y = to_tf32(x.y - h.y);
                        l.z = to_tf32(x.z - h.z); l.w = to_tf32(x.w - h.w);
                        hi[i] = h;
                        lo[i] = l;
                    }
                } else {
                    // raw tile: [128 rows][64 fp32], linear (256 B rows).  A thread takes 16 quads of 4 floats,
                    // quad index = thread + 128 j, so that a warp's 128-bit load covers 512 contiguous bytes
                    // (conflict-free; a lane stride of 32 bytes cost two wavefronts per quarter-warp, and
                    // shared-memory cycles are what this kernel is short of).  Once every thread holds its share
                    // (the tile is converted in place) it scales each quad into fp16 range, splits it into
                    // hi + lo halves and writes the two 8-byte results where the 128-byte-swizzled K-major fp16
                    // tiles want them: hi tile at the stage base, lo tile 16 KB above.
                    unsigned char* stage = a_hi + as * A_STAGE;
                    constexpr int QUADS = 2048 / kConvThreads;
                    float4 xr[QUADS];
#pragma unroll
                    for (int j = 0; j < QUADS; ++j)
                        xr[j] = *reinterpret_cast<const float4*>(stage + (t + kConvThreads * j) * 16);
                    asm volatile("bar.sync 2, %0;" ::"n"(kConvThreads) : "memory");      // all raw data is in registers
#pragma unroll
                    for (int j = 0; j < QUADS; ++j) {
                        const int qd = t + kConvThreads * j;
                        const int r = qd >> 4, q4 = qd & 15;
                        uint32_t h0, l0, h1, l1;
                        split_pair(xr[j].x, xr[j].y, cs.x, cs.y, h0, l0);
                        split_pair(xr[j].z, xr[j].w, cs.z, cs.w, h1, l1);
                        // SWIZZLE_128B: 16-byte piece ^ (row % 8); this quad is half a piece
                        const int off = r * 128 + (((q4 >> 1) ^ (r & 7)) << 4) + (q4 & 1) * 8;
                        *reinterpret_cast<uint2*>(stage + off) = make_uint2(h0, h1);
                        *reinterpret_cast<uint2*>(stage + A_TILE_BYTES + off) = make_uint2(l0, l1);
                    }
                }
                // generic-proxy writes -> visible to the tensor core.  (One arrival per thread: electing one lane per
                // warp behind a __syncwarp -- 4 arrivals per k-chunk instead of 128, remote ones in a pair's second
                // CTA -- measured 3-4 % slower on cfg 3 and on every other shape tried.)
                fence_proxy_async();
                if (lead) mbar_arrive(a_ready + as);
                else mbar_arrive_cluster(map_to_cta(smem_u32(a_ready + as), lead_rank));
                if (++as == A_STAGES) { as = 0; aph ^= 1; }
            }
        }
    } else if (warp < 2 + kConvWarps + 8) {
        // ===================== epilogue (8 warps) =====================
        // Warp w may read TMEM lanes 32*(w%4)..+31 only, so the eight warps form two groups of four:
        // group h covers half of this CTA's class columns for all 128 rows.  Two passes over TMEM:
        //   1. per 32-column chunk: logits (decision values out), chunk maximum, e = exp(z - chunk maximum) written
        //      back over the accumulator, running row maximum (first argmax) and rescaled sum
        //                                                      -> combine halves, then CTAs (DSMEM), once
        //   2. p = e * exp(chunk maximum - row maximum) / sum -> global memory, near-tie count
        // One exp per class, two reads and one write of TMEM, one exchange (the max / exp+sum / normalise form this
        // replaces read TMEM three times and exchanged twice; recomputing the exponentials in pass 2 instead of
        // keeping them saturated the special-function unit).  Each pass keeps the next TMEM load in flight.
        // Logit error model of this call: the truncating accumulation leaves ~kappa times the size of the
        // partial sums -- for a row that is the larger of its top logit and what this batch's rows and this
        // model's classes typically add up to on the way ([5]) -- plus what fp16's subnormal floor can cost ([6]).
        const float psum_scale = p.x_scale[5];
        const float floor_err = p.x_scale[6];
        const int ew = warp - (2 + kConvWarps);      // 0..7
        const int etid = ew * 32 + lane;             // 0..255
        // TMEM column of class `cb` of this CTA (see n_cols_cta): blocks of mma_n classes, BN columns apart
        const int mma_n = p.mma_n;
#define LRS_TCOL(cb) ((uint32_t)(mma_n == BN ? (cb) : ((cb) / mma_n) * BN + (cb) % mma_n))
        float* chunk_ref = reinterpret_cast<float*>(smem + SmemLayout::chunk_ref);   // [8 chunks][256 threads]
        const int q = warp & 3;                      // TMEM lane quadrant this warp may read
        const int h = ew >> 2;                       // column half
        const int r_in_tile = q * 32 + lane;
        const uint32_t t_lane0 = tmem_base + ((uint32_t)(q * 32) << 16);
        int n_real = p.n_classes - col0;             // real (unpadded) classes held by this CTA
        n_real = n_real < 0 ? 0 : (n_real > n_cols_cta ? n_cols_cta : n_real);
        const int n_chunks = (n_real + 31) / 32;
        const int cb_lo = h == 0 ? 0 : ((n_chunks + 1) / 2) * 32;
        const int cb_hi = h == 0 ? ((n_chunks + 1) / 2) * 32 : n_chunks * 32;
        // store tiles of the general form (rows whose pitch rules vector stores out): the host left the last 16 KB of
        // the ring area free for them (GemmParams::stage_off)
        float* const stg_general = p.stage_off ? reinterpret_cast<float*>(smem + p.stage_off) + ew * (kStoreTileBytes / 4) : nullptr;
        uint32_t it = 0;
        for (int64_t grp = cluster_id; grp < n_groups; grp += n_clusters, ++it) {
            const int64_t tile = grp * gsz + rt;      // may lie past the end: TMA zero-fills, nothing is stored
            const int64_t row = tile * BM + r_in_tile;
            const bool live = row < p.n_rows;
            const uint32_t par = it & 1;
            const uint32_t buf = p.acc_bufs == 2 ? (it & 1u) : 0u;
            const uint32_t t_lane = t_lane0 + buf * (uint32_t)p.acc_stride;
            // rows of this warp that exist, and where its block starts in the (n_rows x pitch) outputs
            const int64_t row_w = tile * BM + q * 32;
            const int rows_valid = p.n_rows - row_w >= 32 ? 32 : (p.n_rows > row_w ? (int)(p.n_rows - row_w) : 0);
            const int64_t blk_w = row_w * p.out_pitch + p.out_col0 + col0;
            mbar_wait(acc_full + buf, (p.acc_bufs == 2 ? (it >> 1) : it) & 1u);
            tc_fence_after();

            // ---- at most 64 classes: whole rows stay in registers (epilogue_in_registers) ----
            if (!PAIR && n_chunks <= 2 && p.ncc == 1 && !(p.debug_skip & 2)) {
                float* stg_r = reinterpret_cast<float*>(smem + SmemLayout::xchg) + ew * (kStoreTileBytes / 4);   // (no cluster exchange here)
                epilogue_in_registers<F16>(p, t_lane, h, q, lane, n_real, row, live, rows_valid, blk_w, cscale_s, bias_s,
                                           chunk_ref, stg_r, acc_empty + buf, psum_scale, floor_err);
                continue;
            }
            if constexpr (EPI2) {                     // 65..128 classes (launch_gemm_kernel picks this instantiation for them)
                if (n_chunks <= 4 && !(p.debug_skip & 6)) {
                    float* stg_r = reinterpret_cast<float*>(smem + SmemLayout::xchg) + ew * (kStoreTileBytes / 4);
                    epilogue_in_registers2<F16>(p, t_lane, h, q, lane, n_real, row, live, rows_valid, blk_w, cscale_s, bias_s,
                                                chunk_ref, stg_r, acc_empty + buf, psum_scale, floor_err);
                    continue;
                }
            }

            // ---- pass 1: running maximum (first argmax) and sum of exp(z - running maximum) over my columns.
            // The load of the next 32-column chunk is in flight while this one is worked on.
            const int cb_hi1 = (p.debug_skip & 2) ? cb_lo : cb_hi;
            float* const stg = stg_general;
            float m = -INFINITY, ssum = 0.0f;
            int arg = col0 + cb_lo;
            // raw accumulator chunk -> logits, in place
#define LRS_CHUNK_LOGITS(v, cb)                                                                        \
            do {                                                                                       \
                if (p.split) {                                                                         \
                    _Pragma("unroll") for (int i8 = 0; i8 < 32; i8 += 16)                              \
                        tmem_add16(t_lane + (uint32_t)(BN + (cb) + i8), (v) + i8);                     \
                    for (int sg = 1; sg < p.ksplit; ++sg)                                              \
                        _Pragma("unroll") for (int i8 = 0; i8 < 32; i8 += 16)                          \
                            tmem_add16(t_lane + seg_offset(sg, p.seg_w, p.seg_slot) + (uint32_t)((cb) + i8), (v) + i8);     \
                }                                                                                      \
                /* eight columns at a time, fenced: with all 64 scale / bias values of a chunk loaded */ \
                /* ahead of the FMAs the second (in-flight) TMEM buffer no longer fits in registers */    \
                _Pragma("unroll") for (int i8 = 0; i8 < 32; i8 += 8) {                                 \
                    _Pragma("unroll") for (int i = i8; i < i8 + 8; ++i) {                              \
                        if constexpr (F16) (v)[i] = fmaf((v)[i], cscale_s[(cb) + i], bias_s[(cb) + i]); \
                        else (v)[i] += bias_s[(cb) + i]; /* padding carries -inf */                    \
                    }                                                                                  \
                    asm volatile("" ::: "memory");                                                     \
                }                                                                                      \
            } while (0)
#define LRS_PASS1_CHUNK(v, cb)                                                                         \
            do {                                                                                       \
                LRS_CHUNK_LOGITS(v, cb);                                                               \
                if (p.decision && stg) {                                                               \
                    store_block_by_rows(stg, v, lane, p.decision + blk_w + (cb), p.out_pitch, rows_valid, \
                                        n_real - (cb) < 32 ? n_real - (cb) : 32);                      \
                } else if (p.decision && live) {                                                       \
                    const int64_t off = row * p.out_pitch + p.out_col0 + col0 + (cb);                  \
                    if ((cb) + 32 <= n_real && p.vec_store) {                                          \
                        _Pragma("unroll") for (int i = 0; i < 32; i += 4)                              \
                            __stcs(reinterpret_cast<float4*>(p.decision + off + i),                    \
                                   make_float4((v)[i], (v)[i + 1], (v)[i + 2], (v)[i + 3]));           \
                    } else {                                                                           \
                        _Pragma("unroll") for (int i = 0; i < 32; ++i)                                 \
                            if ((cb) + i < n_real) __stcs(p.decision + off + i, (v)[i]);               \
                    }                                                                                  \
                }                                                                                      \
                /* four independent chains for the maximum (and for the sum below): one 32-long chain of */ \
                /* dependent operations per chunk is latency this warp has nobody to hide behind         */ \
                float cm0 = -INFINITY, cm1 = -INFINITY, cm2 = -INFINITY, cm3 = -INFINITY;              \
                _Pragma("unroll") for (int i = 0; i < 32; i += 4) {                                    \
                    cm0 = fmaxf(cm0, (v)[i]); cm1 = fmaxf(cm1, (v)[i + 1]);                            \
                    cm2 = fmaxf(cm2, (v)[i + 2]); cm3 = fmaxf(cm3, (v)[i + 3]);                        \
                }                                                                                      \
                const float cm = fmaxf(fmaxf(cm0, cm1), fmaxf(cm2, cm3));                              \
                if (cm > m) { /* strict: an earlier chunk keeps a tie */                               \
                    int idx = 31;                                                                      \
                    _Pragma("unroll") for (int i = 30; i >= 0; --i) idx = ((v)[i] == cm) ? i : idx;    \
                    arg = col0 + (cb) + idx; /* first index wins */                                    \
                }                                                                                      \
                /* exponentials relative to the chunk's own maximum (<= 1), kept in TMEM for pass 2 */ \
                /* (z - cm) first: exactly 0 for the maximum itself, whatever its size (an fma against a */ \
                /* pre-multiplied cm * log2e leaves the product's rounding error in the exponent)        */ \
                float cs0 = 0.0f, cs1 = 0.0f, cs2 = 0.0f, cs3 = 0.0f;                                  \
                _Pragma("unroll") for (int i = 0; i < 32; i += 4) {                                    \
                    (v)[i] = ex2(((v)[i] - cm) * kLog2e); cs0 += (v)[i];                               \
                    (v)[i + 1] = ex2(((v)[i + 1] - cm) * kLog2e); cs1 += (v)[i + 1];                   \
                    (v)[i + 2] = ex2(((v)[i + 2] - cm) * kLog2e); cs2 += (v)[i + 2];                   \
                    (v)[i + 3] = ex2(((v)[i + 3] - cm) * kLog2e); cs3 += (v)[i + 3];                   \
                }                                                                                      \
                const float cs = (cs0 + cs1) + (cs2 + cs3);                                            \
                tmem_st32(t_lane + LRS_TCOL(cb), v);                                                   \
                chunk_ref[(((cb) - cb_lo) >> 5) * kEpiThreads + etid] = cm;                            \
                if (cm > -INFINITY) { /* an all-padding chunk: -inf - -inf */                          \
                    const float nm = fmaxf(m, cm);                                                     \
                    ssum = ssum * ex2((m - nm) * kLog2e) + cs * ex2((cm - nm) * kLog2e);               \
                    m = nm;                                                                            \
                }                                                                                      \
            } while (0)
            if (cb_lo < cb_hi1) {
                // two register buffers, the next chunk's load always in flight (past the end: the last chunk
                // again, unused -- a load under a condition made the compiler keep copies of both buffers)
                float ra[32], rb[32];
                const int cb_last = cb_hi1 - 32;
                tmem_ld32_issue(t_lane + LRS_TCOL(cb_lo), ra);
                for (int cb = cb_lo; cb < cb_hi1; cb += 64) {
                    tmem_ld_wait(ra);
                    tmem_ld32_issue(t_lane + LRS_TCOL(min(cb + 32, cb_last)), rb);
                    LRS_PASS1_CHUNK(ra, cb);
                    tmem_ld_wait(rb);
                    tmem_ld32_issue(t_lane + LRS_TCOL(min(cb + 64, cb_last)), ra);
                    if (cb + 32 < cb_hi1) LRS_PASS1_CHUNK(rb, cb + 32);
                }
                tmem_ld_wait(ra);
                tmem_wait_st();
            }
            // combine the two column halves of this CTA, then the class slices of the cluster: one exchange of
            // (maximum, its class, sum of exponentials relative to that maximum) per row
            float gm, gs;
            int garg;
            {
                if (h == 1) halves4[r_in_tile] = make_float4(m, __int_as_float(arg), ssum, 0.0f);
                epi_bar_sync();
                if (h == 0) {
                    const float4 o = halves4[r_in_tile];
                    const float nm = fmaxf(m, o.x);
                    float ns = 0.0f;
                    if (m > -INFINITY) ns += ssum * ex2((m - nm) * kLog2e);
                    if (o.x > -INFINITY) ns += o.z * ex2((o.x - nm) * kLog2e);
                    if (o.x > m) arg = __float_as_int(o.y);                       // lower columns keep a tie
                    m = nm;
                    ssum = ns;
                    gm = m;
                    garg = arg;
                    gs = ssum;
                    if (p.ncc > 1) {
                        float4* buf = xchg4 + (size_t)(par * kMaxCluster) * BM;
                        const uint32_t slot = smem_u32(buf + rank * BM + r_in_tile);
                        const uint32_t bar = smem_u32(xchg_bar + par);
                        for (uint32_t dst = 0; dst < (uint32_t)p.ncc; ++dst) {
                            st_cluster_v4(map_to_cta(slot, PAIR ? 2 * dst + rt : rt * p.ncc + dst), m, __int_as_float(arg),
                                          ssum, 0.0f);
                            mbar_arrive_cluster(map_to_cta(bar, PAIR ? 2 * dst + rt : rt * p.ncc + dst));
                        }
                        mbar_wait_cluster(xchg_bar + par, (it >> 1) & 1);
                        gm = -INFINITY;
                        for (int src = 0; src < p.ncc; ++src) {
                            const float4 e = buf[src * BM + r_in_tile];
                            if (e.x > gm) { gm = e.x; garg = __float_as_int(e.y); }  // lower slice keeps a tie
                        }
                        gs = 0.0f;
                        for (int src = 0; src < p.ncc; ++src) {      // fixed order: every CTA gets the same sum
                            const float4 e = buf[src * BM + r_in_tile];
                            if (e.x > -INFINITY) gs += e.z * ex2((e.x - gm) * kLog2e);
                        }
                    }
                    glob4[r_in_tile] = make_float4(gm, __int_as_float(garg), gs, 0.0f);
                }
                epi_bar_sync();
                const float4 gsv = glob4[r_in_tile];
                gm = gsv.x;
                garg = __float_as_int(gsv.y);
                gs = gsv.z;
            }
            const float inv = __frcp_rn(gs);

            // ---- pass 2: p = e * exp(chunk max - max) / sum -> global memory; near-tie count.
            // A runner-up this close to the maximum may be the true winner, and |dp| ~ p(1-p)|dz|: the row is
            // flagged when a class of middling probability could move by ~8e-6 (see `ez` above the loop).
            const float ez = fmaxf(p.kappa * fmaxf(psum_scale, fabsf(gm)) + floor_err, 1e-30f);
            const float risk_thr = p.need_prob ? 8e-6f / ez : 1.0f;      // (p (1 - p) <= 1/4: never)
            const float near_thr = inv * ex2(-2.0f * ez * kLog2e);   // z > max - 2 ez, as a probability
            int near = 0;
            const bool want_rows = p.proba != nullptr;
#define LRS_PASS2_CHUNK(v, cb)                                                                         \
            do {                                                                                       \
                const float fj = ex2((chunk_ref[(((cb) - cb_lo) >> 5) * kEpiThreads + etid] - gm) * kLog2e) * inv; \
                _Pragma("unroll") for (int i = 0; i < 32; ++i) {                                       \
                    (v)[i] *= fj;                                                                      \
                    near += (v)[i] > near_thr ? 1 : 0;                                                 \
                }                                                                                      \
                if (p.proba && stg) {                                                                  \
                    store_block_by_rows(stg, v, lane, p.proba + blk_w + (cb), p.out_pitch, rows_valid, \
                                        n_real - (cb) < 32 ? n_real - (cb) : 32);                      \
                } else if (p.proba && live) {                                                          \
                    const int64_t off = row * p.out_pitch + p.out_col0 + col0 + (cb);                  \
                    if ((cb) + 32 <= n_real && p.vec_store == 2) {                                     \
                        _Pragma("unroll") for (int i = 0; i < 32; i += 8) st_cs_v8(p.proba + off + i, (v) + i); \
                    } else if ((cb) + 32 <= n_real && p.vec_store) {                                   \
                        _Pragma("unroll") for (int i = 0; i < 32; i += 4)                              \
                            __stcs(reinterpret_cast<float4*>(p.proba + off + i),                       \
                                   make_float4((v)[i], (v)[i + 1], (v)[i + 2], (v)[i + 3]));           \
                    } else {                                                                           \
                        _Pragma("unroll") for (int i = 0; i < 32; ++i)                                 \
                            if ((cb) + i < n_real) __stcs(p.proba + off + i, (v)[i]);                  \
                    }                                                                                  \
                }                                                                                      \
            } while (0)
            if ((want_rows || p.flags) && cb_lo < cb_hi1) {
                // two register buffers, the next chunk's load always in flight (past the end: the last chunk
                // again, unused -- a load under a condition made the compiler keep copies of both buffers)
                float ra[32], rb[32];
                const int cb_last = cb_hi1 - 32;
                tmem_ld32_issue(t_lane + LRS_TCOL(cb_lo), ra);
                for (int cb = cb_lo; cb < cb_hi1; cb += 64) {
                    tmem_ld_wait(ra);
                    tmem_ld32_issue(t_lane + LRS_TCOL(min(cb + 32, cb_last)), rb);
                    LRS_PASS2_CHUNK(ra, cb);
                    tmem_ld_wait(rb);
                    tmem_ld32_issue(t_lane + LRS_TCOL(min(cb + 64, cb_last)), ra);
                    if (cb + 32 < cb_hi1) LRS_PASS2_CHUNK(rb, cb + 32);
                }
                tmem_ld_wait(ra);
            }
#undef LRS_PASS2_CHUNK
#undef LRS_PASS1_CHUNK
#undef LRS_CHUNK_LOGITS
#undef LRS_TCOL
            tc_fence_before();
            if (lead) mbar_arrive(acc_empty + buf);   // this buffer may be overwritten
            else mbar_arrive_cluster(map_to_cta(smem_u32(acc_empty + buf), lead_rank));

            if (live) {
                // the thread whose columns hold the maximum always counts it; anything else above tau is a near-tie
                const bool owner = garg >= col0 + cb_lo && garg < col0 + cb_hi;
                if (p.flags && near >= (owner ? 2 : 1)) flag_row(p.flags, p.flag_list, p.n_flagged, p.ezr, row, ez);
                if (rank == 0 && h == 0) {
                    if (p.label) __stcs(p.label + row, garg);
                    if (p.maxprob) __stcs(p.maxprob + row, inv);      // exp(gm - gm) == 1
                    const bool finite_row = finite_f(gm) && finite_f(gs);
                    if (!finite_row) {
                        // With a re-scoring pass the row is handed over (flag 2): an fp16 overflow of a
                        // finite input would otherwise be reported as non-finite data.
                        if (p.flags && !p.count_nonfinite) flag_row(p.flags, p.flag_list, p.n_flagged, p.ezr, row, ez);
                        else atomicAdd(p.status, 1ull);
                    }
                    if (p.lse) p.lse[row] = gm + logf(gs);    // non-finite rows store a non-finite value here
                    // |dp| ~ p (1 - p) |dz| is largest for the class nearest to one half, which is always the top
                    // one (p (1 - p) rises up to 1/2 and no class but the top one can pass it): p_top = inv
                    if (p.flags && finite_row && inv * (1.0f - inv) > risk_thr)
                        flag_row(p.flags, p.flag_list, p.n_flagged, p.ezr, row, ez);
                }
            }
        }
    }

    // ---- teardown ----
    tc_fence_before();
    __syncthreads();
    if (csize > 1) cluster_sync_all();               // no CTA leaves while a peer may still write its smem
    if (warp == 1) {
        if constexpr (PAIR)
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
        else
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

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

// ---- short rows, 13..32 classes -------------------------------------------------------------------
// Rows of at most 64 floats against at most 32 classes (covertype: d=54, C=7..20; letters: d=16, C=26): one
// 128-row tile is ONE k-chunk and one tiny MMA group, so the kernel above -- a W ring, an X ring converted in
// place, one mbarrier round trip per tile and stage -- spends its time on hand-overs (d=16, C=26: 1.1 TB/s),
// and the fp32 register-tile kernel is issue-bound (27 TFLOP/s: 1.7-1.9 TB/s at 20 classes).  Here W (8 KB as
// fp16 hi + lo) stays in shared memory for the whole launch, dense rows arrive as plain 1-D bulk copies of whole
// tiles (any d: no 16-byte row pitch needed), several raw tiles, three operand tiles and eight 32-column
// accumulators are in flight, and each stage is its own set of warps:
//   warp 0       bulk-copy producer (raw fp32 tiles, ring of up to 8)
//   warps 1-4    converters: raw tile -> column balance, scale, fp16 hi / lo halves, K-major 128B-swizzled tiles
//   warp 5       MMA issuer: 3 x (dk / 16) MMAs of M=128, N=32 per tile into one of 8 TMEM buffers
//   warps 6-13   epilogue, two sets of four (TMEM lane quadrants) on alternate tiles: one TMEM read, softmax in
//                registers, outputs, re-scoring flags -- the single-pass form of the kernel above
// Same arithmetic (3xFP16 split, truncating TMEM accumulation) and the same error band as that kernel; flagged
// rows are re-scored whole in fp32 (32 classes x 64 features is nothing) by short_fixup_kernel.
// Sixteen worker warps in sets of four, split at launch between the two long stages: rows of up to 32 floats
// convert quickly and want three epilogue sets (1 + 3); longer rows want two converter sets (2 + 2).  Sets of a
// stage take the tiles in turn.
constexpr int kSrWorkWarps = 16;
constexpr int kSrMaxEpiWarps = 12;
constexpr int kSrThreads = (2 + kSrWorkWarps) * 32;                        // 576
constexpr int kSrMaxOpStages = 4;                        // operand tiles in flight: 2 or 4 (runtime: shared memory)
constexpr int kSrAccBufs = 8;
constexpr int kSrMaxRaw = 8;
constexpr int kSrOpBytes = 2 * A_TILE_BYTES;             // hi | lo, 128-byte rows
struct SrSmem {
    static constexpr int w = 0;                                          // W hi 4 KB | lo 4 KB
    static constexpr int cs = w + 2 * 32 * 128;                          // [64] per-feature scale x call scale
    static constexpr int bias = cs + 64 * 4;                             // [32]
    static constexpr int cscale = bias + 32 * 4;                         // [32]
    static constexpr int out = cscale + 32 * 4;                          // [epilogue warps][32 rows x 32] floats
    static constexpr int bars = out + kSrMaxEpiWarps * 32 * 32 * 4;      // raw_full/empty[8], op_full/empty[4], acc_full/empty[8], w_full
    static constexpr int n_bars = 2 * kSrMaxRaw + 2 * kSrMaxOpStages + 2 * kSrAccBufs + 1;
    static constexpr int tmem_slot = bars + n_bars * 8;
    static constexpr int op = (tmem_slot + 16 + 1023) / 1024 * 1024;     // [op_stages][32 KB], then [raw_stages][stage_bytes]
};
struct ShortParams {
    int need_prob;          // see GemmParams::need_prob
    const float* X;
    int64_t n_rows;
    int d;                      // features (dk = features rounded up to 16 is the kernel's template argument)
    int n_classes;
    int raw_stages, stage_bytes;
    int op_stages;              // 2 or 4 operand tiles per CTA
    int conv_sets;              // 1 or 2 sets of four converter warps; the other 4 - conv_sets sets are epilogue warps
    const float* bias;
    const float* col_inv_scale;
    const float* x_scale;
    const float* col_scale;
    float kappa;
    int count_nonfinite, vec_store;
    float* proba;
    float* decision;
    int32_t* label;
    float* maxprob;
    unsigned long long* status;
    unsigned* flags;
    unsigned* flag_list;
    unsigned* n_flagged;
    float* ezr;                 // unused here (whole rows are re-scored): nullptr
};

// A warp's 32 rows x C values, one row per lane in registers, to their place in an [n][C] matrix: staged in
// shared memory exactly as they will lie in global memory (row pitch C) and written out as the one contiguous
// span they are, 512 bytes per instruction.  (Each lane storing its own row put 32 different cache lines
// under every store instruction.)  With C a multiple of 4 the rows go in as 16-byte pieces whose position is
// XOR-swizzled by the piece's 128-byte line, so that neither the row-wise writes nor the span-wise reads
// collide on banks; other C are written word by word (at most 2-way conflicts).
__device__ __forceinline__ void store_rows_coalesced(float* __restrict__ stage, const float (&v)[32], int C, int lane,
                                                     float* __restrict__ dst /* first row of the warp */, int rows_valid,
                                                     bool aligned16) {
    const bool quads = (C & 3) == 0;
    if (quads) {
        const int u0 = lane * (C >> 2);
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (4 * j < C) {
                const int u = u0 + j;
                *reinterpret_cast<float4*>(stage + 4 * (u ^ ((u >> 3) & 7))) =
                    make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            }
    } else {
#pragma unroll
        for (int i = 0; i < 32; ++i)
            if (i < C) stage[lane * C + i] = v[i];
    }
    __syncwarp();
    const int total = rows_valid * C;
    if (aligned16 && rows_valid == 32) {
        for (int e = lane * 4; e < total; e += 128) {
            const int u = e >> 2;
            const float4 o = *reinterpret_cast<const float4*>(stage + 4 * (quads ? u ^ ((u >> 3) & 7) : u));
            __stcs(reinterpret_cast<float4*>(dst + e), o);
        }
    } else {
        for (int e = lane; e < total; e += 32) {
            const int u = e >> 2;
            __stcs(dst + e, stage[quads ? 4 * (u ^ ((u >> 3) & 7)) + (e & 3) : e]);
        }
    }
    __syncwarp();
}

// DK = features rounded up to 16 (the K extent of a tile); VEC: d % 4 == 0, rows of a raw tile are 16-byte aligned
template <int DK, bool VEC>
__global__ void __launch_bounds__(kSrThreads, 1)
short_rows_kernel(const __grid_constant__ CUtensorMap map_whi, const __grid_constant__ CUtensorMap map_wlo,
                  const ShortParams p) {
    if (reinterpret_cast<const uint32_t*>(p.x_scale)[4] != 0u) return;        // this call was given to the fp32 fallback
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    float* cs_s = reinterpret_cast<float*>(smem + SrSmem::cs);
    float* bias_s = reinterpret_cast<float*>(smem + SrSmem::bias);
    float* cscale_s = reinterpret_cast<float*>(smem + SrSmem::cscale);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SrSmem::bars);
    uint64_t* raw_full = bars;
    uint64_t* raw_empty = raw_full + kSrMaxRaw;
    uint64_t* op_full = raw_empty + kSrMaxRaw;
    uint64_t* op_empty = op_full + kSrMaxOpStages;
    uint64_t* acc_full = op_empty + kSrMaxOpStages;
    uint64_t* acc_empty = acc_full + kSrAccBufs;
    uint64_t* w_full = acc_empty + kSrAccBufs;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + SrSmem::tmem_slot);
    unsigned char* raw = smem + SrSmem::op + p.op_stages * kSrOpBytes;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int d = p.d;
    const int kMmaWarp = 1 + 4 * p.conv_sets;
    const int epi_sets = kSrWorkWarps / 4 - p.conv_sets;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kSrMaxRaw; ++i) { mbar_init(raw_full + i, 1); mbar_init(raw_empty + i, 128); }
        for (int i = 0; i < kSrMaxOpStages; ++i) { mbar_init(op_full + i, 128); mbar_init(op_empty + i, 1); }
        for (int i = 0; i < kSrAccBufs; ++i) { mbar_init(acc_full + i, 1); mbar_init(acc_empty + i, 128); }
        mbar_init(w_full, 1);
        mbar_fence_init();
    }
    if (threadIdx.x < 64) cs_s[threadIdx.x] = threadIdx.x < d ? p.col_scale[threadIdx.x] * p.x_scale[0] : 0.0f;
    if (threadIdx.x < 32) {
        bias_s[threadIdx.x] = p.bias[threadIdx.x];
        cscale_s[threadIdx.x] = p.col_inv_scale[threadIdx.x] * p.x_scale[1];
    }
    if (warp == kMmaWarp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(256)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int64_t n_tiles = (p.n_rows + BM - 1) / BM;
    const int64_t tile_floats = (int64_t)BM * d;

    if (warp == 0) {
        // ===================== producer: W once, then raw tiles =====================
        const bool leader = elect_one();
        if (leader) {
            mbar_expect_tx(w_full, 2 * 32 * 128);
            tma_load_2d(smem + SrSmem::w, &map_whi, 0, 0, w_full);
            tma_load_2d(smem + SrSmem::w + 32 * 128, &map_wlo, 0, 0, w_full);
        }
        int st = 0;
        uint32_t ph = 0;
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            mbar_wait(raw_empty + st, ph ^ 1);
            if (leader) {
                const int64_t left = (p.n_rows - tile * BM) * d;                 // floats of X from this tile on
                // whole 16-byte pieces only; the <= 3 floats a last, ragged tile may have beyond them are read
                // from global memory by the converters
                const uint32_t bytes = (uint32_t)((left < tile_floats ? left : tile_floats) * 4) & ~15u;
                if (bytes) {
                    mbar_expect_tx(raw_full + st, bytes);
                    bulk_g2s(raw + (size_t)st * p.stage_bytes, p.X + tile * tile_floats, bytes, raw_full + st);
                } else {
                    mbar_arrive(raw_full + st);
                }
            }
            __syncwarp();
            if (++st == p.raw_stages) { st = 0; ph ^= 1; }
        }
    } else if (warp < kMmaWarp) {
        // ===================== converters: conv_sets sets of four warps take the tiles in turn =====================
        const int set = (warp - 1) >> 2;
        const int t = (threadIdx.x - 32) & 127;                  // 0..127 inside the set
        constexpr int QPR = DK / 4;                              // 4-float pieces per row
        constexpr int ITEMS = QPR;                               // per thread: 128 rows x QPR pieces / 128 threads
        int st = 0, os = 0;
        uint32_t ph = 0, oph = 0;
        int turn = 0;
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            if (turn == set) {
                mbar_wait(raw_full + st, ph);
                mbar_wait(op_empty + os, oph ^ 1);
                const float* src = reinterpret_cast<const float*>(raw + (size_t)st * p.stage_bytes);
                unsigned char* dst = smem + SrSmem::op + os * kSrOpBytes;
                const int64_t left = (p.n_rows - tile * BM) * d;
                const bool ragged = left < tile_floats;
                float x[ITEMS][4];
                if (!ragged) {
                    // all the loads of the tile first, then the arithmetic
#pragma unroll
                    for (int j = 0; j < ITEMS; ++j) {
                        const int it = t + 128 * j;
                        const int r = it / QPR, q = it % QPR;
                        const int e0 = r * d + 4 * q;
                        if constexpr (VEC) {
                            const float4 v = 4 * q < d ? *reinterpret_cast<const float4*>(src + e0) : make_float4(0.f, 0.f, 0.f, 0.f);
                            x[j][0] = v.x; x[j][1] = v.y; x[j][2] = v.z; x[j][3] = v.w;
                        } else {
#pragma unroll
                            for (int i = 0; i < 4; ++i) x[j][i] = 4 * q + i < d ? src[e0 + i] : 0.0f;
                        }
                    }
                } else {
                    const int n_in_smem = (int)(((left * 4) & ~15ll) >> 2), n_valid = (int)left;
                    const float* gsrc = p.X + tile * tile_floats;
#pragma unroll
                    for (int j = 0; j < ITEMS; ++j) {
                        const int it = t + 128 * j;
                        const int r = it / QPR, q = it % QPR;
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int e = r * d + 4 * q + i;
                            x[j][i] = (4 * q + i < d && e < n_valid) ? (e < n_in_smem ? src[e] : __ldg(gsrc + e)) : 0.0f;
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < ITEMS; ++j) {
                    const int it = t + 128 * j;
                    const int r = it / QPR, q = it % QPR;
                    const float4 cs = *reinterpret_cast<const float4*>(cs_s + 4 * q);
                    uint32_t h0, l0, h1, l1;
                    split_pair(x[j][0], x[j][1], cs.x, cs.y, h0, l0);
                    split_pair(x[j][2], x[j][3], cs.z, cs.w, h1, l1);
                    const int off = r * 128 + (((q >> 1) ^ (r & 7)) << 4) + (q & 1) * 8;     // SWIZZLE_128B
                    *reinterpret_cast<uint2*>(dst + off) = make_uint2(h0, h1);
                    *reinterpret_cast<uint2*>(dst + A_TILE_BYTES + off) = make_uint2(l0, l1);
                }
                fence_proxy_async();
                mbar_arrive(op_full + os);
                mbar_arrive(raw_empty + st);
            }
            if (++turn == p.conv_sets) turn = 0;
            if (++st == p.raw_stages) { st = 0; ph ^= 1; }
            if (++os == p.op_stages) { os = 0; oph ^= 1; }
        }
    } else if (warp == kMmaWarp) {
        // ===================== MMA issuer =====================
        const bool leader = elect_one();
        const uint64_t desc0 = make_smem_desc(0);
        const uint64_t a0 = desc0 | (uint64_t)((smem_u32(smem + SrSmem::op) & 0x3FFFF) >> 4);
        const uint64_t whi = desc0 | (uint64_t)((smem_u32(smem + SrSmem::w) & 0x3FFFF) >> 4);
        const uint64_t wlo = whi + (uint64_t)((32 * 128) >> 4);
        const uint32_t idesc = (kIdesc16 & ~(0x3Fu << 17)) | ((uint32_t)(32 >> 3) << 17);
        mbar_wait(w_full, 0);
        int ab = 0, os = 0;
        uint32_t aph = 0, oph = 0;
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            mbar_wait(op_full + os, oph);
            mbar_wait(acc_empty + ab, aph ^ 1);
            tc_fence_after();
            if (leader) {
                const uint64_t ahi = a0 + (uint64_t)((os * kSrOpBytes) >> 4);
                const uint64_t alo = ahi + (uint64_t)(A_TILE_BYTES >> 4);
                const uint32_t d_tmem = tmem_base + (uint32_t)(ab * 32);
#pragma unroll
                for (int k = 0; k < DK / 16; ++k) {
                    tc_mma_f16(d_tmem, ahi + 2 * k, whi + 2 * k, idesc, k > 0 ? 1u : 0u);
                    tc_mma_f16(d_tmem, alo + 2 * k, whi + 2 * k, idesc, 1u);
                    tc_mma_f16(d_tmem, ahi + 2 * k, wlo + 2 * k, idesc, 1u);
                }
                tc_commit(op_empty + os);
                tc_commit(acc_full + ab);
            }
            __syncwarp();
            if (++ab == kSrAccBufs) { ab = 0; aph ^= 1; }
            if (++os == p.op_stages) { os = 0; oph ^= 1; }
        }
    } else {
        // ===================== epilogue: epi_sets sets of four warps take the tiles in turn =====================
        const int ew = warp - (kMmaWarp + 1);                    // 0..11
        const int set = ew >> 2;
        const int q = warp & 3;                                  // TMEM lane quadrant this warp may read
        float* stage = reinterpret_cast<float*>(smem + SrSmem::out) + ew * 32 * 32;
        const float psum_scale = p.x_scale[5];
        const float floor_err = p.x_scale[6];
        const int C = p.n_classes;
        const bool p_al = (reinterpret_cast<uintptr_t>(p.proba) & 15u) == 0, z_al = (reinterpret_cast<uintptr_t>(p.decision) & 15u) == 0;
        int64_t seq = 0;
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++seq) {
            if ((int)(seq % epi_sets) != set) continue;
            const int ab = (int)(seq % kSrAccBufs);
            mbar_wait(acc_full + ab, (uint32_t)((seq / kSrAccBufs) & 1));
            tc_fence_after();
            float v[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * 32), v);
            tc_fence_before();
            mbar_arrive(acc_empty + ab);
            const int64_t row0 = tile * BM + q * 32;             // first row of this warp
            const int64_t row = row0 + lane;
            const bool live = row < p.n_rows;
            const int rows_valid = (int)(p.n_rows - row0 >= 32 ? 32 : (p.n_rows - row0 > 0 ? p.n_rows - row0 : 0));
            float gm = -INFINITY;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                v[i] = fmaf(v[i], cscale_s[i], bias_s[i]);       // padding carries -inf
                gm = fmaxf(gm, v[i]);
            }
            int garg = 31;
#pragma unroll
            for (int i = 30; i >= 0; --i) garg = (v[i] == gm) ? i : garg;   // first index wins
            if (p.decision) store_rows_coalesced(stage, v, C, lane, p.decision + row0 * C, rows_valid, z_al);
            const float ez = fmaxf(p.kappa * fmaxf(psum_scale, fabsf(gm)) + floor_err, 1e-30f);
            const float tau = gm - 2.0f * ez;
            float gs = 0.0f;
            int gnear = 0;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                gnear += v[i] > tau ? 1 : 0;
                v[i] = i < C ? ex2((v[i] - gm) * kLog2e) : 0.0f;
                gs += v[i];
            }
            const float inv = __frcp_rn(gs);
            // |dp| ~ p (1 - p) |dz| is largest for the class nearest to one half, and that is always the top one
            // (p (1 - p) rises up to 1/2, and no class but the top one can pass it): its probability is `inv`
            const bool risk = p.need_prob && inv * (1.0f - inv) > 8e-6f / ez;
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] *= inv;
            if (p.proba) store_rows_coalesced(stage, v, C, lane, p.proba + row0 * C, rows_valid, p_al);
            if (live) {
                if (p.label) __stcs(p.label + row, garg);
                if (p.maxprob) __stcs(p.maxprob + row, inv);
                const bool finite_row = finite_f(gm) && finite_f(gs);
                if (!finite_row) {
                    if (p.flags && !p.count_nonfinite) flag_row(p.flags, p.flag_list, p.n_flagged, p.ezr, row, ez);
                    else atomicAdd(p.status, 1ull);
                } else if (p.flags && (risk || gnear >= 2)) {
                    flag_row(p.flags, p.flag_list, p.n_flagged, p.ezr, row, ez);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256) : "memory");
}

// Flagged rows of the kernel above, re-scored whole in fp32: one warp per row, lane c owns class c; W (at most
// 32 x 64 floats) sits transposed in shared memory (read from global per row, the 64 dependent trips to L2 cost
// 35 us a row).
__global__ void __launch_bounds__(256)
short_fixup_kernel(const float* __restrict__ X, int d, int C, const float* __restrict__ W, const float* __restrict__ b,
                   const unsigned* __restrict__ list, const unsigned* __restrict__ n_list, float* proba, float* decision,
                   int32_t* label, float* maxprob, unsigned long long* rescored, unsigned long long* status) {
    __shared__ float Ws[64 * 33];
    const unsigned n = *n_list;
    if (blockIdx.x * 8 >= n) return;
    for (int i = threadIdx.x; i < C * d; i += 256) {
        const int c = i / d, k = i - c * d;
        Ws[k * 33 + c] = __ldg(W + i);
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cl = lane < C ? lane : 0;
    for (unsigned li = blockIdx.x * 8 + warp; li < n; li += gridDim.x * 8) {
        const int64_t row = list[li];
        const float* x = X + row * d;
        const float x0 = lane < d ? __ldg(x + lane) : 0.0f, x1 = lane + 32 < d ? __ldg(x + lane + 32) : 0.0f;
        float z = 0.0f;
        for (int k = 0; k < d; ++k) {
            const float xk = __shfl_sync(0xffffffffu, k < 32 ? x0 : x1, k & 31);
            z = fmaf(xk, Ws[k * 33 + cl], z);
        }
        z = lane < C ? z + __ldg(b + lane) : -CUDART_INF_F;
        float m = z;
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
        const unsigned win = __ballot_sync(0xffffffffu, z == m);
        const float e = lane < C ? expf(z - m) : 0.0f;
        float ssum = e;
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) ssum += __shfl_xor_sync(0xffffffffu, ssum, off);
        const float inv = __frcp_rn(ssum);
        if (lane < C) {
            if (proba) proba[row * C + lane] = e * inv;
            if (decision) decision[row * C + lane] = z;
        }
        if (lane == 0) {
            if (label) label[row] = __ffs(win) - 1;                     // first index wins
            if (maxprob) maxprob[row] = inv;
            atomicAdd(rescored, 1ull);
            if (!(finite_f(m) && finite_f(ssum))) atomicAdd(status, 1ull);
        }
    }
}

// What scores a short-row call whose dynamic range ruled out 3xFP16: plain fp32, one warp per row (rare; the
// batch is counted in status[2] like the 3xTF32 fallback of the kernel above).
__global__ void __launch_bounds__(256)
short_fallback_kernel(const float* __restrict__ X, int64_t n_rows, int d, int C, const float* __restrict__ W,
                      const float* __restrict__ b, const float* __restrict__ x_scale, float* proba, float* decision,
                      int32_t* label, float* maxprob, unsigned long long* status) {
    if (reinterpret_cast<const uint32_t*>(x_scale)[4] != 1u) return;
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(status + 2, (unsigned long long)n_rows);
    const int lane = threadIdx.x & 31;
    for (int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); row < n_rows; row += (int64_t)gridDim.x * 8)
        score_row_warp(X + row * d, row, lane, d, C, W, b, proba, decision, label, maxprob, status);
}

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
    p.conv_sets = d <= 32 ? 1 : 2;
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
