// Tile constants, shared-memory plan, kernel parameters and the PTX wrappers (mbarrier, TMA, tcgen05, TMEM)
// of the tensor-core scoring path; the store tile and segment helpers its epilogues share.
// One translation unit with gemm_tcgen05.cu (included there, in this order); not a stand-alone header.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>

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
// Packed fp32 arithmetic (two lanes of a 64-bit register pair per instruction; same IEEE results as the scalar forms):
// an epilogue that is bound by its instruction count does its multiplies and adds two columns at a time.
__device__ __forceinline__ void fma2(float& a0, float& a1, float b0, float b1, float c0, float c1) {   // a = a * b + c
    asm("{\n\t.reg .b64 a, b, c;\n\tmov.b64 a, {%0, %1};\n\tmov.b64 b, {%2, %3};\n\tmov.b64 c, {%4, %5};\n\t"
        "fma.rn.f32x2 a, a, b, c;\n\tmov.b64 {%0, %1}, a;\n\t}"
        : "+f"(a0), "+f"(a1) : "f"(b0), "f"(b1), "f"(c0), "f"(c1));
}
__device__ __forceinline__ void mul2(float& a0, float& a1, float b0, float b1) {                        // a = a * b
    asm("{\n\t.reg .b64 a, b;\n\tmov.b64 a, {%0, %1};\n\tmov.b64 b, {%2, %3};\n\tmul.rn.f32x2 a, a, b;\n\tmov.b64 {%0, %1}, a;\n\t}"
        : "+f"(a0), "+f"(a1) : "f"(b0), "f"(b1));
}
__device__ __forceinline__ void sub2(float& a0, float& a1, float b0, float b1) {                        // a = a - b
    asm("{\n\t.reg .b64 a, b;\n\tmov.b64 a, {%0, %1};\n\tmov.b64 b, {%2, %3};\n\tsub.rn.f32x2 a, a, b;\n\tmov.b64 {%0, %1}, a;\n\t}"
        : "+f"(a0), "+f"(a1) : "f"(b0), "f"(b1));
}
__device__ __forceinline__ void add2(float& a0, float& a1, float b0, float b1) {                        // a = a + b
    asm("{\n\t.reg .b64 a, b;\n\tmov.b64 a, {%0, %1};\n\tmov.b64 b, {%2, %3};\n\tadd.rn.f32x2 a, a, b;\n\tmov.b64 {%0, %1}, a;\n\t}"
        : "+f"(a0), "+f"(a1) : "f"(b0), "f"(b1));
}
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

}  // namespace
}  // namespace lrs
