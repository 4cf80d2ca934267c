// gemm_softmax_kernel: TMA-fed tcgen05 GEMM with the softmax fused into its epilogue (see gemm_tcgen05.cu).
// One translation unit with gemm_tcgen05.cu (included there, in this order); not a stand-alone header.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "generic_row.cuh"

namespace lrs {
namespace {

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

}  // namespace
}  // namespace lrs
