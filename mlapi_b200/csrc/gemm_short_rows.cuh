// short_rows_kernel and its re-scoring kernels: 13..32 classes on dense rows of at most 64 floats.
// One translation unit with gemm_tcgen05.cu (included there, in this order); not a stand-alone header.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "generic_row.cuh"

namespace lrs {
namespace {

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
    } else if ((C & 1) == 0) {                       // pairs: half the staging stores of the scalar form
#pragma unroll
        for (int i = 0; i < 32; i += 2)
            if (i < C) *reinterpret_cast<float2*>(stage + lane * C + i) = make_float2(v[i], v[i + 1]);
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
            for (int i = 0; i < 32; i += 4) {                    // (128-bit broadcast loads: 16 instead of 64 per tile)
                const float4 cs = *reinterpret_cast<const float4*>(cscale_s + i), bi = *reinterpret_cast<const float4*>(bias_s + i);
                fma2(v[i], v[i + 1], cs.x, cs.y, bi.x, bi.y);    // padding carries -inf
                fma2(v[i + 2], v[i + 3], cs.z, cs.w, bi.z, bi.w);
                gm = fmaxf(fmaxf(gm, v[i]), fmaxf(v[i + 1], fmaxf(v[i + 2], v[i + 3])));
            }
            int garg = 31;
#pragma unroll
            for (int i = 30; i >= 0; --i) garg = (v[i] == gm) ? i : garg;   // first index wins
            if (p.decision) store_rows_coalesced(stage, v, C, lane, p.decision + row0 * C, rows_valid, z_al);
            const float ez = fmaxf(p.kappa * fmaxf(psum_scale, fabsf(gm)) + floor_err, 1e-30f);
            const float tau = gm - 2.0f * ez;
            float gs0 = 0.0f, gs1 = 0.0f;
            int gnear = 0;
#pragma unroll
            for (int i = 0; i < 32; i += 2) {                    // two columns per multiply / add (packed fp32)
                gnear += (v[i] > tau ? 1 : 0) + (v[i + 1] > tau ? 1 : 0);
                sub2(v[i], v[i + 1], gm, gm);
                mul2(v[i], v[i + 1], kLog2e, kLog2e);
                v[i] = ex2(v[i]);                                // (padding: exp(-inf) = 0)
                v[i + 1] = ex2(v[i + 1]);
                add2(gs0, gs1, v[i], v[i + 1]);
            }
            const float gs = gs0 + gs1;
            const float inv = __frcp_rn(gs);
            // |dp| ~ p (1 - p) |dz| is largest for the class nearest to one half, and that is always the top one
            // (p (1 - p) rises up to 1/2, and no class but the top one can pass it): its probability is `inv`
            const bool risk = p.need_prob && inv * (1.0f - inv) > 8e-6f / ez;
#pragma unroll
            for (int i = 0; i < 32; i += 2) mul2(v[i], v[i + 1], inv, inv);
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

}  // namespace
}  // namespace lrs
