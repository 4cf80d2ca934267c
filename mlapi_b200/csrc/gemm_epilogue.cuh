// Register epilogues of gemm_softmax_kernel: models of at most 128 classes, whole rows in registers.
// One translation unit with gemm_tcgen05.cu (included there, in this order); not a stand-alone header.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "generic_row.cuh"

namespace lrs {
namespace {

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

}  // namespace
}  // namespace lrs
