// Refit row, multinomial models: the transposed product of the gradient,
//     grad_W = R^T X        R = softmax(z) - onehot(y)  [n x C],   X [n x d]
// (`grad_pointwise.T @ X` in LinearModelLoss.loss_gradient, SK/linear_model/_linear_loss.py:353-357, which
// `LogisticRegression().fit` -- REF/Logistic Regression.ipynb:33-34 -- evaluates once per lbfgs iteration),
// on the tcgen05 tensor cores.  The contraction runs over the ROWS, so both operands are "MN-major" in UMMA
// terms: a tile of R is [64 rows (K)][classes], a tile of X is [64 rows (K)][features], each row of a tile
// 128 bytes of consecutive classes / features -- which is exactly how the two matrices lie in memory.  No
// transposition anywhere: TMA boxes of 64 x 64 halves with the 128-byte swizzle are the canonical MN-major
// operand tiles.
//
// Precision: fp32-grade, as in the scoring GEMM (gemm_tcgen05.cu): every product is split,
//     r x ~= r_hi x_hi + r_lo x_hi + r_hi x_lo          (fp16 halves, fp32 accumulation in TMEM)
// X is constant over a fit, so it is scaled (per feature, by the power of two that puts the column's largest
// magnitude at the top of fp16's range) and split ONCE, when the plan is made; R changes every evaluation and
// is split (scale 2^13: |r| <= 1) by a streaming kernel that also produces the column sums (grad_b).
// TMEM accumulation truncates instead of rounding, and the error grows with the number of accumulation steps:
// a CTA therefore flushes its accumulator every 32 k-chunks (2048 rows) into an fp32 tile in global memory
// that it owns (plain read-modify-write by the same thread in the same order: deterministic), through two
// accumulator buffers so that the MMAs of the next 2048 rows run under the flush.  Row slabs (one per CTA of
// a tile) are summed in float64 in a fixed order at the end.
//
// One CTA = 128 classes x 256 features x one slab of rows.  Warps: TMA producer, MMA issuer, four epilogue
// warps (TMEM lane quadrants).  Two stages of [R_hi 16 KB | R_lo 16 KB | X_hi 32 KB | X_lo 32 KB].
// Tensor roofline: 2 n C d flops counted once, executed three times.
#include <cuda.h>
#include <cuda_fp16.h>

#include <cstdio>
#include <vector>

#include "common.cuh"

namespace lrs {
namespace {

constexpr int GM = 128;                 // classes per CTA (TMEM lanes)
constexpr int GN = 256;                 // features per CTA (one MMA)
constexpr int GK = 64;                  // rows per k-chunk
constexpr int kAtomBytes = 64 * 128;    // one MN-major swizzle atom column: 64 rows (K) x 128 B (64 halves along MN)
constexpr int kAHalf = GM / 64 * kAtomBytes;     // 16 KB: R_hi (or R_lo) tile
constexpr int kBHalf = GN / 64 * kAtomBytes;     // 32 KB: X_hi (or X_lo) tile
constexpr int kStageBytes = 2 * kAHalf + 2 * kBHalf;   // 96 KB
constexpr int kStages = 2;
constexpr int kFlushChunks = 32;        // k-chunks accumulated in TMEM between flushes
constexpr int kGradThreads = 6 * 32;
constexpr float kRScale = 8192.0f;      // 2^13

struct GradParams {
    int n_chunks;            // k-chunks of 64 rows in this call's block of rows
    int tiles_n;             // feature tiles
    int slabs;               // row slabs per output tile
    int pm, pn;              // rows / columns of one slab's partial matrix (tiles_m * 128, tiles_n * 256)
    float* partial;          // [slabs][pm][pn]
};

__device__ __forceinline__ void g_tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ bool g_elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void g_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void g_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void g_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void g_mma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void g_tmem_ld32(uint32_t taddr, float (&v)[32]) {
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

// MN-major, 128-byte-swizzled operand tile: rows of the tile are K (here: matrix rows), 128 B = 64 halves
// along M / N per row, 8-row groups of 1024 B (stride byte offset), 64-wide MN blocks 8 KB apart (leading
// byte offset).  Canonical form ((8,8,m),(8,k)) : ((1,8,LBO),(64,SBO)) in halves -- cute/atom/mma_traits_sm100.hpp.
__device__ __forceinline__ uint64_t make_mn_desc(uint32_t smem_addr) {
    uint64_t desc = 0;
    desc |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);          // start address
    desc |= (uint64_t)(kAtomBytes >> 4) << 16;               // leading byte offset: next 64 columns of M / N
    desc |= (uint64_t)(1024 >> 4) << 32;                     // stride byte offset: next 8 rows of K
    desc |= (uint64_t)1 << 46;                               // descriptor version (sm_100)
    desc |= (uint64_t)2 << 61;                               // SWIZZLE_128B
    return desc;
}
constexpr uint32_t kGradIdesc = (1u << 4)                    // D = f32, A = B = f16 (format 0)
                                | (1u << 15) | (1u << 16)    // A and B are MN-major
                                | ((uint32_t)(GN >> 3) << 17) | ((uint32_t)(GM >> 4) << 24);

struct GradSmem {
    // [kStages][kStageBytes] from the 1024-aligned base, then:
    static constexpr int bars = kStages * kStageBytes;                 // full[kStages], empty[kStages], acc_full[2], acc_empty[2]
    static constexpr int tmem_slot = bars + (2 * kStages + 4) * 8;
    static constexpr int total = tmem_slot + 16;
};
static_assert(GradSmem::total + 1024 <= 227 * 1024, "shared memory budget");

__global__ void __launch_bounds__(kGradThreads, 1)
grad_gemm_kernel(const __grid_constant__ CUtensorMap map_rhi, const __grid_constant__ CUtensorMap map_rlo,
                 const __grid_constant__ CUtensorMap map_xhi, const __grid_constant__ CUtensorMap map_xlo,
                 const GradParams p) {
    extern __shared__ unsigned char gsmem_dyn[];
    unsigned char* smem = gsmem_dyn + ((1024u - (smem_u32(gsmem_dyn) & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + GradSmem::bars);
    uint64_t* full = bars;
    uint64_t* empty = full + kStages;
    uint64_t* acc_full = empty + kStages;
    uint64_t* acc_empty = acc_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + GradSmem::tmem_slot);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // blockIdx.x = (tile_m * tiles_n + tile_n) * slabs + slab
    const int slab = blockIdx.x % p.slabs;
    const int tile = blockIdx.x / p.slabs;
    const int tile_n = tile % p.tiles_n, tile_m = tile / p.tiles_n;
    // this CTA's k-chunks: a contiguous range, sizes differ by at most one
    const int kc_lo = (int)((int64_t)p.n_chunks * slab / p.slabs);
    const int kc_hi = (int)((int64_t)p.n_chunks * (slab + 1) / p.slabs);
    const int n_kc = kc_hi - kc_lo;
    const int n_flush = (n_kc + kFlushChunks - 1) / kFlushChunks;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kStages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(acc_full + i, 1); mbar_init(acc_empty + i, 128); }
        mbar_fence_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    g_fence_before();
    __syncthreads();
    g_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        const bool leader = g_elect_one();
        int st = 0;
        uint32_t ph = 0;
        for (int kc = kc_lo; kc < kc_hi; ++kc) {
            mbar_wait(empty + st, ph ^ 1);
            if (leader) {
                unsigned char* s = smem + st * kStageBytes;
                mbar_expect_tx(full + st, kStageBytes);
                const int row = kc * GK;
#pragma unroll
                for (int j = 0; j < GM / 64; ++j) {
                    g_tma_load_2d(s + j * kAtomBytes, &map_rhi, tile_m * GM + j * 64, row, full + st);
                    g_tma_load_2d(s + kAHalf + j * kAtomBytes, &map_rlo, tile_m * GM + j * 64, row, full + st);
                }
#pragma unroll
                for (int j = 0; j < GN / 64; ++j) {
                    g_tma_load_2d(s + 2 * kAHalf + j * kAtomBytes, &map_xhi, tile_n * GN + j * 64, row, full + st);
                    g_tma_load_2d(s + 2 * kAHalf + kBHalf + j * kAtomBytes, &map_xlo, tile_n * GN + j * 64, row, full + st);
                }
            }
            __syncwarp();
            if (++st == kStages) { st = 0; ph ^= 1; }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (whole warp walks the loop, one lane issues) =====================
        const bool leader = g_elect_one();
        const uint64_t base = make_mn_desc(smem_u32(smem));
        int st = 0;
        uint32_t ph = 0;
        int kc = 0;
        for (int f = 0; f < n_flush; ++f) {
            const uint32_t buf = (uint32_t)f & 1u;
            mbar_wait(acc_empty + buf, (((uint32_t)f >> 1) & 1u) ^ 1u);
            g_fence_after();
            const uint32_t d_tmem = tmem_base + buf * (uint32_t)GN;
            const int kc_end = min(n_kc, kc + kFlushChunks);
            for (uint32_t first = 0; kc < kc_end; ++kc, first = 1) {
                mbar_wait(full + st, ph);
                g_fence_after();
                const uint64_t ahi = base + (uint64_t)((st * kStageBytes) >> 4);
                const uint64_t alo = ahi + (uint64_t)(kAHalf >> 4);
                const uint64_t bhi = ahi + (uint64_t)((2 * kAHalf) >> 4);
                const uint64_t blo = bhi + (uint64_t)(kBHalf >> 4);
                if (leader) {
#pragma unroll
                    for (int ks = 0; ks < GK / 16; ++ks) {          // 16 rows of K = two 8-row groups = 2048 B
                        const uint64_t o = (uint64_t)(ks * (2048 >> 4));
                        g_mma_f16(d_tmem, ahi + o, bhi + o, kGradIdesc, (ks > 0) ? 1u : first);
                        g_mma_f16(d_tmem, alo + o, bhi + o, kGradIdesc, 1u);
                        g_mma_f16(d_tmem, ahi + o, blo + o, kGradIdesc, 1u);
                    }
                    g_commit(empty + st);
                }
                __syncwarp();
                if (++st == kStages) { st = 0; ph ^= 1; }
            }
            if (leader) g_commit(acc_full + buf);
            __syncwarp();
        }
    } else {
        // ===================== epilogue: flush the accumulator into this CTA's fp32 tile =====================
        const int q = warp & 3;                                     // TMEM lane quadrant this warp may read
        const int r_in_tile = q * 32 + lane;                        // class row of the tile
        float* my = p.partial + ((size_t)slab * p.pm + (size_t)tile_m * GM + r_in_tile) * p.pn + (size_t)tile_n * GN;
        for (int f = 0; f < n_flush; ++f) {
            const uint32_t buf = (uint32_t)f & 1u;
            mbar_wait(acc_full + buf, ((uint32_t)f >> 1) & 1u);
            g_fence_after();
            const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16) + buf * (uint32_t)GN;
#pragma unroll 1
            for (int cb = 0; cb < GN; cb += 32) {
                float v[32];
                g_tmem_ld32(t_lane + (uint32_t)cb, v);
                float4* dst = reinterpret_cast<float4*>(my + cb);
                if (f > 0) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float4 o = dst[i];
                        v[4 * i] += o.x; v[4 * i + 1] += o.y; v[4 * i + 2] += o.z; v[4 * i + 3] += o.w;
                    }
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            }
            g_fence_before();
            mbar_arrive(acc_empty + buf);
        }
        if (n_flush == 0) {                                          // a slab without rows: its tile is zero
            for (int cb = 0; cb < GN; cb += 4) *reinterpret_cast<float4*>(my + cb) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    g_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
}

// ---- streaming helpers ----------------------------------------------------------------------------
// largest |x| per feature over all rows (non-finite values are ignored: the forward pass reports those)
__global__ void __launch_bounds__(256)
col_absmax_kernel(const float* __restrict__ X, int64_t n_rows, int64_t ldx, int d, unsigned* __restrict__ colmax_bits) {
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= d) return;
    float mx = 0.0f;
    for (int64_t i = blockIdx.y; i < n_rows; i += gridDim.y) {
        const float a = fabsf(ldg_stream1(X + i * ldx + k));
        if (a <= 3.0e38f) mx = fmaxf(mx, a);
    }
    atomicMax(colmax_bits + k, __float_as_uint(mx));               // non-negative floats order like their bits
}
// scale[k] = power of two that puts colmax[k] in [2^14, 2^15); inv[k] = 1 / scale[k]
__global__ void col_scale_kernel(const unsigned* __restrict__ colmax_bits, int d, float* __restrict__ scale,
                                 double* __restrict__ inv) {
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= d) return;
    const float mx = __uint_as_float(colmax_bits[k]);
    int e = 0;
    float sc = 1.0f;
    if (mx > 0.0f) {
        frexpf(mx, &e);
        e = 15 - e;
        e = e > 100 ? 100 : (e < -100 ? -100 : e);
        sc = ldexpf(1.0f, e);
    }
    scale[k] = sc;
    inv[k] = 1.0 / (double)sc;
}
// X -> fp16 hi / lo at pitch d8 (zero beyond d); a value that still overflows fp16 (a non-finite input) stays
// non-finite and shows up in the gradient, as it does in sklearn's
__global__ void __launch_bounds__(256)
x_split_kernel(const float* __restrict__ X, int64_t n_rows, int64_t ldx, int d, int d8, const float* __restrict__ scale,
               __half* __restrict__ hi, __half* __restrict__ lo) {
    const int64_t total = n_rows * (int64_t)(d8 / 2);
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int64_t i = t / (d8 / 2);
        const int k = (int)(t % (d8 / 2)) * 2;
        float a = 0.0f, b = 0.0f;
        if (k < d) a = ldg_stream1(X + i * ldx + k) * scale[k];
        if (k + 1 < d) b = ldg_stream1(X + i * ldx + k + 1) * scale[k + 1];
        const __half2 h = __floats2half2_rn(a, b);
        const float2 hf = __half22float2(h);
        const __half2 l = __floats2half2_rn(a - hf.x, b - hf.y);
        *reinterpret_cast<__half2*>(hi + i * d8 + k) = h;
        *reinterpret_cast<__half2*>(lo + i * d8 + k) = l;
    }
}
// R -> fp16 hi / lo (scale 2^13) at pitch c8, zero beyond C, and per-CTA column sums (float64) for grad_b.
// A CTA takes a contiguous range of rows; a thread owns column pairs threadIdx.x + 256 j.
constexpr int kMaxPairsPerThread = 4;       // 2048 classes
__global__ void __launch_bounds__(256)
r_split_kernel(const float* __restrict__ R, int64_t n_rows, int64_t ldr, int C, int c8, __half* __restrict__ hi,
               __half* __restrict__ lo, double* __restrict__ colsum_part /* [grid][c8] */) {
    const int64_t r_lo = n_rows * blockIdx.x / gridDim.x, r_hi = n_rows * (blockIdx.x + 1) / gridDim.x;
    double s0[kMaxPairsPerThread], s1[kMaxPairsPerThread];
#pragma unroll
    for (int j = 0; j < kMaxPairsPerThread; ++j) s0[j] = s1[j] = 0.0;
    for (int64_t i = r_lo; i < r_hi; ++i) {
        const float* r = R + i * ldr;
#pragma unroll
        for (int j = 0; j < kMaxPairsPerThread; ++j) {
            const int k = (threadIdx.x + 256 * j) * 2;
            if (k >= c8) break;
            const float a = k < C ? ldg_stream1(r + k) : 0.0f;
            const float b = k + 1 < C ? ldg_stream1(r + k + 1) : 0.0f;
            s0[j] += (double)a;
            s1[j] += (double)b;
            const __half2 h = __floats2half2_rn(a * kRScale, b * kRScale);
            const float2 hf = __half22float2(h);
            const __half2 l = __floats2half2_rn(a * kRScale - hf.x, b * kRScale - hf.y);
            *reinterpret_cast<__half2*>(hi + i * c8 + k) = h;
            *reinterpret_cast<__half2*>(lo + i * c8 + k) = l;
        }
    }
#pragma unroll
    for (int j = 0; j < kMaxPairsPerThread; ++j) {
        const int k = (threadIdx.x + 256 * j) * 2;
        if (k >= c8) break;
        colsum_part[(size_t)blockIdx.x * c8 + k] = s0[j];
        colsum_part[(size_t)blockIdx.x * c8 + k + 1] = s1[j];
    }
}
// out[c][k] = sum over slabs (fixed order, float64) of partial[s][c][k], unscaled; out[c][d] = sum_i R[i][c]
__global__ void __launch_bounds__(256)
grad_combine_kernel(const float* __restrict__ partial, int slabs, int pm, int pn, int C, int d,
                    const double* __restrict__ inv_scale, const double* __restrict__ colsum_part, int n_part, int c8,
                    double* __restrict__ out) {
    const int64_t total = (int64_t)C * (d + 1);
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int c = (int)(t / (d + 1)), k = (int)(t % (d + 1));
        double s = 0.0;
        if (k < d) {
            for (int sl = 0; sl < slabs; ++sl) s += (double)partial[((size_t)sl * pm + c) * pn + k];
            s *= inv_scale[k] * (1.0 / (double)kRScale);
        } else {
            for (int b = 0; b < n_part; ++b) s += colsum_part[(size_t)b * c8 + c];
        }
        out[t] = s;
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode_tiled() {
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
// fp16 [rows][cols] at pitch `ld` halves; box = 64 rows x 64 columns (128 B), 128-byte swizzle
int make_half_map(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld) {
    EncodeTiledFn fn = g_encode_tiled();
    if (!fn) return fail(LR_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)GK};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(LR_ERR_CUDA, "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    return LR_OK;
}

constexpr int kSplitBlocks = 1184;      // CTAs of r_split_kernel = rows of the column-sum scratch

}  // namespace

struct GradPlan {
    int device = 0;
    int num_sms = 148;
    int64_t n_rows = 0;
    int d = 0, d8 = 0;
    int max_scores = 0, c8_max = 0;
    int64_t max_block_rows = 0;
    __half* x_hi = nullptr;
    __half* x_lo = nullptr;
    float* x_scale = nullptr;        // [d]
    double* x_inv_scale = nullptr;   // [d]
    __half* r_hi = nullptr;          // [max_block_rows][c8_max]
    __half* r_lo = nullptr;
    float* partial = nullptr;        // [slabs][pm][pn], sized for the largest shape
    size_t partial_floats = 0;
    double* colsum_part = nullptr;   // [kSplitBlocks][c8_max]
    bool configured = false;
};

}  // namespace lrs

using namespace lrs;

extern "C" {

int lr_grad_plan_create(const float* X_dev, int64_t n_rows, int64_t ldx, int32_t n_features, int32_t max_scores,
                        int64_t max_block_rows, void* stream, lr_grad_plan** out) {
    if (!out) return fail(LR_ERR_INVALID, "out is NULL");
    *out = nullptr;
    if (!X_dev || n_rows < 1 || n_features < 1 || ldx < n_features)
        return fail(LR_ERR_INVALID, "X is NULL, empty, or its pitch is smaller than n_features");
    if (max_scores < 1 || max_scores > 2 * 256 * kMaxPairsPerThread)
        return fail(LR_ERR_LIMIT, "lr_grad_gemm serves 1..2048 score rows");
    if (max_block_rows < 1 || max_block_rows > n_rows) max_block_rows = n_rows;
    if (n_rows >= ((int64_t)1 << 31)) return fail(LR_ERR_LIMIT, "lr_grad_gemm serves fewer than 2^31 rows per plan");
    cudaPointerAttributes attr;
    LRS_CUDA(cudaPointerGetAttributes(&attr, X_dev));
    if (attr.type != cudaMemoryTypeDevice) return fail(LR_ERR_INVALID, "X_dev is not device memory");
    int prev = -1;
    cudaGetDevice(&prev);
    if (prev != attr.device) LRS_CUDA(cudaSetDevice(attr.device));
    struct Restore { int prev, cur; ~Restore() { if (prev >= 0 && prev != cur) cudaSetDevice(prev); } } restore{prev, attr.device};
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    GradPlan* g = new GradPlan();
    g->device = attr.device;
    cudaDeviceGetAttribute(&g->num_sms, cudaDevAttrMultiProcessorCount, attr.device);
    g->n_rows = n_rows;
    g->d = n_features;
    g->d8 = (n_features + 7) / 8 * 8;
    g->max_scores = max_scores;
    g->c8_max = (max_scores + 7) / 8 * 8;
    g->max_block_rows = max_block_rows;
    auto bail = [&](int rc) { lr_grad_plan_destroy(reinterpret_cast<lr_grad_plan*>(g)); return rc; };
    // one fp32 tile per CTA; a call launches at most max(num_sms, tiles) CTAs whatever its class count
    const int tiles_m = (max_scores + GM - 1) / GM, tiles_n = (n_features + GN - 1) / GN;
    const int max_ctas = tiles_m * tiles_n > g->num_sms ? tiles_m * tiles_n : g->num_sms;
    g->partial_floats = (size_t)max_ctas * GM * GN;
    unsigned* colmax = nullptr;
    if (cudaMalloc(&g->x_hi, (size_t)n_rows * g->d8 * sizeof(__half)) != cudaSuccess ||
        cudaMalloc(&g->x_lo, (size_t)n_rows * g->d8 * sizeof(__half)) != cudaSuccess ||
        cudaMalloc(&g->x_scale, n_features * sizeof(float)) != cudaSuccess ||
        cudaMalloc(&g->x_inv_scale, n_features * sizeof(double)) != cudaSuccess ||
        cudaMalloc(&g->r_hi, (size_t)max_block_rows * g->c8_max * sizeof(__half)) != cudaSuccess ||
        cudaMalloc(&g->r_lo, (size_t)max_block_rows * g->c8_max * sizeof(__half)) != cudaSuccess ||
        cudaMalloc(&g->partial, g->partial_floats * sizeof(float)) != cudaSuccess ||
        cudaMalloc(&g->colsum_part, (size_t)kSplitBlocks * g->c8_max * sizeof(double)) != cudaSuccess ||
        cudaMalloc(&colmax, n_features * sizeof(unsigned)) != cudaSuccess) {
        cudaFree(colmax);
        return bail(fail(LR_ERR_CUDA, std::string("cudaMalloc(grad plan): ") + cudaGetErrorString(cudaGetLastError())));
    }
    cudaMemsetAsync(colmax, 0, n_features * sizeof(unsigned), s);
    int64_t gy = n_rows < 4096 ? n_rows : 4096;
    col_absmax_kernel<<<dim3((n_features + 255) / 256, (unsigned)gy), 256, 0, s>>>(X_dev, n_rows, ldx, n_features, colmax);
    col_scale_kernel<<<(n_features + 255) / 256, 256, 0, s>>>(colmax, n_features, g->x_scale, g->x_inv_scale);
    x_split_kernel<<<g->num_sms * 8, 256, 0, s>>>(X_dev, n_rows, ldx, n_features, g->d8, g->x_scale, g->x_hi, g->x_lo);
    count_launch(3);
    const cudaError_t e = cudaStreamSynchronize(s);
    cudaFree(colmax);
    if (e != cudaSuccess) return bail(fail(LR_ERR_CUDA, std::string("grad plan kernels: ") + cudaGetErrorString(e)));
    *out = reinterpret_cast<lr_grad_plan*>(g);
    return LR_OK;
}

void lr_grad_plan_destroy(lr_grad_plan* h) {
    if (!h) return;
    GradPlan* g = reinterpret_cast<GradPlan*>(h);
    int prev = -1;
    cudaGetDevice(&prev);
    if (prev != g->device) cudaSetDevice(g->device);
    cudaDeviceSynchronize();
    cudaFree(g->x_hi); cudaFree(g->x_lo); cudaFree(g->x_scale); cudaFree(g->x_inv_scale);
    cudaFree(g->r_hi); cudaFree(g->r_lo); cudaFree(g->partial); cudaFree(g->colsum_part);
    if (prev >= 0 && prev != g->device) cudaSetDevice(prev);
    delete g;
}

int lr_grad_gemm(lr_grad_plan* h, int64_t row0, int64_t n_rows, const float* R_dev, int64_t ldr, int32_t n_scores,
                 double* out_dev, void* stream) {
    if (!h || !R_dev || !out_dev) return fail(LR_ERR_INVALID, "plan / R / out is NULL");
    GradPlan* g = reinterpret_cast<GradPlan*>(h);
    if (n_scores < 1 || n_scores > g->max_scores) return fail(LR_ERR_LIMIT, "n_scores exceeds the plan's max_scores");
    if (row0 < 0 || n_rows < 1 || row0 + n_rows > g->n_rows || n_rows > g->max_block_rows)
        return fail(LR_ERR_INVALID, "row block outside the plan's rows (or larger than max_block_rows)");
    if (ldr < n_scores) return fail(LR_ERR_INVALID, "row pitch of R is smaller than n_scores");
    int prev = -1;
    cudaGetDevice(&prev);
    if (prev != g->device) LRS_CUDA(cudaSetDevice(g->device));
    struct Restore { int prev, cur; ~Restore() { if (prev >= 0 && prev != cur) cudaSetDevice(prev); } } restore{prev, g->device};
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int C = n_scores, c8 = (C + 7) / 8 * 8, d = g->d;
    const int tiles_m = (C + GM - 1) / GM, tiles_n = (d + GN - 1) / GN;
    const int n_chunks = (int)((n_rows + GK - 1) / GK);
    int slabs = g->num_sms / (tiles_m * tiles_n);
    if (slabs < 1) slabs = 1;
    if (slabs > n_chunks) slabs = n_chunks;
    GradParams p;
    p.n_chunks = n_chunks;
    p.tiles_n = tiles_n;
    p.slabs = slabs;
    p.pm = tiles_m * GM;
    p.pn = tiles_n * GN;
    p.partial = g->partial;
    if ((size_t)slabs * p.pm * p.pn > g->partial_floats) return fail(LR_ERR_LIMIT, "partial buffer too small (internal)");
    int n_part = (int)(n_rows < kSplitBlocks ? n_rows : kSplitBlocks);
    r_split_kernel<<<n_part, 256, 0, s>>>(R_dev, n_rows, ldr, C, c8, g->r_hi, g->r_lo, g->colsum_part);
    count_launch();
    CUtensorMap m_rhi, m_rlo, m_xhi, m_xlo;
    int rc;
    // per-call maps: rows beyond this block read as zero (TMA fills what lies outside the tensor)
    if ((rc = make_half_map(&m_rhi, g->r_hi, (uint64_t)n_rows, (uint64_t)c8, (uint64_t)c8)) != LR_OK) return rc;
    if ((rc = make_half_map(&m_rlo, g->r_lo, (uint64_t)n_rows, (uint64_t)c8, (uint64_t)c8)) != LR_OK) return rc;
    if ((rc = make_half_map(&m_xhi, g->x_hi + row0 * g->d8, (uint64_t)n_rows, (uint64_t)g->d8, (uint64_t)g->d8)) != LR_OK) return rc;
    if ((rc = make_half_map(&m_xlo, g->x_lo + row0 * g->d8, (uint64_t)n_rows, (uint64_t)g->d8, (uint64_t)g->d8)) != LR_OK) return rc;
    if (!g->configured) {
        LRS_CUDA(cudaFuncSetAttribute(grad_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GradSmem::total + 1024));
        g->configured = true;
    }
    grad_gemm_kernel<<<tiles_m * tiles_n * slabs, kGradThreads, GradSmem::total + 1024, s>>>(m_rhi, m_rlo, m_xhi, m_xlo, p);
    count_launch();
    grad_combine_kernel<<<g->num_sms * 4, 256, 0, s>>>(g->partial, slabs, p.pm, p.pn, C, d, g->x_inv_scale,
                                                        g->colsum_part, n_part, c8, out_dev);
    count_launch();
    LRS_CUDA(cudaGetLastError());
    return LR_OK;
}

}  // extern "C"
