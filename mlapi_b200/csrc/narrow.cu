// Narrow path: models with at most 16 score columns (the notebook's iris model d=4, C=3 and the d=32
// binary north-star config among them).  Four kernels share one lane mapping, reduction and epilogue:
//   narrow_kernel       <= 4 columns, W in registers or shared memory, direct 128-bit (or 4-byte) loads
//   mid_kernel          5..16 columns: U rows x NS classes of accumulators per thread, W from shared memory / L1 / L2
//   narrow_flat_kernel  rows that are not 16-byte aligned, and aligned rows of 1 KB and more, in batches of
//                       >= 4096: whole tiles of rows through a cp.async.bulk ring in shared memory
//   narrow_bulk_kernel  the first ring variant (aligned short rows), opt-in, kept as the measured alternative
// For a given row alignment, a row scores to the same bits whichever of them its batch size selects.
//
// Replaces, per row, sklearn's decision_function (SK/linear_model/_base.py:390-391),
// _predict_proba_lr / softmax (SK/linear_model/_base.py:437-440, SK/utils/extmath.py:1012-1024),
// the label rule of predict (SK/linear_model/_base.py:415-418) and the `.max()` of REF/main.py:22
// in a single pass over X.
//
// Mapping.  LANES lanes (a power of two, <= 32) share a row; each lane owns V-float chunks of it
// (V = 4 -> one 128-bit load per chunk).  A warp covers G = 32/LANES rows per load instruction and
// keeps U such instructions in flight, so one warp iteration consumes G*U consecutive rows --
// a contiguous, fully coalesced span of X.  The partial dot products are then summed across the
// LANES lanes with a halving exchange ("transpose reduce"): instead of every lane ending up with
// every sum, each lane ends up owning complete sums of distinct rows, so the sigmoid/softmax
// epilogue and the output stores are spread over all 32 lanes and the P rows a warp writes are
// again one contiguous span.
//
// HBM roofline: algorithmic bytes per row = 4*d read + 4*max(C,2) written (+4 per optional
// label / maxprob / decision column); W and b are read once per CTA and excluded.
#include <cmath>
#include <cstdlib>
#include <map>
#include <mutex>
#include <utility>

#include "common.cuh"

namespace lrs {

namespace {

constexpr int kThreads = 256;
constexpr size_t kWSmemMax = 96 * 1024;    // W staged in shared memory up to here
// 17..24 columns are served up to this many features (beyond, and for more columns: tensor cores).  Measured:
// d=32, C=20 1.27 -> 1.86 TB/s, d=54, C=20 0.86 -> 1.69; a 32-column form lost to the tensor path (d=16, C=32: 0.9 : 1.6).
constexpr int kMid24MaxD = 56;
constexpr size_t kWL1Max = 16 << 20;        // 5..16 columns: beyond shared memory, W is read through L1 (hits up to
                                           // ~192 KB, L2 beyond: d=8000, C=10 still runs 4x faster than the generic kernel)

// Lanes per row of the scalar-load kernels (rows that are not 16-byte aligned), the plain one and the
// flat-tile one alike: one lane per four floats as in the 128-bit kernels, up to 32 -- a lane then makes
// four scalar loads where those make one vector load, and the cross-lane reduction and the epilogue
// stay as cheap per row as there.
inline int scalar_lanes(int d) {
    const int chunks = (d + 3) / 4;
    int lanes = 4;
    while (lanes < chunks && lanes < 32) lanes <<= 1;
    // d just above a power of two (33: 16 lanes x 3 steps = 48 slots): half the lanes waste less (8 x 5 = 40)
    if (lanes > 4 && lanes <= 16) {
        const int full = (d + lanes - 1) / lanes * lanes, half = (d + lanes / 2 - 1) / (lanes / 2) * (lanes / 2);
        if (half < full) lanes /= 2;
    }
    return lanes;
}

// Score columns the kernels are instantiated for (padding columns carry zero weights and a -inf intercept).
inline int padded_scores(int c) {
    return c == 1 ? 1 : c <= 3 ? 3 : c <= 4 ? 4 : c <= 6 ? 6 : c <= 8 ? 8 : c <= 10 ? 10 : c <= 12 ? 12 : c <= 16 ? 16 : 24;
}

template <int V>
struct Vec;
template <>
struct Vec<4> {
    float v[4];
    __device__ __forceinline__ void load(const float* p) {
        float4 t = ldg_stream4(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
};
template <>
struct Vec<1> {
    float v[1];
    __device__ __forceinline__ void load(const float* p) { v[0] = ldg_stream1(p); }
};

struct Outs {
    float* proba;
    float* decision;
    int32_t* label;
    float* maxprob;
    unsigned long long* status;
    int n_scores;   // real score columns (<= NS)
};

// Epilogue for one complete row whose NS scores (intercept added) sit in z[].
template <int NS>
__device__ __forceinline__ void finish_row(const float (&z)[NS], int64_t row, const Outs& o) {
    if constexpr (NS == 1) {
        const float zz = z[0];
        if (!finite_f(zz)) atomicAdd(o.status, 1ull);
        const float2 p = sigmoid_pair(zz);
        if (o.proba) __stcs(reinterpret_cast<float2*>(o.proba) + row, p);
        if (o.decision) __stcs(o.decision + row, zz);
        if (o.label) __stcs(o.label + row, zz > 0.0f ? 1 : 0);
        if (o.maxprob) __stcs(o.maxprob + row, fmaxf(p.x, p.y));
    } else {
        float zmax = z[0];
        int arg = 0;
        bool ok = true;
#pragma unroll
        for (int c = 0; c < NS; ++c) {
            if (c < o.n_scores) ok = ok && finite_f(z[c]);
            if (c > 0 && z[c] > zmax) { zmax = z[c]; arg = c; }   // strict: first maximum wins
        }
        if (!ok) atomicAdd(o.status, 1ull);
        float e[NS];
        float sum = 0.0f;
#pragma unroll
        for (int c = 0; c < NS; ++c) { e[c] = expf(z[c] - zmax); sum += e[c]; }
        const float inv = __frcp_rn(sum);
        if (o.proba) {
            float* p = o.proba + row * o.n_scores;
#pragma unroll
            for (int c = 0; c < NS; ++c)
                if (c < o.n_scores) __stcs(p + c, e[c] * inv);
        }
        if (o.decision) {
            float* q = o.decision + row * o.n_scores;
#pragma unroll
            for (int c = 0; c < NS; ++c)
                if (c < o.n_scores) __stcs(q + c, z[c]);
        }
        if (o.label) __stcs(o.label + row, arg);
        if (o.maxprob) __stcs(o.maxprob + row, inv);   // exp(zmax - zmax) == 1
    }
}

// SINGLE: every lane owns exactly one chunk of the row, so its NS*V coefficients live in registers.
template <int V, int LANES, int NS, int U, bool SINGLE>
__global__ void __launch_bounds__(kThreads)
narrow_kernel(const float* __restrict__ X, int64_t n_rows, int64_t ldx, int d, int dpad,
              const float* __restrict__ Wn, const float* __restrict__ bn, Outs o) {
    constexpr int G = 32 / LANES;                 // rows per load instruction per warp
    constexpr int M = (U < LANES) ? U : LANES;    // lanes that end up owning distinct rows
    constexpr int ROWS_OWNED = U / M;
    constexpr int ROWS_PER_ITER = G * U;

    extern __shared__ __align__(16) float w_s[];  // [NS][dpad]
    if constexpr (!SINGLE) {
        for (int i = threadIdx.x; i < NS * dpad; i += kThreads) w_s[i] = Wn[i];
        __syncthreads();
    }

    const int lane = threadIdx.x & 31;
    const int g = lane / LANES;
    const int s = lane % LANES;
    const int k0 = s * V;

    float wreg[NS][V];
    if constexpr (SINGLE) {
#pragma unroll
        for (int c = 0; c < NS; ++c)
#pragma unroll
            for (int v = 0; v < V; ++v) wreg[c][v] = Wn[c * dpad + k0 + v];
    }
    float bias[NS];
#pragma unroll
    for (int c = 0; c < NS; ++c) bias[c] = bn[c];

    const int64_t warps_total = (int64_t)gridDim.x * (kThreads / 32);
    const int64_t warp_id = (int64_t)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
    const int64_t n_tiles = (n_rows + ROWS_PER_ITER - 1) / ROWS_PER_ITER;

    for (int64_t tile = warp_id; tile < n_tiles; tile += warps_total) {
        const int64_t base = tile * ROWS_PER_ITER;
        float acc[U][NS];
#pragma unroll
        for (int j = 0; j < U; ++j)
#pragma unroll
            for (int c = 0; c < NS; ++c) acc[j][c] = 0.0f;

        // All U loads of a chunk are issued back to back and unpredicated.  Full tiles (all but the
        // last) step a single pointer by a fixed stride; the last tile clamps row indices to the
        // final row (those results are never stored).  Lanes past the end of a row re-read column 0
        // against zero weights, so nothing here can fault or branch.
        const bool full_tile = base + ROWS_PER_ITER <= n_rows;
        const int64_t jstride = (int64_t)G * ldx;
        if constexpr (SINGLE) {
            Vec<V> x[U];
            const int kc = k0 < d ? k0 : 0;
            if (full_tile) {
                const float* p = X + (base + g) * ldx + kc;
#pragma unroll
                for (int j = 0; j < U; ++j) x[j].load(p + j * jstride);
            } else {
#pragma unroll
                for (int j = 0; j < U; ++j) {
                    int64_t row = base + j * G + g;
                    row = row < n_rows ? row : n_rows - 1;
                    x[j].load(X + row * ldx + kc);
                }
            }
#pragma unroll
            for (int j = 0; j < U; ++j)
#pragma unroll
                for (int c = 0; c < NS; ++c)
#pragma unroll
                    for (int v = 0; v < V; ++v) acc[j][c] = fmaf(x[j].v[v], wreg[c][v], acc[j][c]);
        } else {
            const int dloop = (d + LANES * V - 1) / (LANES * V) * (LANES * V);   // <= dpad
            const float* rowp[U];
#pragma unroll
            for (int j = 0; j < U; ++j) {
                int64_t row = base + j * G + g;
                row = row < n_rows ? row : n_rows - 1;
                rowp[j] = X + row * ldx;
            }
            for (int k = k0; k < dloop; k += LANES * V) {
                Vec<V> x[U];
                const int kc = k < d ? k : 0;      // past-the-end chunks meet zero weights
#pragma unroll
                for (int j = 0; j < U; ++j) x[j].load(rowp[j] + kc);
#pragma unroll
                for (int c = 0; c < NS; ++c) {
                    float w[V];
                    if constexpr (V == 4) {
                        const float4 t = *reinterpret_cast<const float4*>(w_s + c * dpad + k);
                        w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w;
                    } else {
                        w[0] = w_s[c * dpad + k];
                    }
#pragma unroll
                    for (int j = 0; j < U; ++j)
#pragma unroll
                        for (int v = 0; v < V; ++v) acc[j][c] = fmaf(x[j].v[v], w[v], acc[j][c]);
                }
            }
        }

        lane_reduce<LANES / 2, U, NS, U>(acc, s);

        // lane s now owns rows u = sel*ROWS_OWNED + r (r < ROWS_OWNED) of its group, complete.
        const int sel = s / (LANES / M);
        const bool owner = (s % (LANES / M)) == 0;
        if (owner) {
#pragma unroll
            for (int r = 0; r < ROWS_OWNED; ++r) {
                const int u = sel * ROWS_OWNED + r;
                const int64_t row = base + u * G + g;
                if (row < n_rows) {
                    float z[NS];
#pragma unroll
                    for (int c = 0; c < NS; ++c) z[c] = acc[r][c] + bias[c];
                    finish_row<NS>(z, row, o);
                }
            }
        }
    }
}

// ---- 5..16 score columns -------------------------------------------------------------------------
// With more than four classes the kernel above runs out of shared-memory bandwidth: every 128-bit
// piece of W it reads feeds four FMAs of ONE row (measured on B200: C=10, d=784 at 1.4 TB/s).  Here a
// thread keeps U rows x NS classes of accumulators (U = 4: a piece of W feeds 16 FMAs; U = 2 at 12
// classes), and, up to 12 classes, the X pieces of the next k-step are requested before the FMAs of
// the current one (at 16 classes 4 x 16 accumulators plus two X buffers do not fit 128 registers).  NS is the exact padded class count (6, 8, 10, 12, 16):
// at C=10 the kernel needs 31 of the ~70 fp32 TFLOP/s to keep up with HBM, at 16 it would need 51.
// W comes from shared memory while it fits 96 KB, else through L1 (`__ldg`; X bypasses L1).
// fp32 FMA arithmetic throughout, same as the kernel above.
template <int NS>
struct MidRows { static constexpr int value = (NS == 12 || NS > 16) ? 2 : 4; };    // rows per thread (U)

// AL = false (with V = 4): rows that are not 16-byte aligned -- a lane still owns four consecutive floats per
// step (W pieces stay 128-bit loads), fetched with four 4-byte loads, each clamped on its own at the row end.
template <int V, int LANES, int NS, bool WSMEM, bool AL = true>
__global__ void __launch_bounds__(kThreads, 2)
mid_kernel(const float* __restrict__ X, int64_t n_rows, int64_t ldx, int d, int dpad,
           const float* __restrict__ Wn, const float* __restrict__ bn, Outs o) {
    constexpr int U = MidRows<NS>::value;
    constexpr int G = 32 / LANES;
    constexpr int M = (U < LANES) ? U : LANES;
    constexpr int ROWS_OWNED = U / M;
    constexpr int ROWS_PER_ITER = G * U;
    constexpr int STEP = LANES * V;

    extern __shared__ __align__(16) float w_s[];  // [NS][dpad] when WSMEM
    if constexpr (WSMEM) {
        for (int i = threadIdx.x; i < NS * dpad; i += kThreads) w_s[i] = Wn[i];
        __syncthreads();
    }
    const float* __restrict__ wsrc = WSMEM ? w_s : Wn;

    const int lane = threadIdx.x & 31;
    const int g = lane / LANES;
    const int s = lane % LANES;
    const int k0 = s * V;
    const int dloop = (d + STEP - 1) / STEP * STEP;   // <= dpad; W is zero beyond d

    const int64_t warps_total = (int64_t)gridDim.x * (kThreads / 32);
    const int64_t warp_id = (int64_t)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
    const int64_t n_tiles = (n_rows + ROWS_PER_ITER - 1) / ROWS_PER_ITER;

    for (int64_t tile = warp_id; tile < n_tiles; tile += warps_total) {
        const int64_t base = tile * ROWS_PER_ITER;
        const float* rowp[U];
#pragma unroll
        for (int j = 0; j < U; ++j) {
            int64_t row = base + j * G + g;
            row = row < n_rows ? row : n_rows - 1;     // clamped rows are computed and never stored
            rowp[j] = X + row * ldx;
        }
        float acc[U][NS];
#pragma unroll
        for (int j = 0; j < U; ++j)
#pragma unroll
            for (int c = 0; c < NS; ++c) acc[j][c] = 0.0f;

        auto load_piece = [&](Vec<V>& dst, const float* row, int k) {
            if constexpr (V == 4 && !AL) {
#pragma unroll
                for (int v = 0; v < 4; ++v) dst.v[v] = ldg_stream1(row + (k + v < d ? k + v : 0));
            } else {
                dst.load(row + (k < d ? k : 0));       // past-the-end pieces meet zero weights
            }
        };
        // One k-step: every class's piece of W against the four rows' pieces of X.
        auto fma_step = [&](const Vec<V> (&xv)[U], int k) {
#pragma unroll
            for (int c = 0; c < NS; ++c) {
                float w[V];
                if constexpr (V == 4) {
                    float4 t;
                    if constexpr (WSMEM) t = *reinterpret_cast<const float4*>(wsrc + c * dpad + k);
                    else t = __ldg(reinterpret_cast<const float4*>(wsrc + c * dpad + k));
                    w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w;
                } else {
                    w[0] = WSMEM ? wsrc[c * dpad + k] : __ldg(wsrc + c * dpad + k);
                }
#pragma unroll
                for (int j = 0; j < U; ++j)
#pragma unroll
                    for (int v = 0; v < V; ++v) acc[j][c] = fmaf(xv[j].v[v], w[v], acc[j][c]);
            }
        };
        // Software pipeline of depth one over the k-steps: the loads of step i+1 sit in the same basic
        // block as the FMAs of step i and ahead of them (a conditional prefetch was sunk below the
        // FMAs by the compiler, which left nothing in flight while they ran).  Every load is a real
        // piece of the row (or column 0 against zero weights past d), none is issued twice.
        // Measured (B200, tools/shape_sweep.py): 12 classes do best with two rows per thread and the
        // pipeline (4.46 against 4.04 TB/s at d=1024); 16 classes with four rows and one step at a time
        // (two rows + pipeline: 3.0 against 3.5 TB/s at d=128) -- neither reaches past ~27 fp32 TFLOP/s.
        if constexpr (NS <= 12 || NS > 16) {
            Vec<V> xa[U], xb[U];
#pragma unroll
            for (int j = 0; j < U; ++j) load_piece(xa[j], rowp[j], k0);
            int k = k0;
            for (; k + 2 * STEP < dloop; k += 2 * STEP) {
#pragma unroll
                for (int j = 0; j < U; ++j) load_piece(xb[j], rowp[j], k + STEP);
                fma_step(xa, k);
#pragma unroll
                for (int j = 0; j < U; ++j) load_piece(xa[j], rowp[j], k + 2 * STEP);
                fma_step(xb, k + STEP);
            }
            if (k + STEP < dloop) {                    // two steps left (same for the whole warp)
                const int kn = k + STEP;
#pragma unroll
                for (int j = 0; j < U; ++j) load_piece(xb[j], rowp[j], kn);
                fma_step(xa, k);
                fma_step(xb, kn);
            } else {
                fma_step(xa, k);
            }
        } else {
            for (int k = k0; k < dloop; k += STEP) {
                Vec<V> xa[U];
#pragma unroll
                for (int j = 0; j < U; ++j) load_piece(xa[j], rowp[j], k);
                fma_step(xa, k);
            }
        }

        lane_reduce<LANES / 2, U, NS, U>(acc, s);

        const int sel = s / (LANES / M);
        const bool owner = (s % (LANES / M)) == 0;
        if (owner) {
#pragma unroll
            for (int r = 0; r < ROWS_OWNED; ++r) {
                const int u = sel * ROWS_OWNED + r;
                const int64_t row = base + u * G + g;
                if (row < n_rows) {
                    float z[NS];
#pragma unroll
                    for (int c = 0; c < NS; ++c) z[c] = acc[r][c] + __ldg(bn + c);
                    finish_row<NS>(z, row, o);
                }
            }
        }
    }
}

// ---- bulk-copy ring variant ------------------------------------------------------------------------
// Same arithmetic as narrow_kernel<4, LANES, NS, U, true>, but X reaches the SM through the TMA
// engine: one producer lane streams contiguous row tiles (kConsumerWarps * G * U rows) into a ring of
// shared-memory stages with cp.async.bulk, completion counted on an mbarrier per stage; the eight
// consumer warps read their rows with conflict-free 128-bit shared loads and release the stage
// through a second mbarrier.  Bytes in flight per SM = stages x tile bytes (~190 KB), independent of
// register allocation and occupancy.  Measured on B200 (profiles/r01_narrow.md): 6553 GB/s against
// 6837 GB/s for the plain-load kernel once its eight loads are issued back to back, so this variant
// is opt-in (LRS_NARROW_BULK=1) and kept as the measured alternative.
// Needs dense rows (ldx == d), d % 4 == 0 and a 16-byte aligned X.
constexpr int kConsumerWarps = 8;
constexpr int kBulkThreads = (kConsumerWarps + 1) * 32;
constexpr int kBulkSmemBudget = 200 * 1024;

template <int LANES, int NS, int U>
__global__ void __launch_bounds__(kBulkThreads, 1)
narrow_bulk_kernel(const float* __restrict__ X, int64_t n_rows, int d, int dpad, int stages,
                   const float* __restrict__ Wn, const float* __restrict__ bn, Outs o) {
    constexpr int V = 4;
    constexpr int G = 32 / LANES;
    constexpr int M = (U < LANES) ? U : LANES;
    constexpr int ROWS_OWNED = U / M;
    constexpr int ROWS_PER_WARP = G * U;
    constexpr int TILE_ROWS = ROWS_PER_WARP * kConsumerWarps;

    extern __shared__ __align__(128) unsigned char smem_raw[];
    const uint32_t stage_bytes = (uint32_t)TILE_ROWS * d * sizeof(float);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);          // [stages]
    uint64_t* empty = full + stages;                                  // [stages]
    unsigned char* ring = smem_raw + 128 * ((2 * stages * 8 + 127) / 128);

    if (threadIdx.x == 0) {
        for (int i = 0; i < stages; ++i) {
            mbar_init(full + i, 1);
            mbar_init(empty + i, kConsumerWarps);
        }
        mbar_fence_init();
    }
    __syncthreads();

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int64_t n_tiles = (n_rows + TILE_ROWS - 1) / TILE_ROWS;

    if (warp == kConsumerWarps) {
        // ===== producer: one lane feeds the ring =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                mbar_wait(empty + stage, phase ^ 1);
                const int64_t row0 = tile * TILE_ROWS;
                const int64_t rows = (n_rows - row0 < TILE_ROWS) ? (n_rows - row0) : TILE_ROWS;
                const uint32_t bytes = (uint32_t)(rows * d * sizeof(float));
                mbar_expect_tx(full + stage, bytes);
                bulk_g2s(ring + (size_t)stage * stage_bytes, X + row0 * d, bytes, full + stage);
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

    // ===== consumers =====
    const int g = lane / LANES;
    const int s = lane % LANES;
    const int k0 = s * V;
    const int kc = k0 < d ? k0 : 0;
    float wreg[NS][V];
#pragma unroll
    for (int c = 0; c < NS; ++c)
#pragma unroll
        for (int v = 0; v < V; ++v) wreg[c][v] = Wn[c * dpad + k0 + v];
    float bias[NS];
#pragma unroll
    for (int c = 0; c < NS; ++c) bias[c] = bn[c];

    int stage = 0;
    uint32_t phase = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        mbar_wait(full + stage, phase);
        const float* xs = reinterpret_cast<const float*>(ring + (size_t)stage * stage_bytes) +
                          (size_t)(warp * ROWS_PER_WARP + g) * d + kc;
        float4 x[U];
#pragma unroll
        for (int j = 0; j < U; ++j) x[j] = *reinterpret_cast<const float4*>(xs + (size_t)j * G * d);
        float acc[U][NS];
#pragma unroll
        for (int j = 0; j < U; ++j)
#pragma unroll
            for (int c = 0; c < NS; ++c) {
                float a = x[j].x * wreg[c][0];
                a = fmaf(x[j].y, wreg[c][1], a);
                a = fmaf(x[j].z, wreg[c][2], a);
                acc[j][c] = fmaf(x[j].w, wreg[c][3], a);
            }
        // the stage is free as soon as every lane of this warp holds its data in registers
        __syncwarp();
        if (lane == 0) mbar_arrive(empty + stage);
        if (++stage == stages) { stage = 0; phase ^= 1; }

        lane_reduce<LANES / 2, U, NS, U>(acc, s);
        const int sel = s / (LANES / M);
        const bool owner = (s % (LANES / M)) == 0;
        const int64_t base = tile * TILE_ROWS + warp * ROWS_PER_WARP;
        if (owner) {
#pragma unroll
            for (int r = 0; r < ROWS_OWNED; ++r) {
                const int u = sel * ROWS_OWNED + r;
                const int64_t row = base + u * G + g;
                if (row < n_rows) {
                    float z[NS];
#pragma unroll
                    for (int c = 0; c < NS; ++c) z[c] = acc[r][c] + bias[c];
                    finish_row<NS>(z, row, o);
                }
            }
        }
    }
}

// ---- rows that are not 16-byte aligned ---------------------------------------------------------------
// Odd feature counts (30, 33, 785 ...), odd pitches or an unaligned base leave the kernels above with
// scalar 4-byte loads: 2.3-2.7 TB/s measured at d = 30 / 33.  The ROWS may be unaligned, but a block of
// them is still one contiguous span of memory: the producer lane copies whole tiles of rows
// (TILE_ROWS x ldx floats, a multiple of 16 bytes) into a shared-memory ring with cp.async.bulk -- the
// 16-byte aligned middle of the span in one bulk copy, the up to three floats before and after it
// with plain loads, so nothing outside [X, X + n*ldx) is touched -- and the consumer warps (16, or 12 for 5..24 classes) read
// their rows back with scalar shared loads, which do not care about alignment.  Same lane mapping,
// cross-lane reduction and epilogue as above; W in shared memory.  Serves ldx <= kFlatMaxLdx.
constexpr int kFlatMaxLdx = 2048;
constexpr int kFlatSmemBudget = 200 * 1024;
// consumer warps: 16 (scalar shared loads were latency-bound with 8); 12 for the U x NS register tiles of
// 5..16 classes, which need more than the 96 registers 17 warps leave each thread
template <int NS>
struct FlatWarps { static constexpr int value = NS <= 4 ? 16 : 12; };

// AL = false with V = 4 (5..16 classes, unaligned rows): the lane mapping of mid_kernel<4, ..., AL = false>
// -- four consecutive floats per lane and step, read with four scalar shared loads clamped one by one at
// the row end, against 128-bit pieces of W.
template <int V, int LANES, int NS, int U, bool AL = true>
__global__ void __launch_bounds__((FlatWarps<NS>::value + 1) * 32, 1)
narrow_flat_kernel(const float* __restrict__ X, int64_t n_rows, int64_t ldx, int d, int dpad, int tile_rows,
                   int stages, uint32_t stage_bytes, const float* __restrict__ Wn, const float* __restrict__ bn,
                   Outs o) {
    constexpr int G = 32 / LANES;
    constexpr int M = (U < LANES) ? U : LANES;
    constexpr int ROWS_OWNED = U / M;
    constexpr int GROUP = G * U;                      // rows one warp takes per pass
    constexpr int kFlatWarps = FlatWarps<NS>::value;
    constexpr int kFlatThreads = (kFlatWarps + 1) * 32;

    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);          // [stages]
    uint64_t* empty = full + stages;                                  // [stages]
    float* w_s = reinterpret_cast<float*>(smem_raw + 128 * ((2 * stages * 8 + 127) / 128));   // [NS][dpad]
    unsigned char* ring = reinterpret_cast<unsigned char*>(w_s + NS * dpad);                  // dpad % 32 == 0

    if (threadIdx.x == 0) {
        for (int i = 0; i < stages; ++i) {
            mbar_init(full + i, 1);
            mbar_init(empty + i, kFlatWarps);
        }
        mbar_fence_init();
    }
    for (int i = threadIdx.x; i < NS * dpad; i += kFlatThreads) w_s[i] = Wn[i];
    __syncthreads();

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int64_t n_tiles = (n_rows + tile_rows - 1) / tile_rows;
    // tile_rows * ldx is a multiple of 4 floats, so every tile starts at the same offset inside a 16-byte line
    const uint32_t a0 = (uint32_t)((reinterpret_cast<uintptr_t>(X) >> 2) & 3u);

    if (warp == kFlatWarps) {
        // ===== producer: one lane feeds the ring =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                mbar_wait(empty + stage, phase ^ 1);
                const int64_t row0 = tile * tile_rows;
                const int64_t rows = (n_rows - row0 < tile_rows) ? (n_rows - row0) : tile_rows;
                // floats of this tile that exist: the last row of the matrix ends after d, not ldx
                const int64_t nfl = (row0 + rows == n_rows) ? (rows - 1) * ldx + d : rows * ldx;
                const float* src = X + row0 * ldx;
                float* dst = reinterpret_cast<float*>(ring + (size_t)stage * stage_bytes) + a0;   // tile float i -> dst[i]
                const int64_t head = nfl < (int64_t)((4 - a0) & 3u) ? nfl : (int64_t)((4 - a0) & 3u);
                const int64_t mid = (nfl - head) & ~int64_t(3);
                for (int64_t i = 0; i < head; ++i) dst[i] = src[i];
                for (int64_t i = head + mid; i < nfl; ++i) dst[i] = src[i];
                if (mid > 0) {
                    mbar_expect_tx(full + stage, (uint32_t)(mid * 4));     // release: the plain stores above are ordered before it
                    bulk_g2s(dst + head, src + head, (uint32_t)(mid * 4), full + stage);
                } else {
                    mbar_arrive(full + stage);
                }
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

    // ===== consumers =====
    const int g = lane / LANES;
    const int s = lane % LANES;
    constexpr int STEP = LANES * V;
    const int k0 = s * V;
    const int dloop = (d + STEP - 1) / STEP * STEP;          // <= dpad; W is zero beyond d
    float bias[NS];
#pragma unroll
    for (int c = 0; c < NS; ++c) bias[c] = bn[c];

    int stage = 0;
    uint32_t phase = 0;
    // A tile holds tile_rows / GROUP row groups -- fewer than there are warps when rows are long -- so
    // the groups of successive tiles go round the warps: group number (tile sequence * groups per tile
    // + index in tile) modulo kFlatWarps is the warp that takes it.
    const int gpt = tile_rows / GROUP;
    int first = warp;                                 // index in the tile of this warp's first group
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        mbar_wait(full + stage, phase);
        const float* ts = reinterpret_cast<const float*>(ring + (size_t)stage * stage_bytes) + a0;
        const int64_t row0 = tile * tile_rows;
        const int rows = (int)((n_rows - row0 < tile_rows) ? (n_rows - row0) : tile_rows);
        const int gb0 = first * GROUP;
        first = (first + kFlatWarps - gpt % kFlatWarps) % kFlatWarps;
        for (int gb = gb0; gb < rows; gb += kFlatWarps * GROUP) {
            const float* rp[U];
#pragma unroll
            for (int j = 0; j < U; ++j) {
                int r = gb + j * G + g;
                r = r < rows ? r : rows - 1;                  // clamped rows are computed and never stored
                rp[j] = ts + (int64_t)r * ldx;
            }
            float acc[U][NS];
#pragma unroll
            for (int j = 0; j < U; ++j)
#pragma unroll
                for (int c = 0; c < NS; ++c) acc[j][c] = 0.0f;
            // Same lanes, same pieces per lane and same order of FMAs as narrow_kernel<V, LANES, NS, U>:
            // a row scores to the same bits whichever of the two kernels its batch size selects.
            for (int k = k0; k < dloop; k += STEP) {
                const int kc = k < d ? k : 0;                 // past-the-end lanes meet zero weights
                float x[U][V];
#pragma unroll
                for (int j = 0; j < U; ++j) {
                    if constexpr (V == 4 && AL) {             // aligned rows: conflict-free 128-bit shared loads
                        const float4 t = *reinterpret_cast<const float4*>(rp[j] + kc);
                        x[j][0] = t.x; x[j][1] = t.y; x[j][2] = t.z; x[j][3] = t.w;
                    } else if constexpr (V == 4) {
#pragma unroll
                        for (int v = 0; v < 4; ++v) x[j][v] = rp[j][k + v < d ? k + v : 0];
                    } else {
                        x[j][0] = rp[j][kc];
                    }
                }
#pragma unroll
                for (int c = 0; c < NS; ++c) {
                    float w[V];
                    if constexpr (V == 4) {
                        const float4 t = *reinterpret_cast<const float4*>(w_s + c * dpad + k);
                        w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w;
                    } else {
                        w[0] = w_s[c * dpad + k];
                    }
#pragma unroll
                    for (int j = 0; j < U; ++j)
#pragma unroll
                        for (int v = 0; v < V; ++v) acc[j][c] = fmaf(x[j][v], w[v], acc[j][c]);
                }
            }
            lane_reduce<LANES / 2, U, NS, U>(acc, s);
            const int sel = s / (LANES / M);
            const bool owner = (s % (LANES / M)) == 0;
            if (owner) {
#pragma unroll
                for (int r = 0; r < ROWS_OWNED; ++r) {
                    const int u = sel * ROWS_OWNED + r;
                    const int rt = gb + u * G + g;
                    if (rt < rows) {
                        float z[NS];
#pragma unroll
                        for (int c = 0; c < NS; ++c) z[c] = acc[r][c] + bias[c];
                        finish_row<NS>(z, row0 + rt, o);
                    }
                }
            }
        }
        // every lane of this warp is done reading the stage
        __syncwarp();
        if (lane == 0) mbar_arrive(empty + stage);
        if (++stage == stages) { stage = 0; phase ^= 1; }
    }
}

// Resident CTAs per SM of one kernel instantiation for a given dynamic shared-memory size.  The size
// depends on the model (its padded W), not on the instantiation, so the answer is cached per kernel,
// device and size; the opt-in above 48 KB is made once per kernel and device.  (Keyed by the kernel's
// address: every instantiation has the same function-pointer type.)
template <typename K>
int resident_blocks(K kern, int dev, size_t smem, int* occ_out) {
    struct Slot { bool opted = false; size_t smem = ~size_t(0); int occ = 0; };
    static std::map<std::pair<const void*, int>, Slot> slots;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    Slot& sl = slots[{reinterpret_cast<const void*>(kern), dev}];
    if (!sl.opted) {
        LRS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWSmemMax));
        sl.opted = true;
    }
    if (sl.smem != smem) {
        int occ = 0;
        LRS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kThreads, smem));
        sl.occ = occ > 0 ? occ : 1;
        sl.smem = smem;
    }
    *occ_out = sl.occ;
    return LR_OK;
}

template <int NS>
constexpr int unroll_for() { return NS == 1 ? 8 : 4; }

template <int V, int LANES, int NS, bool SINGLE>
int launch_inst(const Model& m, const ScoreArgs& a) {
    constexpr int U = unroll_for<NS>();
    auto kern = narrow_kernel<V, LANES, NS, U, SINGLE>;
    const size_t smem = SINGLE ? 0 : (size_t)NS * m.narrow_dpad * sizeof(float);
    int dev = m.device & 63;
    int occ = 0;
    int rc = resident_blocks(kern, dev, smem, &occ);
    if (rc != LR_OK) return rc;
    constexpr int rows_per_iter = (32 / LANES) * U;
    const int64_t tiles = (a.n_rows + rows_per_iter - 1) / rows_per_iter;
    const int64_t need = (tiles + (kThreads / 32) - 1) / (kThreads / 32);
    const int64_t cap = (int64_t)m.num_sms * occ;
    const int grid = (int)(need < cap ? need : cap);
    Outs o{a.proba, a.decision, a.label, a.maxprob, m.status, m.n_scores};
    kern<<<grid, kThreads, smem, a.stream>>>(a.X, a.n_rows, a.ldx, m.n_features, m.narrow_dpad,
                                             m.Wn, m.bn, o);
    count_launch();
    LRS_CUDA(cudaGetLastError());
    return LR_OK;
}

template <int V, int LANES, int NS>
int launch_ns(const Model& m, const ScoreArgs& a, bool single) {
    if constexpr (V == 4 && NS <= 4) {
        if (single) return launch_inst<V, LANES, NS, true>(m, a);
    }
    return launch_inst<V, LANES, NS, false>(m, a);
}

template <int V, int LANES>
int launch_lanes(const Model& m, const ScoreArgs& a, bool single) {
    switch (m.narrow_ns) {
        case 1: return launch_ns<V, LANES, 1>(m, a, single);
        case 3: return launch_ns<V, LANES, 3>(m, a, single);
        case 4: return launch_ns<V, LANES, 4>(m, a, single);
    }
    return fail(LR_ERR_INVALID, "narrow path: unsupported score count");
}

template <int V, int LANES, int NS, bool WSMEM, bool AL>
int launch_mid_inst(const Model& m, const ScoreArgs& a) {
    auto kern = mid_kernel<V, LANES, NS, WSMEM, AL>;
    const size_t smem = WSMEM ? (size_t)NS * m.narrow_dpad * sizeof(float) : 0;
    const int dev = m.device & 63;
    int occ = 0;
    const int rc = resident_blocks(kern, dev, smem, &occ);
    if (rc != LR_OK) return rc;
    constexpr int rows_per_iter = (32 / LANES) * MidRows<NS>::value;
    const int64_t tiles = (a.n_rows + rows_per_iter - 1) / rows_per_iter;
    const int64_t need = (tiles + (kThreads / 32) - 1) / (kThreads / 32);
    const int64_t cap = (int64_t)m.num_sms * occ;
    const int grid = (int)(need < cap ? need : cap);
    Outs o{a.proba, a.decision, a.label, a.maxprob, m.status, m.n_scores};
    kern<<<grid, kThreads, smem, a.stream>>>(a.X, a.n_rows, a.ldx, m.n_features, m.narrow_dpad, m.Wn, m.bn, o);
    count_launch();
    LRS_CUDA(cudaGetLastError());
    return LR_OK;
}

template <int LANES, int NS, bool AL>
int launch_mid_w(const Model& m, const ScoreArgs& a) {
    const bool wsmem = (size_t)NS * m.narrow_dpad * sizeof(float) <= kWSmemMax;
    return wsmem ? launch_mid_inst<4, LANES, NS, true, AL>(m, a) : launch_mid_inst<4, LANES, NS, false, AL>(m, a);
}

// Lanes per row for the 5..16-class kernel.  Fewer lanes = more rows per warp instruction, and lanes
// of different rows read the same piece of W (one shared-memory wavefront instead of four at 32
// lanes: with 32 the W reads cost as many cycles as the FMAs they feed); more lanes = longer
// contiguous pieces of each row per request.  Measured (lanes 4, 8, 16, 32 swept): 4 lanes win up to d=256,
// 8 from there on.  AL = rows are 16-byte aligned (else the same mapping with 4-byte loads).
template <int NS, bool AL>
int launch_mid(const Model& m, const ScoreArgs& a) {
    const int chunks = (m.n_features + 3) / 4;
    return chunks <= 64 ? launch_mid_w<4, NS, AL>(m, a) : launch_mid_w<8, NS, AL>(m, a);
}

template <int V>
int launch_v(const Model& m, const ScoreArgs& a) {
    switch (m.narrow_ns) {
        case 6: return launch_mid<6, V == 4>(m, a);
        case 8: return launch_mid<8, V == 4>(m, a);
        case 10: return launch_mid<10, V == 4>(m, a);
        case 12: return launch_mid<12, V == 4>(m, a);
        case 16: return launch_mid<16, V == 4>(m, a);
        case 24: return launch_mid<24, V == 4>(m, a);
    }
    int chunks = (m.n_features + V - 1) / V;
    int lanes = 1;
    while (lanes < chunks && lanes < 32) lanes <<= 1;
    if (V == 1) lanes = scalar_lanes(m.n_features);
    const bool single = chunks <= lanes;
    switch (lanes) {
        case 1: return launch_lanes<V, 1>(m, a, single);
        case 2: return launch_lanes<V, 2>(m, a, single);
        case 4: return launch_lanes<V, 4>(m, a, single);
        case 8: return launch_lanes<V, 8>(m, a, single);
        case 16: return launch_lanes<V, 16>(m, a, single);
        default: return launch_lanes<V, 32>(m, a, single);
    }
}

template <int LANES, int NS>
int launch_bulk_inst(const Model& m, const ScoreArgs& a) {
    constexpr int U = unroll_for<NS>();
    constexpr int tile_rows = (32 / LANES) * U * kConsumerWarps;
    auto kern = narrow_bulk_kernel<LANES, NS, U>;
    const size_t stage_bytes = (size_t)tile_rows * m.n_features * sizeof(float);
    int stages = (int)((kBulkSmemBudget - 256) / stage_bytes);
    if (stages > 12) stages = 12;
    if (stages < 2) return -1;
    const size_t smem = 128 * ((2 * stages * 8 + 127) / 128) + stages * stage_bytes;
    static bool configured[64] = {false};
    const int dev = m.device & 63;
    if (!configured[dev]) {
        LRS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kBulkSmemBudget));
        configured[dev] = true;
    }
    const int64_t tiles = (a.n_rows + tile_rows - 1) / tile_rows;
    const int grid = (int)(tiles < m.num_sms ? tiles : m.num_sms);
    Outs o{a.proba, a.decision, a.label, a.maxprob, m.status, m.n_scores};
    kern<<<grid, kBulkThreads, smem, a.stream>>>(a.X, a.n_rows, m.n_features, m.narrow_dpad, stages,
                                                 m.Wn, m.bn, o);
    count_launch();
    LRS_CUDA(cudaGetLastError());
    return LR_OK;
}

template <int LANES>
int launch_bulk_lanes(const Model& m, const ScoreArgs& a) {
    switch (m.narrow_ns) {
        case 1: return launch_bulk_inst<LANES, 1>(m, a);
        case 3: return launch_bulk_inst<LANES, 3>(m, a);
        case 4: return launch_bulk_inst<LANES, 4>(m, a);
    }
    return -1;
}

// Returns -1 when the bulk variant does not apply (caller falls through to the plain-load kernel).
int launch_bulk(const Model& m, const ScoreArgs& a) {
    const int chunks = m.n_features / 4;
    int lanes = 1;
    while (lanes < chunks) lanes <<= 1;
    switch (lanes) {
        case 1: return launch_bulk_lanes<1>(m, a);
        case 2: return launch_bulk_lanes<2>(m, a);
        case 4: return launch_bulk_lanes<4>(m, a);
        case 8: return launch_bulk_lanes<8>(m, a);
        case 16: return launch_bulk_lanes<16>(m, a);
        case 32: return launch_bulk_lanes<32>(m, a);
    }
    return -1;
}

}  // namespace

// Score columns the model runs with on this path, 0 if it does not serve it.  <= 4 columns want W in
// shared memory (<= 96 KB, d <= 4096); a 2..4-class model with longer rows runs as a 6-column one on the
// register-tile kernel (W through L1 / L2, padding columns cost FMAs it has to spare); 5..16 columns take
// any W up to 16 MB; 17..24 columns with rows of at most 56 floats run here too (as 24
// columns, two rows per thread): the tensor-core path cannot go below one 128-row tile per k-chunk and
// measured 0.9-1.3 TB/s on d=54, C=20 / d=32, C=20, where plain FMAs reach 1.7-1.9.
// A binary model with long rows belongs to the wide path.
static int narrow_columns(const Model& m) {
    // (aligned rows: the tensor path's single-pass epilogue overtakes from d ~ 48: d=60, C=24 2.5 : 1.9 TB/s)
    if (m.n_scores > 24 || (m.n_scores > 16 && m.n_features > (m.n_features % 4 ? kMid24MaxD : 40))) return 0;
    int ns = padded_scores(m.n_scores);
    const int dpad = (m.n_features + 127) / 128 * 128;
    if (ns <= 4 && !((size_t)ns * dpad * sizeof(float) <= kWSmemMax && m.n_features <= 4096)) {
        if (ns == 1) return 0;
        ns = 6;
    }
    if (ns > 4 && (size_t)ns * dpad * sizeof(float) > kWL1Max) return 0;
    return ns;
}

bool narrow_supports(const Model& m) { return narrow_columns(m) != 0; }

int narrow_prepare(Model& m, const float* W_host, const float* b_host) {
    const int ns = narrow_columns(m);
    if (ns == 0) return LR_OK;
    // dpad covers whole lane-chunk sweeps for both V=4 and V=1 mappings (a multiple of 32*4).
    const int dpad = (m.n_features + 127) / 128 * 128;
    std::string w((size_t)ns * dpad * sizeof(float), '\0');
    float* wp = reinterpret_cast<float*>(&w[0]);
    float bp[24];
    for (int c = 0; c < ns; ++c) {
        bp[c] = c < m.n_scores ? b_host[c] : -INFINITY;   // padded classes never win, exp -> 0
        if (c < m.n_scores)
            for (int k = 0; k < m.n_features; ++k) wp[(size_t)c * dpad + k] = W_host[(size_t)c * m.n_features + k];
    }
    if (!m.Wn) {                                  // lr_model_update reloads into the same buffers
        LRS_CUDA(cudaMalloc(&m.Wn, w.size()));
        LRS_CUDA(cudaMalloc(&m.bn, ns * sizeof(float)));
    }
    LRS_CUDA(cudaMemcpy(m.Wn, wp, w.size(), cudaMemcpyHostToDevice));
    LRS_CUDA(cudaMemcpy(m.bn, bp, ns * sizeof(float), cudaMemcpyHostToDevice));
    m.narrow_ns = ns;
    m.narrow_dpad = dpad;
    if (const char* e = getenv("LRS_NARROW_BULK")) m.narrow_bulk = atoi(e) != 0;
    if (const char* e = getenv("LRS_NARROW_BULK_MIN_ROWS")) m.narrow_bulk_min_rows = atoll(e);
    return LR_OK;
}

namespace {

template <int V, int LANES, int NS, bool AL = true>
int launch_flat_inst(const Model& m, const ScoreArgs& a) {
    constexpr int U = NS <= 4 ? unroll_for<NS>() : MidRows<NS>::value;     // as the direct-load kernel of the shape
    constexpr int GROUP = (32 / LANES) * U;
    auto kern = narrow_flat_kernel<V, LANES, NS, U, AL>;
    // a tile = whole row groups (GROUP rows, a multiple of 4), about 32 KB of them
    const int64_t group_bytes = (int64_t)GROUP * a.ldx * 4;
    const int64_t gpt = group_bytes >= 32768 ? 1 : 32768 / group_bytes;
    const int tile_rows = (int)(GROUP * gpt);
    const uint32_t stage_bytes = (uint32_t)(((int64_t)tile_rows * a.ldx * 4 + 16 + 15) & ~int64_t(15));
    const size_t wbytes = (size_t)NS * m.narrow_dpad * sizeof(float);
    int stages = (int)((kFlatSmemBudget - 256 - (int64_t)wbytes) / stage_bytes);
    if (stages > 12) stages = 12;
    if (stages < 3) return -1;
    const size_t smem = 128 * ((2 * stages * 8 + 127) / 128) + wbytes + (size_t)stages * stage_bytes;
    {
        static std::map<std::pair<const void*, int>, bool> opted;
        static std::mutex mu;
        std::lock_guard<std::mutex> lock(mu);
        bool& done = opted[{reinterpret_cast<const void*>(kern), m.device}];
        if (!done) {
            LRS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kFlatSmemBudget));
            done = true;
        }
    }
    const int64_t tiles = (a.n_rows + tile_rows - 1) / tile_rows;
    const int grid = (int)(tiles < m.num_sms ? tiles : m.num_sms);
    Outs o{a.proba, a.decision, a.label, a.maxprob, m.status, m.n_scores};
    kern<<<grid, (FlatWarps<NS>::value + 1) * 32, smem, a.stream>>>(a.X, a.n_rows, a.ldx, m.n_features, m.narrow_dpad, tile_rows, stages,
                                                 stage_bytes, m.Wn, m.bn, o);
    count_launch();
    LRS_CUDA(cudaGetLastError());
    return LR_OK;
}

template <int V, int LANES>
int launch_flat_lanes(const Model& m, const ScoreArgs& a) {
    switch (m.narrow_ns) {
        case 1: return launch_flat_inst<V, LANES, 1>(m, a);
        case 3: return launch_flat_inst<V, LANES, 3>(m, a);
        case 4: return launch_flat_inst<V, LANES, 4>(m, a);
    }
    return -1;      // 5..16 classes: the scalar-load form of the register-tile kernel measured faster
}

template <int LANES>
int launch_flat_mid(const Model& m, const ScoreArgs& a) {
    switch (m.narrow_ns) {
        case 6: return launch_flat_inst<4, LANES, 6, false>(m, a);
        case 8: return launch_flat_inst<4, LANES, 8, false>(m, a);
        case 10: return launch_flat_inst<4, LANES, 10, false>(m, a);
        case 12: return launch_flat_inst<4, LANES, 12, false>(m, a);
        case 16: return launch_flat_inst<4, LANES, 16, false>(m, a);
        case 24: return launch_flat_inst<4, LANES, 24, false>(m, a);
    }
    return -1;
}

// Returns -1 when the flat-tile kernel does not apply (caller falls through to the plain-load kernels).
int launch_flat(const Model& m, const ScoreArgs& a, bool vec) {
    if (m.narrow_ns > 4) {
        // 5..16 classes: unaligned rows up to d = 256 only (measured: d=54, C=7 2.5 -> 4.8 TB/s, d=61, C=10
        // 2.5 -> 5.0; from d = 785 up the direct 4-byte loads are as fast or faster, and aligned rows never
        // gained from the ring).  Four lanes per row, as launch_mid picks for these sizes.
        if (vec || (m.n_features + 3) / 4 > 64) return -1;
        return launch_flat_mid<4>(m, a);
    }
    if (vec) {                                           // lanes as launch_v<4> picks them
        const int chunks = m.n_features / 4;
        if (chunks <= 8) return launch_flat_lanes<4, 8>(m, a);
        if (chunks <= 16) return launch_flat_lanes<4, 16>(m, a);
        return launch_flat_lanes<4, 32>(m, a);
    }
    switch (scalar_lanes(m.n_features)) {
        case 4: return launch_flat_lanes<1, 4>(m, a);
        case 8: return launch_flat_lanes<1, 8>(m, a);
        case 16: return launch_flat_lanes<1, 16>(m, a);
        default: return launch_flat_lanes<1, 32>(m, a);
    }
}

}  // namespace

int narrow_launch(const Model& m, const ScoreArgs& a) {
    if (!m.Wn) return fail(LR_ERR_INVALID, "narrow path does not serve this model shape");
    const bool vec = (m.n_features % 4 == 0) && (a.ldx % 4 == 0) &&
                     ((reinterpret_cast<uintptr_t>(a.X) & 15u) == 0);
    // Opt-in: large dense batches through the bulk-copy ring (see narrow_bulk_kernel).
    if (vec && m.narrow_bulk && a.ldx == m.n_features && m.n_features <= 128 && a.n_rows >= m.narrow_bulk_min_rows) {
        const int rc = launch_bulk(m, a);
        if (rc >= 0) return rc;
    }
    // Rows that are not 16-byte aligned: whole tiles of rows through shared memory (see narrow_flat_kernel);
    // batches of a few rows keep the plain scalar-load kernel (one launch, no ring to fill).
    // Aligned rows of 1 KB and more do better through the same ring with 128-bit shared loads (measured,
    // binary: d=256 6.05 -> 6.68 TB/s, d=1000 5.96 -> 7.06; shorter rows are faster on direct loads).
    static const bool flat_on = [] { const char* e = getenv("LRS_NARROW_FLAT"); return !e || atoi(e) != 0; }();
    static const int flat_min_d = [] { const char* e = getenv("LRS_NARROW_FLAT_MIN_D"); return e ? atoi(e) : 256; }();
    if (flat_on && (!vec || (m.narrow_ns <= 4 && m.n_features >= flat_min_d && m.n_features >= 20)) && a.ldx <= kFlatMaxLdx &&
        a.n_rows >= 4096) {
        const int rc = launch_flat(m, a, vec);
        if (rc >= 0) return rc;
    }
    return vec ? launch_v<4>(m, a) : launch_v<1>(m, a);
}

}  // namespace lrs
