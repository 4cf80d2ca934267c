// Refit row: the loss and gradient that scikit-learn's lbfgs solver
// evaluates each iteration of `LogisticRegression().fit` (REF/Logistic Regression.ipynb:33-34 ->
// SK/linear_model/_logistic.py:584-598 -> LinearModelLoss.loss_gradient,
// SK/linear_model/_linear_loss.py:291-378), as ONE fused pass over X:
//     z = X.W^T + b ;  r = sigmoid(z) - y   (binary, SK/_loss/_loss.pyx.tp:728-751)
//                      r = softmax(z) - onehot(y)   (multinomial, SK/_loss/_loss.pyx.tp:269,1202)
//     loss_sum = sum_i loss_i ;  grad_W = r^T X ;  grad_b = sum_i r_i
// The kernel returns UNNORMALISED sums over the rows it was given (float64): a rank scores its row
// shard, the ranks all-reduce the one [loss | grad_W | grad_b] buffer (the only collective in the
// design), and the host divides by n and adds the L2 term (intercept unpenalised).
//
// Mapping: as in the narrow scoring kernel, LANES lanes share a row and each owns one 4-float (or
// 1-float) chunk of it, which it keeps in registers for both the dot product and the outer product
// r.x^T, so X is read once.  Per-lane gradient accumulators are summed across the CTA through shared
// memory, each CTA writes one partial vector to scratch, and the last CTA to finish (atomic ticket)
// adds the partials in a fixed order in float64 -- deterministic, no float atomics.
//
// HBM roofline: 4*d + 4 bytes per row (X and the int32 label); everything else is O(grid).
#include <cmath>
#include <mutex>

#include "common.cuh"

namespace lrs {
namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kU = 4;                       // rows in flight per lane group

struct RefitArgs {
    const float* X;
    const int32_t* y;
    int64_t n_rows;
    int64_t ldx;
    const float* W;      // [n_scores][d]
    const float* b;      // [n_scores]
    int n_scores;
    int d;
    double* out;         // [1 + n_scores*d + n_scores]
    float* partials;     // scratch: [grid][stride]
    unsigned int* ticket;  // scratch: arrival counter, zero on entry, reset on exit
    int stride;          // 1 + n_scores*(d+1), padded to 4
};

__device__ __forceinline__ float log1pexp_f(float z) {
    // log(1 + exp(z)), stable on both sides (same branches as SK/_loss/_loss.pyx.tp:728-751 merged)
    return z > 0.0f ? z + log1pf(expf(-z)) : log1pf(expf(z));
}

// d in (16, 32] binary -- the cfg-1 shape -- takes a leaner body: 8 rows per lane group, the partial
// dot products are transpose-reduced so that each of the 8 lanes evaluates the sigmoid / log-loss of
// ONE row (instead of all 8 lanes repeating it for every row), then the residuals go round the group
// with 8 shuffles for the outer product.
template <int V, int LANES, int NS>
struct RefitShape {
    static constexpr bool lean = V == 4 && LANES == 8 && NS == 1;
    static constexpr int rows = lean ? 8 : kU;      // rows per lane group per iteration
};

template <int V, int LANES, int NS>
__global__ void __launch_bounds__(kThreads)
refit_kernel(const RefitArgs a) {
    constexpr int G = 32 / LANES;
    constexpr bool LEAN = RefitShape<V, LANES, NS>::lean;
    constexpr int ROWS_PER_ITER = G * RefitShape<V, LANES, NS>::rows;
    extern __shared__ float red[];                 // [kWarps][stride]
    __shared__ bool is_last;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane / LANES, s = lane % LANES;
    const int k0 = s * V;
    const bool active = k0 < a.d;                  // lanes past the end of a row carry zero weights
    const int kc = active ? k0 : 0;

    float w[NS][V];
    float bias[NS];
#pragma unroll
    for (int c = 0; c < NS; ++c) {
        bias[c] = c < a.n_scores ? a.b[c] : -INFINITY;
#pragma unroll
        for (int v = 0; v < V; ++v)
            w[c][v] = (c < a.n_scores && active && k0 + v < a.d) ? a.W[(size_t)c * a.d + k0 + v] : 0.0f;
    }
    float gw[NS][V];
    float gb[NS];
    float loss = 0.0f;
#pragma unroll
    for (int c = 0; c < NS; ++c) {
        gb[c] = 0.0f;
#pragma unroll
        for (int v = 0; v < V; ++v) gw[c][v] = 0.0f;
    }

    const int64_t warps_total = (int64_t)gridDim.x * kWarps;
    const int64_t n_tiles = (a.n_rows + ROWS_PER_ITER - 1) / ROWS_PER_ITER;
    for (int64_t tile = (int64_t)blockIdx.x * kWarps + warp; tile < n_tiles; tile += warps_total) {
        const int64_t base = tile * ROWS_PER_ITER;
        if constexpr (LEAN) {
            float4 x[8];
            float acc[8][1];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                int64_t row = base + j * G + g;
                row = row < a.n_rows ? row : a.n_rows - 1;
                x[j] = ldg_stream4(a.X + row * a.ldx + kc);
            }
            const int64_t my_row = base + s * G + g;             // the row this lane will own
            const bool mine = my_row < a.n_rows;
            const float yy = mine ? (float)__ldg(a.y + my_row) : 0.0f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float t = x[j].x * w[0][0];
                t = fmaf(x[j].y, w[0][1], t);
                t = fmaf(x[j].z, w[0][2], t);
                acc[j][0] = fmaf(x[j].w, w[0][3], t);
            }
            lane_reduce<4, 8, 1, 8>(acc, s);                      // acc[0][0] = full dot product of row s
            const float z = acc[0][0] + bias[0];
            const float e = expf(-fabsf(z));
            const float big = __frcp_rn(1.0f + e);
            const float pr = z > 0.0f ? big : e * big;
            const float r = mine ? pr - yy : 0.0f;
            if (mine) {
                loss += log1pexp_f(z) - yy * z;
                gb[0] += r;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float rj = __shfl_sync(0xffffffffu, r, (lane & ~7) | j);
                gw[0][0] = fmaf(rj, x[j].x, gw[0][0]);
                gw[0][1] = fmaf(rj, x[j].y, gw[0][1]);
                gw[0][2] = fmaf(rj, x[j].z, gw[0][2]);
                gw[0][3] = fmaf(rj, x[j].w, gw[0][3]);
            }
            continue;
        }
        float x[kU][V];
        int yv[kU];
        bool live[kU];
#pragma unroll
        for (int j = 0; j < kU; ++j) {
            int64_t row = base + j * G + g;
            live[j] = row < a.n_rows;
            row = live[j] ? row : a.n_rows - 1;
            const float* p = a.X + row * a.ldx + kc;
            if constexpr (V == 4) {
                const float4 t = ldg_stream4(p);
                x[j][0] = t.x; x[j][1] = t.y; x[j][2] = t.z; x[j][3] = t.w;
            } else {
                x[j][0] = ldg_stream1(p);
            }
            yv[j] = __ldg(a.y + row);
        }
#pragma unroll
        for (int j = 0; j < kU; ++j) {
            float z[NS];
#pragma unroll
            for (int c = 0; c < NS; ++c) {
                float acc = 0.0f;
#pragma unroll
                for (int v = 0; v < V; ++v) acc = fmaf(x[j][v], w[c][v], acc);
#pragma unroll
                for (int off = LANES / 2; off >= 1; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
                z[c] = acc + bias[c];               // every lane of the group holds the full score
            }
            float r[NS];
            float li;
            if constexpr (NS == 1) {
                const float yy = (float)yv[j];
                const float e = expf(-fabsf(z[0]));
                const float big = __frcp_rn(1.0f + e);
                const float pr = z[0] > 0.0f ? big : e * big;
                r[0] = pr - yy;
                li = log1pexp_f(z[0]) - yy * z[0];
            } else {
                float zmax = z[0];
#pragma unroll
                for (int c = 1; c < NS; ++c) zmax = fmaxf(zmax, z[c]);
                float sum = 0.0f, zy = 0.0f;
#pragma unroll
                for (int c = 0; c < NS; ++c) {
                    r[c] = expf(z[c] - zmax);
                    sum += r[c];
                    if (c == yv[j]) zy = z[c];
                }
                const float inv = __frcp_rn(sum);
#pragma unroll
                for (int c = 0; c < NS; ++c) r[c] = r[c] * inv - (c == yv[j] ? 1.0f : 0.0f);
                li = logf(sum) + zmax - zy;
            }
            if (live[j]) {
#pragma unroll
                for (int c = 0; c < NS; ++c) {
#pragma unroll
                    for (int v = 0; v < V; ++v) gw[c][v] = fmaf(r[c], x[j][v], gw[c][v]);
                    if (s == 0) gb[c] += r[c];
                }
                if (s == 0) loss += li;
            }
        }
    }

    if constexpr (LEAN) {   // loss and intercept gradient were accumulated per owning lane: fold the group
#pragma unroll
        for (int off = 4; off >= 1; off >>= 1) {
            loss += __shfl_xor_sync(0xffffffffu, loss, off);
            gb[0] += __shfl_xor_sync(0xffffffffu, gb[0], off);
        }
    }
    // ---- CTA reduction: lanes that own the same chunk (same s) across groups, then across warps ----
#pragma unroll
    for (int off = 16; off >= LANES; off >>= 1) {
#pragma unroll
        for (int c = 0; c < NS; ++c) {
#pragma unroll
            for (int v = 0; v < V; ++v) gw[c][v] += __shfl_xor_sync(0xffffffffu, gw[c][v], off);
            gb[c] += __shfl_xor_sync(0xffffffffu, gb[c], off);
        }
        loss += __shfl_xor_sync(0xffffffffu, loss, off);
    }
    float* mine = red + (size_t)warp * a.stride;
    for (int i = lane; i < a.stride; i += 32) mine[i] = 0.0f;
    __syncwarp();
    if (g == 0) {                                   // lanes 0..LANES-1 now hold the warp's sums
        if (s == 0) {
            mine[0] = loss;
#pragma unroll
            for (int c = 0; c < NS; ++c)
                if (c < a.n_scores) mine[1 + a.n_scores * a.d + c] = gb[c];
        }
        if (active) {
#pragma unroll
            for (int c = 0; c < NS; ++c)
#pragma unroll
                for (int v = 0; v < V; ++v)
                    if (c < a.n_scores && k0 + v < a.d) mine[1 + c * a.d + k0 + v] = gw[c][v];
        }
    }
    __syncthreads();
    float* part = a.partials + (size_t)blockIdx.x * a.stride;
    for (int i = threadIdx.x; i < a.stride; i += kThreads) {
        float t = 0.0f;
#pragma unroll
        for (int wdx = 0; wdx < kWarps; ++wdx) t += red[(size_t)wdx * a.stride + i];
        part[i] = t;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = atomicAdd(a.ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (is_last) {
        __threadfence();
        const int n_out = 1 + a.n_scores * (a.d + 1);
        for (int i = threadIdx.x; i < n_out; i += kThreads) {
            double t = 0.0;
            for (unsigned blk = 0; blk < gridDim.x; ++blk) t += (double)a.partials[(size_t)blk * a.stride + i];
            a.out[i] = t;
        }
        if (threadIdx.x == 0) *a.ticket = 0u;       // ready for the next call on this scratch
    }
}

// Scratch owned by the library: one growing buffer per device ([ticket | partials]).
struct Scratch {
    std::mutex mu;
    void* ptr[64] = {};
    size_t bytes[64] = {};
};
Scratch g_scratch;

template <int V, int LANES, int NS>
int launch_inst(RefitArgs a, int num_sms, size_t scratch_floats_cap, cudaStream_t stream) {
    auto kern = refit_kernel<V, LANES, NS>;
    const size_t smem = (size_t)kWarps * a.stride * sizeof(float);
    if (smem > 48 * 1024) LRS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    LRS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kThreads, smem));
    if (occ < 1) occ = 1;
    constexpr int rows_per_iter = (32 / LANES) * RefitShape<V, LANES, NS>::rows;
    const int64_t tiles = (a.n_rows + rows_per_iter - 1) / rows_per_iter;
    int64_t grid = (tiles + kWarps - 1) / kWarps;
    const int64_t cap = (int64_t)num_sms * occ;
    if (grid > cap) grid = cap;
    const int64_t fit = (int64_t)(scratch_floats_cap / (size_t)a.stride);
    if (grid > fit) grid = fit;                     // fewer CTAs, each strides over more tiles
    if (grid < 1) return fail(LR_ERR_LIMIT, "refit: scratch cannot hold one partial gradient vector");
    kern<<<(unsigned)grid, kThreads, smem, stream>>>(a);
    count_launch();
    LRS_CUDA(cudaGetLastError());
    return LR_OK;
}

template <int V, int LANES>
int launch_ns(const RefitArgs& a, int num_sms, size_t cap, cudaStream_t stream) {
    if (a.n_scores == 1) return launch_inst<V, LANES, 1>(a, num_sms, cap, stream);
    if (a.n_scores <= 3) return launch_inst<V, LANES, 3>(a, num_sms, cap, stream);
    if (a.n_scores <= 4) return launch_inst<V, LANES, 4>(a, num_sms, cap, stream);
    if (a.n_scores <= 8) return launch_inst<V, LANES, 8>(a, num_sms, cap, stream);
    return fail(LR_ERR_LIMIT, "refit kernel serves at most 8 score columns");
}

template <int V>
int launch_v(const RefitArgs& a, int num_sms, size_t cap, cudaStream_t stream) {
    const int chunks = (a.d + V - 1) / V;
    if (chunks > 32)
        return fail(LR_ERR_LIMIT, "refit kernel serves n_features <= 128 (16-byte aligned rows) or <= 32 (any rows)");
    int lanes = 1;
    while (lanes < chunks) lanes <<= 1;
    switch (lanes) {
        case 1: return launch_ns<V, 1>(a, num_sms, cap, stream);
        case 2: return launch_ns<V, 2>(a, num_sms, cap, stream);
        case 4: return launch_ns<V, 4>(a, num_sms, cap, stream);
        case 8: return launch_ns<V, 8>(a, num_sms, cap, stream);
        case 16: return launch_ns<V, 16>(a, num_sms, cap, stream);
        default: return launch_ns<V, 32>(a, num_sms, cap, stream);
    }
}

int check_refit_args(const float* X, const int32_t* y, int64_t n_rows, int64_t ldx, const float* W, const float* b,
                     int n_scores, int d, double* out) {
    if (!X || !y || !W || !b || !out) return fail(LR_ERR_INVALID, "refit: NULL pointer argument");
    if (n_rows < 1) return fail(LR_ERR_INVALID, "refit: n_rows must be >= 1");
    if (n_scores < 1 || d < 1) return fail(LR_ERR_INVALID, "refit: n_scores and n_features must be >= 1");
    if (ldx < d) return fail(LR_ERR_INVALID, "refit: row pitch ldx is smaller than n_features");
    return LR_OK;
}

int run_refit(RefitArgs a, void* scratch, size_t scratch_bytes, cudaStream_t stream) {
    int dev = 0, num_sms = 148;
    LRS_CUDA(cudaGetDevice(&dev));
    LRS_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
    a.stride = (1 + a.n_scores * (a.d + 1) + 3) / 4 * 4;
    a.ticket = reinterpret_cast<unsigned int*>(scratch);
    a.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
    const size_t cap_floats = (scratch_bytes - 256) / sizeof(float);
    const bool vec = (a.d % 4 == 0) && (a.ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.X) & 15u) == 0);
    return vec ? launch_v<4>(a, num_sms, cap_floats, stream) : launch_v<1>(a, num_sms, cap_floats, stream);
}

}  // namespace
}  // namespace lrs

using namespace lrs;

extern "C" int lr_loss_grad(const float* X_dev, const int32_t* y_dev, int64_t n_rows, int64_t ldx,
                            const float* W_dev, const float* b_dev, int32_t n_scores, int32_t n_features,
                            double* out_dev, void* stream) {
    int rc = check_refit_args(X_dev, y_dev, n_rows, ldx, W_dev, b_dev, n_scores, n_features, out_dev);
    if (rc != LR_OK) return rc;
    // the shape limits come first: nothing is allocated (and kept) for a call that is then refused
    if (n_scores > 8) return fail(LR_ERR_LIMIT, "refit kernel serves at most 8 score columns");
    {
        const bool vec = (n_features % 4 == 0) && (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(X_dev) & 15u) == 0);
        if (n_features > (vec ? 128 : 32))
            return fail(LR_ERR_LIMIT, "refit kernel serves n_features <= 128 (16-byte aligned rows) or <= 32 (any rows)");
    }
    int dev = 0, prev = -1;
    LRS_CUDA(cudaGetDevice(&dev));
    prev = dev;
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, X_dev) == cudaSuccess && attr.type == cudaMemoryTypeDevice && attr.device != dev) {
        LRS_CUDA(cudaSetDevice(attr.device));
        dev = attr.device;
    }
    struct Restore { int prev, cur; ~Restore() { if (prev != cur) cudaSetDevice(prev); } } restore{prev, dev};
    // scratch: ticket word + one partial vector per CTA (at most SMs x 8 CTAs), grown on demand
    const size_t stride = ((size_t)1 + (size_t)n_scores * ((size_t)n_features + 1) + 3) / 4 * 4;
    const size_t want = 256 + stride * sizeof(float) * 148 * 8;
    void* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_scratch.mu);
        const int di = dev & 63;
        if (g_scratch.bytes[di] < want) {
            if (g_scratch.ptr[di]) LRS_CUDA(cudaFree(g_scratch.ptr[di]));
            g_scratch.ptr[di] = nullptr;
            g_scratch.bytes[di] = 0;
            LRS_CUDA(cudaMalloc(&g_scratch.ptr[di], want));
            LRS_CUDA(cudaMemset(g_scratch.ptr[di], 0, 256));
            g_scratch.bytes[di] = want;
        }
        scratch = g_scratch.ptr[di];
    }
    RefitArgs a{X_dev, y_dev, n_rows, ldx, W_dev, b_dev, n_scores, n_features, out_dev, nullptr, nullptr, 0};
    return run_refit(a, scratch, want, reinterpret_cast<cudaStream_t>(stream));
}

// ---- lr_loss_grad_static: the same reduction with no allocation anywhere ---------------------------
//
// Who this is for.  lr_loss_grad above grows a cudaMalloc'ed scratch buffer the first time it runs
// on a device.  Some hosts cannot afford that call: a serving process that hands all of its device
// memory to its own pool at start-up (a later cudaMalloc can simply fail there), a refit step that
// runs beside latency-sensitive scoring streams (cudaMalloc/cudaFree may stall the whole device while
// they run), or an embedding runtime whose rules forbid third-party code from touching the device
// heap.  For those callers the per-CTA partial vectors live in a fixed array that is part of the
// library image: every GPU that loads liblrscore holds kStaticScratchBytes (about 0.6 MB) of it
// until the process exits, whether this entry point is used or not.
//
// Limits that follow from the fixed size: one partial vector [loss | grad_W | grad_b] may hold at
// most kStaticMaxVector floats, and the array has room for kStaticMaxCtas such vectors (two CTAs per
// SM on a B200; shorter vectors leave room for more CTAs, and when fewer CTAs fit than the GPU could
// run, each one strides over more row tiles).  Larger problems are refused with LR_ERR_LIMIT; use
// lr_loss_grad for them.
//
// Sharing.  There is ONE array per device and every call uses it, so it still holds the previous
// call's partial sums and arrival counter when the next call starts.  The kernel overwrites every
// partial slot it later reads, and the counter -- the one word that has to start at zero -- is
// cleared on the caller's stream before each launch.  Two calls whose GPU work overlaps (different
// streams, or different host threads without ordering) would write the same slots and race: keep
// calls on one device ordered one after the other.
constexpr int kStaticMaxCtas = 296;
constexpr int kStaticMaxVector = 512;
constexpr size_t kStaticScratchBytes = 256 + (size_t)kStaticMaxCtas * kStaticMaxVector * sizeof(float);

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

namespace lrs {
namespace {
// Device address of g_static_scratch, per device.  In host code the symbol's name is only a
// stand-in; the address a kernel may use comes from cudaGetSymbolAddress, asked once per device
// while that device is current.
std::mutex g_static_mu;
void* g_static_addr[64] = {};
}  // namespace
}  // namespace lrs

extern "C" int lr_loss_grad_static(const float* X_dev, const int32_t* y_dev, int64_t n_rows, int64_t ldx,
                                   const float* W_dev, const float* b_dev, int32_t n_scores,
                                   int32_t n_features, double* out_dev, void* stream) {
    int rc = check_refit_args(X_dev, y_dev, n_rows, ldx, W_dev, b_dev, n_scores, n_features, out_dev);
    if (rc != LR_OK) return rc;
    const int64_t vec = 1 + (int64_t)n_scores * ((int64_t)n_features + 1);
    if (vec > kStaticMaxVector)
        return fail(LR_ERR_LIMIT, "lr_loss_grad_static: one partial vector needs " + std::to_string(vec) +
                                      " floats (1 + n_scores*(n_features+1)) but the built-in scratch holds at most " +
                                      std::to_string(kStaticMaxVector) + "; use lr_loss_grad for this shape");
    int dev = 0;
    LRS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(LR_ERR_LIMIT, "lr_loss_grad_static: device index above 63");
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, X_dev) == cudaSuccess && attr.type == cudaMemoryTypeDevice && attr.device != dev)
        return fail(LR_ERR_INVALID, "lr_loss_grad_static: X_dev lives on device " + std::to_string(attr.device) +
                                        " but device " + std::to_string(dev) + " is current");
    void* scratch = nullptr;
    {
        // One-time set-up for this device (the first call here should happen before any stream
        // capture is started): look up where the array lives on the current device.
        std::lock_guard<std::mutex> lock(g_static_mu);
        if (!g_static_addr[dev]) {
            void* p = nullptr;
            LRS_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
            g_static_addr[dev] = p;
        }
        scratch = g_static_addr[dev];
    }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    LRS_CUDA(cudaMemsetAsync(scratch, 0, 256, s));          // the arrival counter must start at zero
    RefitArgs a{X_dev, y_dev, n_rows, ldx, W_dev, b_dev, n_scores, n_features, out_dev, nullptr, nullptr, 0};
    return run_refit(a, scratch, kStaticScratchBytes, s);
}

// ---- lr_weighted_row_sum: out[k] = sum_i r[i] * X[i][k] -----------------------------------------------
// The transposed product of a binary model's gradient, `X.T @ grad_pointwise`
// (SK/linear_model/_linear_loss.py:353-357), for feature counts beyond lr_loss_grad: the forward pass of
// such a model (d = 100k in the BASELINE config) is the wide scoring kernel, and the residual r is known
// only after a whole row has been read, so X is streamed a second time here.  HBM-bound: 4 bytes per
// element of X, nothing else of size.  Grid = column strips (256 threads x one 128-bit piece) x row
// blocks; each CTA keeps its strip's sums in registers, eight rows in flight per thread; per-block partial
// vectors are then added in float64 in a fixed order (deterministic).
namespace lrs {
namespace {

template <int V>
__global__ void __launch_bounds__(256, 2)
wsum_kernel(const float* __restrict__ X, const float* __restrict__ r, int64_t n_rows, int64_t ldx, int d,
            int64_t rows_per_block, float* __restrict__ partial) {
    constexpr int RB = 16;                             // rows in flight per thread (8: 5.2 TB/s)
    const int k = (blockIdx.x * 256 + threadIdx.x) * V;
    const int64_t r0 = (int64_t)blockIdx.y * rows_per_block;
    const int64_t r1 = r0 + rows_per_block < n_rows ? r0 + rows_per_block : n_rows;
    float acc[V];
#pragma unroll
    for (int v = 0; v < V; ++v) acc[v] = 0.0f;
    if (k < d) {                                       // V == 4: d % 4 == 0, the piece is whole
        const float* p = X + r0 * ldx + k;
        int64_t i = r0;
        for (; i + RB <= r1; i += RB, p += RB * ldx) {
            float x[RB][V], w[RB];
#pragma unroll
            for (int j = 0; j < RB; ++j) {
                if constexpr (V == 4) {
                    const float4 t = ldg_stream4(p + j * ldx);
                    x[j][0] = t.x; x[j][1] = t.y; x[j][2] = t.z; x[j][3] = t.w;
                } else {
                    x[j][0] = ldg_stream1(p + j * ldx);
                }
                w[j] = __ldg(r + i + j);
            }
#pragma unroll
            for (int j = 0; j < RB; ++j)
#pragma unroll
                for (int v = 0; v < V; ++v) acc[v] = fmaf(w[j], x[j][v], acc[v]);
        }
        for (; i < r1; ++i, p += ldx) {
            const float w = __ldg(r + i);
            if constexpr (V == 4) {
                const float4 t = ldg_stream4(p);
                acc[0] = fmaf(w, t.x, acc[0]); acc[1] = fmaf(w, t.y, acc[1]);
                acc[2] = fmaf(w, t.z, acc[2]); acc[3] = fmaf(w, t.w, acc[3]);
            } else {
                acc[0] = fmaf(w, ldg_stream1(p), acc[0]);
            }
        }
        float* o = partial + (size_t)blockIdx.y * d + k;
#pragma unroll
        for (int v = 0; v < V; ++v) o[v] = acc[v];
    }
}

__global__ void __launch_bounds__(256)
wsum_reduce_kernel(const float* __restrict__ partial, int n_blocks, int d, double* __restrict__ out) {
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= d) return;
    double s = 0.0;
    for (int b = 0; b < n_blocks; ++b) s += (double)partial[(size_t)b * d + k];
    out[k] = s;
}

}  // namespace
}  // namespace lrs

extern "C" int lr_weighted_row_sum(const float* X_dev, const float* r_dev, int64_t n_rows, int64_t ldx,
                                   int32_t n_features, double* out_dev, void* stream) {
    using namespace lrs;
    if (!X_dev || !r_dev || !out_dev) return fail(LR_ERR_INVALID, "lr_weighted_row_sum: null pointer");
    if (n_rows < 0 || n_features < 1 || ldx < n_features) return fail(LR_ERR_INVALID, "lr_weighted_row_sum: bad shape");
    int dev = 0;
    LRS_CUDA(cudaGetDevice(&dev));
    const int prev = dev;
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, X_dev) == cudaSuccess && attr.type == cudaMemoryTypeDevice && attr.device != dev) {
        LRS_CUDA(cudaSetDevice(attr.device));
        dev = attr.device;
    }
    struct Restore { int prev, cur; ~Restore() { if (prev != cur) cudaSetDevice(prev); } } restore{prev, dev};   // the caller's device
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int d = n_features;
    if (n_rows == 0) {
        LRS_CUDA(cudaMemsetAsync(out_dev, 0, (size_t)d * sizeof(double), s));
        return LR_OK;
    }
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const bool vec = d % 4 == 0 && ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(X_dev) & 15u) == 0;
    const int per = vec ? 1024 : 256;                      // columns per strip
    const int strips = (d + per - 1) / per;
    // enough row blocks for ~6 CTAs per SM, at least 64 rows each
    int64_t blocks = ((int64_t)sms * 6 + strips - 1) / strips;
    if (blocks > (n_rows + 63) / 64) blocks = (n_rows + 63) / 64;
    if (blocks < 1) blocks = 1;
    if (blocks > 65535) blocks = 65535;
    const int64_t rows_per_block = (n_rows + blocks - 1) / blocks;
    blocks = (n_rows + rows_per_block - 1) / rows_per_block;
    float* partial = nullptr;
    LRS_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&partial), (size_t)blocks * d * sizeof(float), s));
    const dim3 grid((unsigned)strips, (unsigned)blocks);
    if (vec) wsum_kernel<4><<<grid, 256, 0, s>>>(X_dev, r_dev, n_rows, ldx, d, rows_per_block, partial);
    else wsum_kernel<1><<<grid, 256, 0, s>>>(X_dev, r_dev, n_rows, ldx, d, rows_per_block, partial);
    count_launch();
    wsum_reduce_kernel<<<(d + 255) / 256, 256, 0, s>>>(partial, (int)blocks, d, out_dev);
    count_launch();
    LRS_CUDA(cudaGetLastError());
    LRS_CUDA(cudaFreeAsync(partial, s));
    return LR_OK;
}
