// C ABI of liblrscore (see include/lrscore.h): model loader, path dispatch, device and host entry
// points.  Everything here is host code; the kernels live in narrow.cu / wide.cu / generic.cu /
// gemm_tcgen05.cu.
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"

namespace lrs {

static thread_local std::string t_last_error;
std::atomic<uint64_t> g_launches{0};

void set_error(const std::string& msg) { t_last_error = msg; }
int fail(int code, const std::string& msg) {
    t_last_error = msg;
    return code;
}

// Keeps the caller's current device intact across a call.
struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; }
        if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard() {
        int cur = -1;
        if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
    }
};

// A few host threads that copy between pageable and pinned memory together: one thread moves ~8-10 GB/s, the
// PCIe link 55.  copy2d blocks until all rows are copied; the caller takes a share itself.
class CopyPool {
public:
    explicit CopyPool(int workers) {
        for (int i = 0; i < workers; ++i) th_.emplace_back([this, i] { loop(i); });
    }
    ~CopyPool() {
        { std::lock_guard<std::mutex> l(mu_); stop_ = true; ++gen_; }
        cv_.notify_all();
        for (auto& t : th_) t.join();
    }
    void copy2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, int64_t rows) {
        if (rows <= 0 || width == 0) return;
        const int parts = (int)th_.size() + 1;
        if ((size_t)rows * width < ((size_t)1 << 20) || th_.empty()) { slice(dst, dpitch, src, spitch, width, rows, 0, 1); return; }
        {
            std::lock_guard<std::mutex> l(mu_);
            job_ = {static_cast<char*>(dst), static_cast<const char*>(src), dpitch, spitch, width, rows, parts, false};
            left_ = (int)th_.size();
            ++gen_;
        }
        cv_.notify_all();
        slice(dst, dpitch, src, spitch, width, rows, parts - 1, parts);
        std::unique_lock<std::mutex> l(mu_);
        done_.wait(l, [this] { return left_ == 0; });
    }
    // rows of doubles -> rows of floats (pitches and width in ELEMENTS), the same way
    void convert2d(float* dst, size_t dpitch, const double* src, size_t spitch, size_t width, int64_t rows) {
        if (rows <= 0 || width == 0) return;
        const int parts = (int)th_.size() + 1;
        if ((size_t)rows * width < ((size_t)1 << 18) || th_.empty()) { cslice(dst, dpitch, src, spitch, width, rows, 0, 1); return; }
        {
            std::lock_guard<std::mutex> l(mu_);
            job_ = {reinterpret_cast<char*>(dst), reinterpret_cast<const char*>(src), dpitch, spitch, width, rows, parts, true};
            left_ = (int)th_.size();
            ++gen_;
        }
        cv_.notify_all();
        cslice(dst, dpitch, src, spitch, width, rows, parts - 1, parts);
        std::unique_lock<std::mutex> l(mu_);
        done_.wait(l, [this] { return left_ == 0; });
    }
    // floats -> doubles, contiguous (`count` elements), the same way
    void widen1d(double* dst, const float* src, size_t count) {
        const size_t blk = 16384, n = count / blk;               // 64 KB of floats per share
        if (n >= 16 && !th_.empty()) {
            const int parts = (int)th_.size() + 1;
            {
                std::lock_guard<std::mutex> l(mu_);
                job_ = {reinterpret_cast<char*>(dst), reinterpret_cast<const char*>(src), blk, blk, blk, (int64_t)n, parts, false, true};
                left_ = (int)th_.size();
                ++gen_;
            }
            cv_.notify_all();
            wslice(dst, src, blk, (int64_t)n, parts - 1, parts);
            std::unique_lock<std::mutex> l(mu_);
            done_.wait(l, [this] { return left_ == 0; });
            for (size_t k = n * blk; k < count; ++k) dst[k] = (double)src[k];
        } else {
            for (size_t k = 0; k < count; ++k) dst[k] = (double)src[k];
        }
    }
    void copy1d(void* dst, const void* src, size_t bytes) {      // contiguous: shared out in 64 KB blocks
        const size_t blk = (size_t)64 << 10, n = bytes / blk;
        copy2d(dst, blk, src, blk, blk, (int64_t)n);
        std::memcpy(static_cast<char*>(dst) + n * blk, static_cast<const char*>(src) + n * blk, bytes - n * blk);
    }

private:
    struct Job { char* dst; const char* src; size_t dpitch, spitch, width; int64_t rows; int parts; bool convert; bool widen = false; };
    static void wslice(double* dst, const float* src, size_t blk, int64_t blocks, int part, int parts) {
        const size_t lo = (size_t)(blocks * part / parts) * blk, hi = (size_t)(blocks * (part + 1) / parts) * blk;
        for (size_t k = lo; k < hi; ++k) dst[k] = (double)src[k];
    }
    static void cslice(float* dst, size_t dpitch, const double* src, size_t spitch, size_t width, int64_t rows, int part, int parts) {
        const int64_t lo = rows * part / parts, hi = rows * (part + 1) / parts;
        for (int64_t r = lo; r < hi; ++r) {
            float* d = dst + (size_t)r * dpitch;
            const double* s = src + (size_t)r * spitch;
            for (size_t k = 0; k < width; ++k) d[k] = (float)s[k];      // (values beyond float range become inf: a non-finite row)
        }
    }
    static void slice(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, int64_t rows, int part, int parts) {
        const int64_t lo = rows * part / parts, hi = rows * (part + 1) / parts;
        char* d = static_cast<char*>(dst) + (size_t)lo * dpitch;
        const char* s = static_cast<const char*>(src) + (size_t)lo * spitch;
        if (width == dpitch && width == spitch) { std::memcpy(d, s, (size_t)(hi - lo) * width); return; }
        for (int64_t r = lo; r < hi; ++r, d += dpitch, s += spitch) std::memcpy(d, s, width);
    }
    void loop(int me) {
        unsigned long long seen = 0;
        for (;;) {
            Job j;
            {
                std::unique_lock<std::mutex> l(mu_);
                cv_.wait(l, [&] { return gen_ != seen; });
                seen = gen_;
                if (stop_) return;
                j = job_;
            }
            if (j.widen) wslice(reinterpret_cast<double*>(j.dst), reinterpret_cast<const float*>(j.src), j.width, j.rows, me, j.parts);
            else if (j.convert) cslice(reinterpret_cast<float*>(j.dst), j.dpitch, reinterpret_cast<const double*>(j.src), j.spitch, j.width, j.rows, me, j.parts);
            else slice(j.dst, j.dpitch, j.src, j.spitch, j.width, j.rows, me, j.parts);
            { std::lock_guard<std::mutex> l(mu_); --left_; }
            done_.notify_one();
        }
    }
    std::vector<std::thread> th_;
    std::mutex mu_;
    std::condition_variable cv_, done_;
    Job job_{};
    unsigned long long gen_ = 0;
    int left_ = 0;
    bool stop_ = false;
};

// Device staging for lr_score_host: kSlots chunks in flight, each on its own stream
// (H2D copy -> kernel -> D2H copies), so the two copy engines and the SMs overlap.
struct HostStage {
    static constexpr int kSlots = 3;
    std::mutex mu;
    // pageable caller memory goes through pinned bounce buffers, filled / drained by a few threads (made on first use)
    CopyPool* pool = nullptr;
    float* hx[kSlots] = {};
    float* hproba[kSlots] = {};
    float* hdecision[kSlots] = {};
    int32_t* hlabel[kSlots] = {};
    float* hmaxprob[kSlots] = {};
    int64_t chunk_rows = 0;
    int64_t pitch = 0;        // staged row pitch in elements (n_features rounded up to 4)
    float* x[kSlots] = {};
    float* proba[kSlots] = {};
    float* decision[kSlots] = {};
    int32_t* label[kSlots] = {};
    float* maxprob[kSlots] = {};
    cudaStream_t stream[kSlots] = {};
    unsigned long long* status_host = nullptr;   // pinned landing pad for the status block
    void release() {
        if (status_host) { cudaFreeHost(status_host); status_host = nullptr; }
        delete pool;
        pool = nullptr;
        for (int i = 0; i < kSlots; ++i) {
            cudaFreeHost(hx[i]); cudaFreeHost(hproba[i]); cudaFreeHost(hdecision[i]); cudaFreeHost(hlabel[i]); cudaFreeHost(hmaxprob[i]);
            hx[i] = hproba[i] = hdecision[i] = hmaxprob[i] = nullptr;
            hlabel[i] = nullptr;
            cudaFree(x[i]); cudaFree(proba[i]); cudaFree(decision[i]); cudaFree(label[i]); cudaFree(maxprob[i]);
            if (stream[i]) cudaStreamDestroy(stream[i]);
            x[i] = proba[i] = decision[i] = maxprob[i] = nullptr;
            label[i] = nullptr;
            stream[i] = nullptr;
        }
        chunk_rows = 0;
    }
};

static int select_path(const Model& m, const ScoreArgs& a) {
    if (m.forced_path != LR_PATH_AUTO) return m.forced_path;
    if (m.n_scores == 1) {
        if (m.Wn && m.n_features < 1024) return LR_PATH_NARROW;
        // 1024..2047 features in a large batch: one sweep step per row on the wide path; the narrow path's
        // flat-tile ring does better there (measured 6.5 -> 7.0 TB/s at d = 1024)
        if (m.Wn && m.n_features < 2048 && a.n_rows >= 4096 && a.ldx <= 2048) return LR_PATH_NARROW;
        // unaligned rows up to 2048 floats: the narrow path's flat-tile ring beats 4-byte loads
        const bool aligned = a.ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(a.X) & 15u) == 0;
        if (m.Wn && !aligned && a.ldx <= 2048) return LR_PATH_NARROW;
        if (wide_supports(m, a)) return LR_PATH_WIDE;
        if (m.Wn) return LR_PATH_NARROW;
        return LR_PATH_GENERIC;
    }
    if (m.n_scores > 1 && gemm_short_rows(m, a)) return LR_PATH_GEMM;
    if (m.Wn) {
        // 13..16 classes on rows of 64 floats and more: the register-tile kernel is issue-bound there (~27 fp32
        // TFLOP/s: 3.4 TB/s at 16 classes), the tensor path's small-N form is not (4.4-5.1 TB/s measured at 32).
        // Batches large enough to amortise its extra launches, rows TMA can address.
        const bool aligned = a.ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(a.X) & 15u) == 0;
        if (m.n_scores >= 13 && m.n_features >= 64 && a.n_rows >= 16384 && aligned && gemm_supports(m, a))
            return LR_PATH_GEMM;
        return LR_PATH_NARROW;
    }
    // A handful of rows of a small model: the tensor path's fixed cost (batch statistics, two kernels, the re-scoring
    // pass: ~30 us) is three times what one warp per row needs for rows of <= 4096 multiply-adds and <= 32 classes
    // (d=64, 20 classes: 10-14 us up to 4096 rows, d=128, 30 classes: 13-21; with 100 classes of d=32 the warp's serial
    // walk over the classes takes as long as the tensor path).
    if (m.n_scores <= 32 && (int64_t)m.n_features * m.n_scores <= 4096 && a.n_rows <= 4096) return LR_PATH_GENERIC;
    if (gemm_supports(m, a)) return LR_PATH_GEMM;
    if (!m.groups.empty() && a.n_rows < (int64_t)1 << 31) return LR_PATH_GEMM;     // > 2048 classes: class groups
    return LR_PATH_GENERIC;
}

static int dispatch(const Model& m, const ScoreArgs& a) {
    switch (select_path(m, a)) {
        case LR_PATH_NARROW: return narrow_launch(m, a);
        case LR_PATH_WIDE: return wide_launch(m, a);
        case LR_PATH_GEMM: return m.groups.empty() ? gemm_launch(m, a) : gemm_groups_launch(m, a);
        case LR_PATH_GENERIC: return generic_launch(m, a);
    }
    return fail(LR_ERR_INVALID, "unknown path id");
}

static int check_args(const Model* m, const void* X, int64_t n_rows, int64_t ldx) {
    if (!m) return fail(LR_ERR_INVALID, "model is NULL");
    if (n_rows < 0) return fail(LR_ERR_INVALID, "n_rows is negative");
    if (n_rows > 0 && !X) return fail(LR_ERR_INVALID, "X is NULL");
    if (ldx < m->n_features)
        return fail(LR_ERR_INVALID, "row pitch ldx (" + std::to_string(ldx) + ") is smaller than n_features (" +
                                        std::to_string(m->n_features) + ")");
    return LR_OK;
}

}  // namespace lrs

using namespace lrs;

extern "C" {

const char* lr_last_error(void) { return t_last_error.c_str(); }
int lr_abi_version(void) { return 2; }
uint64_t lr_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int lr_model_create(const double* coef, const double* intercept, int32_t n_scores,
                    int32_t n_features, int32_t device, lr_model** out) {
    if (!out) return fail(LR_ERR_INVALID, "out is NULL");
    *out = nullptr;
    if (!coef || !intercept) return fail(LR_ERR_INVALID, "coef / intercept is NULL");
    if (n_scores < 1 || n_features < 1) return fail(LR_ERR_INVALID, "n_scores and n_features must be >= 1");
    int count = 0;
    LRS_CUDA(cudaGetDeviceCount(&count));
    if (device < 0 || device >= count)
        return fail(LR_ERR_INVALID, "device " + std::to_string(device) + " out of range (" + std::to_string(count) + " visible)");
    DeviceGuard guard(device);
    if (!guard.ok) return fail(LR_ERR_CUDA, "cudaSetDevice failed");
    cudaDeviceProp prop;
    LRS_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(LR_ERR_CUDA, std::string("liblrscore is built for sm_100a only; device is ") + prop.name +
                                     " (sm_" + std::to_string(prop.major) + std::to_string(prop.minor) + ")");

    Model* m = new Model();
    m->device = device;
    m->n_scores = n_scores;
    m->n_features = n_features;
    m->n_out = n_scores < 2 ? 2 : n_scores;
    m->num_sms = prop.multiProcessorCount;

    const size_t nw = (size_t)n_scores * n_features;
    std::vector<float> w32(nw), b32(n_scores);
    for (size_t i = 0; i < nw; ++i) w32[i] = (float)coef[i];
    for (int i = 0; i < n_scores; ++i) b32[i] = (float)intercept[i];

    int rc = LR_OK;
    auto bail = [&](int code) { lr_model_destroy(reinterpret_cast<lr_model*>(m)); return code; };
    if (cudaMalloc(&m->W, nw * sizeof(float)) != cudaSuccess ||
        cudaMalloc(&m->b, n_scores * sizeof(float)) != cudaSuccess ||
        cudaMalloc(&m->status, 8 * sizeof(unsigned long long)) != cudaSuccess)      // [4..7]: class groups' counters
        return bail(fail(LR_ERR_CUDA, std::string("cudaMalloc(model): ") + cudaGetErrorString(cudaGetLastError())));
    if (cudaMemcpy(m->W, w32.data(), nw * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy(m->b, b32.data(), n_scores * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemset(m->status, 0, 8 * sizeof(unsigned long long)) != cudaSuccess)
        return bail(fail(LR_ERR_CUDA, std::string("upload(model): ") + cudaGetErrorString(cudaGetLastError())));
    if ((rc = narrow_prepare(*m, w32.data(), b32.data())) != LR_OK) return bail(rc);
    if ((rc = gemm_prepare(*m, coef, intercept)) != LR_OK) return bail(rc);
    if ((rc = gemm_groups_prepare(*m, coef, intercept)) != LR_OK) return bail(rc);
    m->stage = new HostStage();
    *out = reinterpret_cast<lr_model*>(m);
    return LR_OK;
}

int lr_model_update(lr_model* h, const double* coef, const double* intercept) {
    if (!h) return fail(LR_ERR_INVALID, "model is NULL");
    if (!coef || !intercept) return fail(LR_ERR_INVALID, "coef / intercept is NULL");
    Model* m = reinterpret_cast<Model*>(h);
    DeviceGuard guard(m->device);
    if (!guard.ok) return fail(LR_ERR_CUDA, "cudaSetDevice failed");
    LRS_CUDA(cudaDeviceSynchronize());              // nothing may still be reading the old coefficients
    const size_t nw = (size_t)m->n_scores * m->n_features;
    std::vector<float> w32(nw), b32(m->n_scores);
    for (size_t i = 0; i < nw; ++i) w32[i] = (float)coef[i];
    for (int i = 0; i < m->n_scores; ++i) b32[i] = (float)intercept[i];
    LRS_CUDA(cudaMemcpy(m->W, w32.data(), nw * sizeof(float), cudaMemcpyHostToDevice));
    LRS_CUDA(cudaMemcpy(m->b, b32.data(), m->n_scores * sizeof(float), cudaMemcpyHostToDevice));
    int rc;
    if ((rc = narrow_prepare(*m, w32.data(), b32.data())) != LR_OK) return rc;
    if ((rc = gemm_groups_load(*m, coef, intercept)) != LR_OK) return rc;
    return gemm_load(*m, coef, intercept);
}

void lr_model_destroy(lr_model* h) {
    if (!h) return;
    Model* m = reinterpret_cast<Model*>(h);
    DeviceGuard guard(m->device);
    cudaDeviceSynchronize();
    if (m->stage) { m->stage->release(); delete m->stage; }
    gemm_destroy(*m);
    gemm_groups_destroy(*m);
    cudaFree(m->W); cudaFree(m->b); cudaFree(m->Wn); cudaFree(m->bn); cudaFree(m->status);
    delete m;
}

int lr_model_info(const lr_model* h, int32_t* n_scores, int32_t* n_features, int32_t* device) {
    if (!h) return fail(LR_ERR_INVALID, "model is NULL");
    const Model* m = reinterpret_cast<const Model*>(h);
    if (n_scores) *n_scores = m->n_scores;
    if (n_features) *n_features = m->n_features;
    if (device) *device = m->device;
    return LR_OK;
}

int lr_model_set_path(lr_model* h, int32_t path) {
    if (!h) return fail(LR_ERR_INVALID, "model is NULL");
    if (path < LR_PATH_AUTO || path > LR_PATH_GENERIC) return fail(LR_ERR_INVALID, "unknown path id");
    reinterpret_cast<Model*>(h)->forced_path = path;
    return LR_OK;
}

int lr_model_select_path(const lr_model* h, const float* X_dev, int64_t n_rows, int64_t ldx, int32_t* path) {
    if (!h || !path) return fail(LR_ERR_INVALID, "model / path is NULL");
    const Model* m = reinterpret_cast<const Model*>(h);
    ScoreArgs a{X_dev, n_rows, ldx, nullptr, nullptr, nullptr, nullptr, nullptr};
    *path = select_path(*m, a);
    return LR_OK;
}

int lr_score(lr_model* h, const float* X_dev, int64_t n_rows, int64_t ldx, float* proba_dev,
             float* decision_dev, int32_t* label_dev, float* maxprob_dev, void* stream) {
    Model* m = reinterpret_cast<Model*>(h);
    int rc = check_args(m, X_dev, n_rows, ldx);
    if (rc != LR_OK) return rc;
    if (n_rows == 0) return LR_OK;
    DeviceGuard guard(m->device);
    if (!guard.ok) return fail(LR_ERR_CUDA, "cudaSetDevice failed");
    ScoreArgs a{X_dev, n_rows, ldx, proba_dev, decision_dev, label_dev, maxprob_dev,
                reinterpret_cast<cudaStream_t>(stream)};
    return dispatch(*m, a);
}

int lr_model_take_nonfinite(lr_model* h, void* stream, uint64_t* out) {
    if (!h || !out) return fail(LR_ERR_INVALID, "model / out is NULL");
    Model* m = reinterpret_cast<Model*>(h);
    DeviceGuard guard(m->device);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    unsigned long long v = 0;
    LRS_CUDA(cudaMemcpyAsync(&v, m->status, sizeof(v), cudaMemcpyDeviceToHost, s));
    LRS_CUDA(cudaMemsetAsync(m->status, 0, sizeof(v), s));
    LRS_CUDA(cudaStreamSynchronize(s));
    *out = v;
    return LR_OK;
}

int lr_model_take_rescored(lr_model* h, void* stream, uint64_t* out) {
    if (!h || !out) return fail(LR_ERR_INVALID, "model / out is NULL");
    Model* m = reinterpret_cast<Model*>(h);
    DeviceGuard guard(m->device);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    unsigned long long w[8] = {};                  // the device counters only ever grow; [5] = the class groups' share
    LRS_CUDA(cudaMemcpyAsync(w, m->status, sizeof(w), cudaMemcpyDeviceToHost, s));
    LRS_CUDA(cudaStreamSynchronize(s));
    const unsigned long long v = w[1] + w[5];
    *out = v - m->rescored_taken.exchange(v);
    return LR_OK;
}

int lr_model_take_fallback(lr_model* h, void* stream, uint64_t* out) {
    if (!h || !out) return fail(LR_ERR_INVALID, "model / out is NULL");
    Model* m = reinterpret_cast<Model*>(h);
    DeviceGuard guard(m->device);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    unsigned long long w[8] = {};
    LRS_CUDA(cudaMemcpyAsync(w, m->status, sizeof(w), cudaMemcpyDeviceToHost, s));
    LRS_CUDA(cudaStreamSynchronize(s));
    const unsigned long long v = w[2] + w[6];
    *out = v - m->fallback_taken.exchange(v);
    return LR_OK;
}

int lr_model_set_host_chunk_rows(lr_model* h, int64_t rows) {
    if (!h) return fail(LR_ERR_INVALID, "model is NULL");
    if (rows < 0) return fail(LR_ERR_INVALID, "rows is negative");
    Model* m = reinterpret_cast<Model*>(h);
    std::lock_guard<std::mutex> lock(m->stage->mu);
    DeviceGuard guard(m->device);
    cudaDeviceSynchronize();
    m->stage->release();
    m->host_chunk_rows = rows;
    return LR_OK;
}

// True when `p` is page-locked host memory the device can address directly (cudaHostAlloc /
// cudaHostRegister under unified addressing); `*dev_ptr` receives the device-side alias.
static bool mapped_host(const void* p, size_t bytes, void** dev_ptr) {
    if (!p) { *dev_ptr = nullptr; return true; }
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) { cudaGetLastError(); return false; }
    if (attr.type != cudaMemoryTypeHost || !attr.devicePointer) return false;
    *dev_ptr = attr.devicePointer;
    if (bytes > 1) {                               // the last byte must belong to the same mapped range
        cudaPointerAttributes end;
        const char* last = static_cast<const char*>(p) + bytes - 1;
        if (cudaPointerGetAttributes(&end, last) != cudaSuccess) { cudaGetLastError(); return false; }
        if (end.type != cudaMemoryTypeHost || !end.devicePointer ||
            static_cast<const char*>(end.devicePointer) - static_cast<const char*>(attr.devicePointer) != (ptrdiff_t)(bytes - 1))
            return false;
    }
    return true;
}

static int stage_prepare(Model& m, HostStage& st) {
    if (st.chunk_rows) return LR_OK;
    const int64_t pitch = ((int64_t)m.n_features + 3) / 4 * 4;
    int64_t rows = m.host_chunk_rows;
    if (rows <= 0) {
        // ~32 MiB of X per chunk: large enough to run PCIe at full rate, small enough that the
        // first kernel starts early; at least 8 rows so that wide models still batch.
        rows = (32ll << 20) / (pitch * (int64_t)sizeof(float));
        if (rows < 8) rows = 8;
    }
    for (int i = 0; i < HostStage::kSlots; ++i) {
        LRS_CUDA(cudaMalloc(&st.x[i], rows * pitch * sizeof(float)));
        LRS_CUDA(cudaMalloc(&st.proba[i], rows * m.n_out * sizeof(float)));
        LRS_CUDA(cudaMalloc(&st.decision[i], rows * m.n_scores * sizeof(float)));
        LRS_CUDA(cudaMalloc(&st.label[i], rows * sizeof(int32_t)));
        LRS_CUDA(cudaMalloc(&st.maxprob[i], rows * sizeof(float)));
        LRS_CUDA(cudaStreamCreateWithFlags(&st.stream[i], cudaStreamNonBlocking));
    }
    LRS_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&st.status_host), 4 * sizeof(unsigned long long), cudaHostAllocDefault));
    st.pitch = pitch;
    st.chunk_rows = rows;
    return LR_OK;
}

// X_host (fp32) or X64 (float64, converted on its way into the pinned bounce buffers): exactly one is given
// out64: proba / decision / maxprob are arrays of double (widened from the engine's fp32 on their way out of the pinned
// buffers)
static int score_host_impl(lr_model* h, const float* X_host, const double* X64, int64_t n_rows, int64_t ldx, void* proba_v,
                           void* decision_v, int32_t* label_host, void* maxprob_v, bool out64) {
    float* proba_host = static_cast<float*>(proba_v);           // (as fp32 arrays unless out64)
    float* decision_host = static_cast<float*>(decision_v);
    float* maxprob_host = static_cast<float*>(maxprob_v);
    Model* m = reinterpret_cast<Model*>(h);
    int rc = check_args(m, X64 ? static_cast<const void*>(X64) : static_cast<const void*>(X_host), n_rows, ldx);
    if (rc != LR_OK) return rc;
    if (n_rows == 0) return LR_OK;
    DeviceGuard guard(m->device);
    if (!guard.ok) return fail(LR_ERR_CUDA, "cudaSetDevice failed");
    HostStage& st = *m->stage;
    std::lock_guard<std::mutex> lock(st.mu);
    if ((rc = stage_prepare(*m, st)) != LR_OK) { st.release(); return rc; }

    const int64_t d = m->n_features;
    // A request-sized batch in pinned buffers is scored in place over PCIe (zero-copy): one launch and
    // one synchronise instead of an H2D copy, a launch and up to four D2H copies.  Worth it while the
    // batch is a few KB (a micro-batch of /predict rows); bigger batches want HBM-resident staging.
    if (!X64 && !out64 && n_rows * d * (int64_t)sizeof(float) <= 64 * 1024 && (int64_t)n_rows * m->n_out <= 64 * 1024) {
        void *xd, *pd, *dd, *ld, *md;
        const size_t nr = (size_t)n_rows;
        if (mapped_host(X_host, ((nr - 1) * (size_t)ldx + (size_t)d) * sizeof(float), &xd) &&
            mapped_host(proba_host, nr * m->n_out * sizeof(float), &pd) &&
            mapped_host(decision_host, nr * m->n_scores * sizeof(float), &dd) &&
            mapped_host(label_host, nr * sizeof(int32_t), &ld) && mapped_host(maxprob_host, nr * sizeof(float), &md)) {
            cudaStream_t s = st.stream[0];
            ScoreArgs a{static_cast<const float*>(xd), n_rows, ldx, static_cast<float*>(pd), static_cast<float*>(dd),
                        static_cast<int32_t*>(ld), static_cast<float*>(md), s};
            unsigned long long* flag_host = st.status_host;      // pinned copy of the status block
            if (select_path(*m, a) == LR_PATH_GEMM) goto staged; // TMA descriptors and re-scoring want device memory
            if ((rc = dispatch(*m, a)) != LR_OK) { cudaDeviceSynchronize(); return rc; }
            LRS_CUDA(cudaMemcpyAsync(flag_host, m->status, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
            LRS_CUDA(cudaStreamSynchronize(s));
            gemm_feedback(*m, n_rows, flag_host[1]);
            if (flag_host[0]) {
                const unsigned long long bad = flag_host[0];
                LRS_CUDA(cudaMemset(m->status, 0, sizeof(unsigned long long)));
                return fail(LR_ERR_NONFINITE, std::to_string(bad) + " row(s) scored to NaN or infinity (non-finite input)");
            }
            return LR_OK;
        }
    }
staged:
    // Any failure inside the loop first waits for the copies already in flight on the other slots' streams:
    // they write into the caller's host buffers, which the caller is free to release once this returns.
    auto drain = [&]() { for (int i = 0; i < HostStage::kSlots; ++i) cudaStreamSynchronize(st.stream[i]); };
#define LRS_STAGE(expr)                                                                            \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            drain();                                                                               \
            return fail(LR_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));          \
        }                                                                                          \
    } while (0)
    // Pageable caller memory (an ordinary numpy array): cudaMemcpyAsync from it goes through the driver's own bounce
    // buffer at 8-10 GB/s and does not overlap anything.  Batches of 8 MB and more are instead copied by a few host
    // threads into pinned buffers of the library's, chunk by chunk, while the previous chunks are on the link and
    // on the SMs; results come back the same way.  d=32 binary, 20M rows from a numpy array: 75M -> 219M rows/s (9.6 -> 28
    // GB/s of X; more than four threads add nothing, the host's memory is the limit); pinned X with P into a numpy array:
    // 246M -> 396M rows/s.
    auto pageable = [](const void* p) {
        if (!p) return false;
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return true; }
        return at.type == cudaMemoryTypeUnregistered;
    };
    const bool big = (size_t)n_rows * d * sizeof(float) >= ((size_t)8 << 20);
    const bool bx = X64 != nullptr || (big && pageable(X_host));
    const bool bp = proba_host && (out64 || (big && pageable(proba_host))), bd = decision_host && (out64 || (big && pageable(decision_host))),
               bl = big && pageable(label_host), bm = maxprob_host && (out64 || (big && pageable(maxprob_host)));
    if (bx || bp || bd || bl || bm) {
        if (!st.pool) st.pool = new CopyPool(getenv("LRS_HOST_COPY_THREADS") ? atoi(getenv("LRS_HOST_COPY_THREADS")) : 4);
        for (int i = 0; i < HostStage::kSlots; ++i) {
            if (bx && !st.hx[i]) LRS_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&st.hx[i]), st.chunk_rows * st.pitch * sizeof(float), cudaHostAllocDefault));
            if (bp && !st.hproba[i]) LRS_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&st.hproba[i]), st.chunk_rows * m->n_out * sizeof(float), cudaHostAllocDefault));
            if (bd && !st.hdecision[i]) LRS_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&st.hdecision[i]), st.chunk_rows * m->n_scores * sizeof(float), cudaHostAllocDefault));
            if (bl && !st.hlabel[i]) LRS_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&st.hlabel[i]), st.chunk_rows * sizeof(int32_t), cudaHostAllocDefault));
            if (bm && !st.hmaxprob[i]) LRS_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&st.hmaxprob[i]), st.chunk_rows * sizeof(float), cudaHostAllocDefault));
        }
    }
    struct Pending { int64_t r0 = 0, rows = 0; bool active = false; } pend[HostStage::kSlots];
    // a slot's chunk has landed in the pinned buffers: hand the results to the caller's arrays
    auto collect = [&](int sl) -> cudaError_t {
        if (!pend[sl].active) return cudaSuccess;
        const cudaError_t e = cudaStreamSynchronize(st.stream[sl]);
        if (e != cudaSuccess) return e;
        const int64_t r0 = pend[sl].r0, rows = pend[sl].rows;
        if (out64) {
            if (bp) st.pool->widen1d(static_cast<double*>(proba_v) + r0 * m->n_out, st.hproba[sl], (size_t)rows * m->n_out);
            if (bd) st.pool->widen1d(static_cast<double*>(decision_v) + r0 * m->n_scores, st.hdecision[sl], (size_t)rows * m->n_scores);
            if (bm) st.pool->widen1d(static_cast<double*>(maxprob_v) + r0, st.hmaxprob[sl], (size_t)rows);
        } else {
            if (bp) st.pool->copy1d(proba_host + r0 * m->n_out, st.hproba[sl], (size_t)rows * m->n_out * sizeof(float));
            if (bd) st.pool->copy1d(decision_host + r0 * m->n_scores, st.hdecision[sl], (size_t)rows * m->n_scores * sizeof(float));
            if (bm) std::memcpy(maxprob_host + r0, st.hmaxprob[sl], (size_t)rows * sizeof(float));
        }
        if (bl) std::memcpy(label_host + r0, st.hlabel[sl], (size_t)rows * sizeof(int32_t));
        pend[sl].active = false;
        return cudaSuccess;
    };
    int slot = 0;
    for (int64_t r0 = 0; r0 < n_rows; r0 += st.chunk_rows, slot = (slot + 1) % HostStage::kSlots) {
        const int64_t rows = (n_rows - r0 < st.chunk_rows) ? (n_rows - r0) : st.chunk_rows;
        cudaStream_t s = st.stream[slot];
        const float* src = X_host ? X_host + r0 * ldx : nullptr;
        LRS_STAGE(collect(slot));                 // (also: the slot's pinned X has left for the device)
        if (bx) {
            if (X64) st.pool->convert2d(st.hx[slot], (size_t)st.pitch, X64 + r0 * ldx, (size_t)ldx, (size_t)d, rows);
            else st.pool->copy2d(st.hx[slot], st.pitch * sizeof(float), src, ldx * sizeof(float), d * sizeof(float), rows);
            LRS_STAGE(cudaMemcpyAsync(st.x[slot], st.hx[slot], ((rows - 1) * st.pitch + d) * sizeof(float), cudaMemcpyHostToDevice, s));
            pend[slot].active = true;             // (so that the next use of the slot waits for this copy)
        } else if (ldx == st.pitch) {
            // one contiguous copy; the last row ends after its d features (a pitch padded to 4 floats may run
            // past the caller's buffer there)
            LRS_STAGE(cudaMemcpyAsync(st.x[slot], src, ((rows - 1) * ldx + d) * sizeof(float), cudaMemcpyHostToDevice, s));
        } else {
            LRS_STAGE(cudaMemcpy2DAsync(st.x[slot], st.pitch * sizeof(float), src, ldx * sizeof(float),
                                        d * sizeof(float), rows, cudaMemcpyHostToDevice, s));
        }
        ScoreArgs a{st.x[slot], rows, st.pitch,
                    proba_host ? st.proba[slot] : nullptr, decision_host ? st.decision[slot] : nullptr,
                    label_host ? st.label[slot] : nullptr, maxprob_host ? st.maxprob[slot] : nullptr, s};
        if ((rc = dispatch(*m, a)) != LR_OK) { drain(); return rc; }
        if (proba_host)
            LRS_STAGE(cudaMemcpyAsync(bp ? st.hproba[slot] : proba_host + r0 * m->n_out, st.proba[slot], rows * m->n_out * sizeof(float),
                                      cudaMemcpyDeviceToHost, s));
        if (decision_host)
            LRS_STAGE(cudaMemcpyAsync(bd ? st.hdecision[slot] : decision_host + r0 * m->n_scores, st.decision[slot],
                                      rows * m->n_scores * sizeof(float), cudaMemcpyDeviceToHost, s));
        if (label_host)
            LRS_STAGE(cudaMemcpyAsync(bl ? st.hlabel[slot] : label_host + r0, st.label[slot], rows * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
        if (maxprob_host)
            LRS_STAGE(cudaMemcpyAsync(bm ? st.hmaxprob[slot] : maxprob_host + r0, st.maxprob[slot], rows * sizeof(float), cudaMemcpyDeviceToHost, s));
        pend[slot].r0 = r0;
        pend[slot].rows = rows;
        if (bp || bd || bl || bm) pend[slot].active = true;
    }
    for (int i = 0; i < HostStage::kSlots; ++i) LRS_STAGE(collect(i));
#undef LRS_STAGE
    for (int i = 0; i < HostStage::kSlots; ++i) LRS_CUDA(cudaStreamSynchronize(st.stream[i]));
    unsigned long long sw[2] = {0, 0};            // [0] non-finite rows, [1] rows re-scored in fp32 (running)
    LRS_CUDA(cudaMemcpy(sw, m->status, sizeof(sw), cudaMemcpyDeviceToHost));
    gemm_feedback(*m, n_rows, sw[1]);
    const unsigned long long bad = sw[0];
    if (bad) {
        LRS_CUDA(cudaMemset(m->status, 0, sizeof(bad)));
        return fail(LR_ERR_NONFINITE, std::to_string(bad) + " row(s) scored to NaN or infinity (non-finite input)");
    }
    return LR_OK;
}

int lr_score_host(lr_model* h, const float* X_host, int64_t n_rows, int64_t ldx, float* proba_host,
                  float* decision_host, int32_t* label_host, float* maxprob_host) {
    return score_host_impl(h, X_host, nullptr, n_rows, ldx, proba_host, decision_host, label_host, maxprob_host, false);
}

int lr_score_host_f64(lr_model* h, const double* X_host, int64_t n_rows, int64_t ldx, float* proba_host,
                      float* decision_host, int32_t* label_host, float* maxprob_host) {
    return score_host_impl(h, nullptr, X_host, n_rows, ldx, proba_host, decision_host, label_host, maxprob_host, false);
}

int lr_score_host_ex(lr_model* h, const void* X_host, int x_is_f64, int64_t n_rows, int64_t ldx, void* proba_host,
                     void* decision_host, int32_t* label_host, void* maxprob_host, int out_is_f64) {
    return score_host_impl(h, x_is_f64 ? nullptr : static_cast<const float*>(X_host),
                           x_is_f64 ? static_cast<const double*>(X_host) : nullptr, n_rows, ldx, proba_host, decision_host,
                           label_host, maxprob_host, out_is_f64 != 0);
}

int lr_host_alloc(void** ptr, uint64_t bytes) {
    if (!ptr) return fail(LR_ERR_INVALID, "ptr is NULL");
    *ptr = nullptr;
    if (bytes == 0) return LR_OK;
    LRS_CUDA(cudaHostAlloc(ptr, bytes, cudaHostAllocPortable));
    return LR_OK;
}

int lr_host_free(void* ptr) {
    if (ptr) LRS_CUDA(cudaFreeHost(ptr));
    return LR_OK;
}

int lr_host_register(void* ptr, uint64_t bytes) {
    if (!ptr || bytes == 0) return fail(LR_ERR_INVALID, "empty range");
    LRS_CUDA(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return LR_OK;
}

int lr_host_unregister(void* ptr) {
    if (!ptr) return fail(LR_ERR_INVALID, "ptr is NULL");
    LRS_CUDA(cudaHostUnregister(ptr));
    return LR_OK;
}

}  // extern "C"
