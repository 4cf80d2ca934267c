/*
 * liblrscore -- C ABI of the B200 (sm_100a) logistic-regression scoring engine.
 *
 * This is the drop-in boundary for the hot path behind mlAPI's POST /predict.  The reference has
 * no FFI layer of its own: `main.py` calls an unpickled scikit-learn estimator
 * (REF/main.py:19-22).  Each entry point below names the reference call it replaces, so that a
 * maintainer can bind it (ctypes stub in INTEGRATION.md) where `loaded_model.predict(...)` and
 * `loaded_model.predict_proba(...)` are called today.
 *
 * Conventions
 *   - plain C symbols, no exceptions; every call returns an int status (LR_OK == 0) and leaves a
 *     message retrievable with lr_last_error() (thread-local) on failure;
 *   - all matrices are row-major fp32 unless stated; "n_scores" is coef_.shape[0] (1 for a binary
 *     model, C for a C-class multinomial one); a probability row has max(n_scores, 2) columns;
 *   - `*_dev` pointers are device pointers on the model's device (a torch tensor's data_ptr()
 *     works); `*_host` pointers are host pointers (pinned memory makes the copies asynchronous);
 *   - a model is immutable between lr_model_create and lr_model_update: concurrent lr_score calls on distinct
 *     streams are safe; lr_score_host serialises on the model's own staging buffers;
 *   - there is no CPU fallback: without a CUDA device every call fails with LR_ERR_CUDA.
 */
#ifndef LRSCORE_H_
#define LRSCORE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LR_OK             0
#define LR_ERR_INVALID    1   /* bad argument (shape, null pointer, alignment that no path accepts) */
#define LR_ERR_CUDA       2   /* a CUDA runtime/driver call failed; lr_last_error() has the text   */
#define LR_ERR_NONFINITE  3   /* at least one row scored to NaN/inf (non-finite X), see below      */
#define LR_ERR_LIMIT      4   /* request exceeds a documented fixed capacity                         */

/* Which kernel family serves a (model, n_rows) pair; see DESIGN.md "Kernels". */
#define LR_PATH_AUTO     0
#define LR_PATH_NARROW   1    /* <= 16 score columns (<= 24 on short rows): fused row-dot + sigmoid/softmax, few lanes per row;
                                 register-tile form for 5..16 columns, shared-memory ring for unaligned / long rows
                                 (large batches of 13..16 classes on rows of >= 64 floats take LR_PATH_GEMM)        */
#define LR_PATH_WIDE     2    /* binary, rows of 1024 floats and more: row-block streaming, W re-used across 8 rows */
#define LR_PATH_GEMM     3    /* 13..2048 classes: TMA-fed tcgen05 GEMM (3xFP16 split, 3xTF32 when the batch's dynamic
                                 range demands it) + fused softmax; rows TMA cannot address arrive by cp.async; dense rows of
                                 <= 64 floats with <= 32 classes take its short-row form; 2049..65536 classes run as
                                 class groups of <= 2048, combined per row                                        */
#define LR_PATH_GENERIC  4    /* any shape: SIMT row-per-warp fallback (what is left: > 65536 classes)           */

typedef struct lr_model lr_model;

/* Thread-local text of the last failure on this thread ("" if none). */
const char* lr_last_error(void);

/* Library ABI version (bumped on any signature change). */
int lr_abi_version(void);   /* 2 (round 2: additive over 1 -- lr_model_update, lr_grad_*, lr_score_host_f64 / _ex, ...) */

/* Number of kernels this library has launched in this process (for bench.py's gpu_launches). */
uint64_t lr_launch_count(void);

/*
 * Model loader.  Replaces `pickle.load(open('LRClassifier.pkl','rb'))` at REF/main.py:19 as far as
 * the device is concerned: takes the estimator's float64 `coef_` (n_scores x n_features, row-major,
 * contiguous -- note sklearn's lbfgs leaves it F-ordered, copy first) and `intercept_` (n_scores)
 * and makes device-resident fp32 copies (plus the padded / TF32-split layouts the kernels want).
 */
int lr_model_create(const double* coef, const double* intercept, int32_t n_scores,
                    int32_t n_features, int32_t device, lr_model** out);
void lr_model_destroy(lr_model* m);

/* Replace the coefficients of an existing model (same shape) in place: what a refit does once per lbfgs
 * evaluation (REF/Logistic Regression.ipynb:33-34 -> SK/linear_model/_logistic.py:584-598), without
 * re-allocating the device copies or re-encoding the TMA descriptors.  Synchronises the device first;
 * not concurrent with lr_score calls on the same model. */
int lr_model_update(lr_model* m, const double* coef, const double* intercept);

int lr_model_info(const lr_model* m, int32_t* n_scores, int32_t* n_features, int32_t* device);

/* Force a kernel family (tests / profiling); LR_PATH_AUTO restores the dispatcher.  A path that
 * cannot serve the model's shape makes the next lr_score fail with LR_ERR_INVALID. */
int lr_model_set_path(lr_model* m, int32_t path);
/* The path lr_score would take for n_rows rows laid out with row pitch ldx at address X_dev. */
int lr_model_select_path(const lr_model* m, const float* X_dev, int64_t n_rows, int64_t ldx,
                         int32_t* path);

/*
 * The hot path.  One call replaces, for n_rows rows at once, the pair
 *     loaded_model.predict(data_in)                 REF/main.py:21
 *     loaded_model.predict_proba(data_in).max()     REF/main.py:22
 * i.e. sklearn's decision_function (SK/linear_model/_base.py:366-396), predict (:398-427) and
 * predict_proba (SK/linear_model/_logistic.py:1595-1625; sigmoid arm _base.py:429-440, softmax arm
 * SK/utils/extmath.py:985-1024), fused so that X is read once.
 *
 *   X_dev        n_rows x n_features fp32, row pitch ldx elements (ldx >= n_features)
 *   proba_dev    n_rows x max(n_scores,2) fp32, dense          or NULL
 *   decision_dev n_rows (binary) / n_rows x n_scores fp32       or NULL
 *   label_dev    n_rows int32 index into classes_ (binary: score > 0; else first argmax)  or NULL
 *   maxprob_dev  n_rows fp32, the row's largest class probability                          or NULL
 *   stream       cudaStream_t (NULL = legacy default stream)
 *
 * Asynchronous: returns once the work is enqueued.  n_rows == 0 is a no-op.  Rows whose score is
 * not finite (NaN/inf in X -- sklearn raises ValueError for those, _base.py:385-389) are counted in
 * a device-side status word, read with lr_model_take_nonfinite().
 * Device memory: the tensor-core path (LR_PATH_GEMM) takes per-call scratch from the device's default
 * stream-ordered pool (cudaMallocAsync on `stream`): ~16 bytes per row, plus -- when max probabilities are wanted
 * without a P matrix on a model of more than 32 classes -- a P block of at most 512 MB.  The other paths take none.
 */
int lr_score(lr_model* m, const float* X_dev, int64_t n_rows, int64_t ldx,
             float* proba_dev, float* decision_dev, int32_t* label_dev, float* maxprob_dev,
             void* stream);

/* Synchronise `stream`, return the number of non-finite rows seen since the last call, and reset. */
int lr_model_take_nonfinite(lr_model* m, void* stream, uint64_t* n_rows_nonfinite);

/* Synchronise `stream` and return (then reset) how many rows the tensor-core path handed to its fp32
 * re-scoring pass since the last call: rows whose runner-up class sits inside the accumulation error
 * of the maximum, or whose probabilities that error could move by more than the parity bar.
 * Diagnostic; 0 on the other paths. */
int lr_model_take_rescored(lr_model* m, void* stream, uint64_t* n_rows_rescored);

/* Synchronise `stream` and return (then reset) how many rows the tensor-core path scored in 3xTF32
 * instead of its default 3xFP16 since the last call.  The arithmetic is chosen per call on the device:
 * W is balanced per feature at load time so that unstandardised features do not matter, and each batch
 * is sampled; when what is left of its dynamic range would push fp16 `lo` halves into subnormals far
 * enough to move a logit by 2e-6, the whole call runs in 3xTF32 (fp32 exponent range, half the tensor
 * rate).  Same results contract either way (decision_function / predict_proba as lr_score);
 * diagnostic: a model that reports most of its rows here is paying 2x on that path. */
int lr_model_take_fallback(lr_model* m, void* stream, uint64_t* n_rows_tf32);

/*
 * Same contract as lr_score with HOST buffers: rows are streamed through device staging buffers in
 * chunks (H2D copy, kernel, D2H copy overlapped on three streams) and the call returns when every
 * output is in host memory.  This is the call the estimator wrapper and the /predict micro-batcher
 * make, and the one bench.py's `e2e` figure times.  Returns LR_ERR_NONFINITE (outputs still
 * written) when a row scored to NaN/inf.  A batch of a few KB whose X and outputs all live in
 * page-locked memory (lr_host_alloc / lr_host_register) is scored in place over PCIe instead: one
 * launch, no copies (not on the tensor-core path).  Pageable buffers (an ordinary numpy array) of 8 MB and more
 * are copied to and from pinned buffers of the library's by four host threads of its own, chunk by chunk, in step
 * with the transfers (page-locked buffers go straight to the link and are ~2x faster still).
 */
int lr_score_host(lr_model* m, const float* X_host, int64_t n_rows, int64_t ldx,
                  float* proba_host, float* decision_host, int32_t* label_host,
                  float* maxprob_host);

/*
 * lr_score_host for float64 rows (what scikit-learn users hold: numpy's default dtype; the reference itself hands
 * Python floats to predict, REF/main.py:20-22).  The rows are rounded to fp32 -- the engine's input type -- by the
 * library's host threads on their way into its pinned buffers, so no fp32 copy of X is ever made in host memory;
 * values beyond fp32's range become infinities, i.e. non-finite rows.  ldx in float64 elements.
 */
int lr_score_host_f64(lr_model* m, const double* X_host, int64_t n_rows, int64_t ldx,
                      float* proba_host, float* decision_host, int32_t* label_host,
                      float* maxprob_host);

/*
 * The general form: X_host fp32 (x_is_f64 == 0) or float64; proba / decision / maxprob arrays of fp32
 * (out_is_f64 == 0) or of float64 -- scikit-learn's output type -- widened from the engine's fp32 by the same host
 * threads on their way out of the pinned buffers (numpy's astype of a 4 GB P matrix takes seconds on its one thread).
 */
int lr_score_host_ex(lr_model* m, const void* X_host, int x_is_f64, int64_t n_rows, int64_t ldx,
                     void* proba_host, void* decision_host, int32_t* label_host, void* maxprob_host,
                     int out_is_f64);

/* Rows per staging chunk used by lr_score_host (0 = default: about 32 MiB of X per chunk). */
int lr_model_set_host_chunk_rows(lr_model* m, int64_t rows);

/*
 * Refit row: loss and gradient of the log-loss over n_rows rows in ONE fused pass over X -- what
 * `LogisticRegression().fit` (REF/Logistic Regression.ipynb:33-34) evaluates per lbfgs iteration
 * through LinearModelLoss.loss_gradient (SK/linear_model/_linear_loss.py:291-378; pointwise terms
 * SK/_loss/_loss.pyx.tp:728-751 binary, :269/:1202 multinomial).
 *
 *   y_dev    n_rows int32 class indices (binary: 0 / 1)
 *   W_dev    n_scores x n_features fp32 (current iterate), b_dev n_scores fp32
 *   out_dev  float64 [1 + n_scores*n_features + n_scores] = [ sum_i loss_i | sum_i r_i x_i^T | sum_i r_i ]
 *
 * The sums are UNNORMALISED and cover only the rows given: ranks all-reduce `out_dev` (the single
 * collective of the design), then divide by n and add the L2 term on the host (intercept
 * unpenalised).  Serves n_scores <= 8 and n_features <= 128 (16-byte aligned rows) or <= 32
 * (any rows); larger shapes return LR_ERR_LIMIT.  Per-CTA partial vectors go through device scratch
 * owned by the library (allocated on first use per device); calls on one device must be
 * stream-ordered with respect to each other.  Uses the current device unless X_dev says otherwise.
 */
int lr_loss_grad(const float* X_dev, const int32_t* y_dev, int64_t n_rows, int64_t ldx,
                 const float* W_dev, const float* b_dev, int32_t n_scores, int32_t n_features,
                 double* out_dev, void* stream);

/*
 * lr_loss_grad for callers that must not allocate.  (No counterpart in the reference: this variant
 * exists for the embedding side, not for parity.)  Same arguments and results as lr_loss_grad, but
 * the per-CTA partial sums go to a fixed array built into the library (about 0.6 MB per GPU, held
 * until the process exits) instead of a cudaMalloc'ed buffer, and the call makes no cudaMalloc /
 * cudaFree or any other allocating call.  Meant for hosts that give all device memory to their own
 * pool at start-up, that cannot let a refit step stall other streams inside the allocator, or whose
 * rules keep third-party code off the device heap.
 *   - limit: 1 + n_scores*(n_features+1) <= 512 floats per partial vector, else LR_ERR_LIMIT
 *     (the array has room for 296 vectors of that size);
 *   - the first call on each device looks the array up (one-time set-up): make that call before
 *     starting any stream capture;
 *   - X_dev must be on the current device;
 *   - all calls on a device share the one array, which still holds the previous call's data: the
 *     arrival counter is cleared on `stream` before every launch, and calls whose GPU work overlaps
 *     race with each other -- keep them ordered.
 */
int lr_loss_grad_static(const float* X_dev, const int32_t* y_dev, int64_t n_rows, int64_t ldx,
                        const float* W_dev, const float* b_dev, int32_t n_scores, int32_t n_features,
                        double* out_dev, void* stream);

/*
 * out_dev[k] = sum_i r_dev[i] * X_dev[i][k]  (float64 [n_features]): the transposed product of a binary
 * model's gradient, `X.T @ grad_pointwise` in LinearModelLoss.loss_gradient
 * (SK/linear_model/_linear_loss.py:353-357), for feature counts beyond lr_loss_grad (the d = 100 000
 * config).  The forward pass of such a model is lr_score (the residual r = sigmoid(z) - y is known only
 * after a whole row has been read), so X is streamed a second time here; HBM-bound.  Any n_features, any
 * pitch (128-bit loads when rows are 16-byte aligned).  Sums over the rows given only: ranks all-reduce.
 * Per-call scratch is stream-ordered (cudaMallocAsync).  Deterministic.
 */
int lr_weighted_row_sum(const float* X_dev, const float* r_dev, int64_t n_rows, int64_t ldx,
                        int32_t n_features, double* out_dev, void* stream);

/*
 * Gradient of a multinomial model, the transposed product: out_dev[c][k] = sum_i R[i][c] * X[i][k] and
 * out_dev[c][n_features] = sum_i R[i][c] (float64 [n_scores][n_features + 1], the layout of [coef_ | intercept_]),
 * i.e. `grad_pointwise.T @ X` and `grad_pointwise.sum(axis=0)` of LinearModelLoss.loss_gradient
 * (SK/linear_model/_linear_loss.py:353-357, :375) with R = softmax(z) - onehot(y), which lr_score's probability
 * matrix becomes by subtracting 1 at each row's label.  tcgen05 GEMM over MN-major operands (the contraction
 * runs over the rows), 3xFP16 split with fp32-grade products, float64 combine in a fixed order: deterministic.
 *
 * X is constant over a fit: lr_grad_plan_create scales (per feature) and splits it into fp16 halves once and
 * owns those copies (as many bytes as X itself) plus the scratch every call needs; lr_grad_gemm then takes a
 * block of rows [row0, row0 + n_rows) of that X with the matching rows of R (n_rows <= max_block_rows; R_dev
 * n_rows x n_scores fp32, row pitch ldr).  Sums cover the rows given only: blocks are added by the caller and
 * ranks all-reduce.  1..2048 score rows.  Calls on one plan must be stream-ordered.
 */
typedef struct lr_grad_plan lr_grad_plan;
int lr_grad_plan_create(const float* X_dev, int64_t n_rows, int64_t ldx, int32_t n_features, int32_t max_scores,
                        int64_t max_block_rows, void* stream, lr_grad_plan** out);
void lr_grad_plan_destroy(lr_grad_plan* plan);
int lr_grad_gemm(lr_grad_plan* plan, int64_t row0, int64_t n_rows, const float* R_dev, int64_t ldr,
                 int32_t n_scores, double* out_dev, void* stream);

/*
 * Page-locked host memory for the buffers handed to lr_score_host (the /predict micro-batcher's
 * staging rows, bench.py's e2e inputs).  With pageable memory the copies still work but the driver
 * stages them and the three-stream overlap is lost.  No reference counterpart: REF/main.py:20
 * builds a Python list per request.
 */
int lr_host_alloc(void** ptr, uint64_t bytes);
int lr_host_free(void* ptr);
/* Pin / unpin an existing host range in place (e.g. a numpy array's buffer). */
int lr_host_register(void* ptr, uint64_t bytes);
int lr_host_unregister(void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* LRSCORE_H_ */
