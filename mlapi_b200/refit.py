"""Batch-gradient refit: `LogisticRegression().fit` (REF/Logistic Regression.ipynb:33-34) with the
loss/gradient pass on the GPU and, across ranks, ONE NCCL all-reduce per iteration.

scikit-learn minimises  (1/n) sum_i loss_i + 1/(2 C n) ||W||^2  (intercept unpenalised) with
scipy's L-BFGS-B (SK/linear_model/_logistic.py:580-598: maxiter=max_iter, maxls=50, gtol=tol,
ftol=64*eps, zero start).  The same optimiser runs here on every rank in float64; per iteration the
current (W, b) goes to the device, this rank's row shard is reduced to one float64 vector
[loss | grad_W | grad_b], the ranks sum that vector (`sharding.allreduce_sum`: 132 B at d=32 binary,
400 KB at d=100k, 4.1 MB at 1000 x 1024), and every rank takes the identical step.

Two ways to that vector:
  * small models (<= 8 score rows, d <= 128): `lr_loss_grad`, one fused pass over X (refit.cu);
  * every other shape (the d=100k binary and the 1000-class configs among them): the forward pass --
    scores, softmax, labels -- is the scoring engine's own kernels (`lr_score` on the current iterate,
    any path), the residual `softmax - onehot` is formed in place in the probability block, and the
    transposed product `residual^T X` is `lr_weighted_row_sum` (a second streaming pass over X) for a
    binary model and `lr_grad_gemm` (the library's own tcgen05 kernel over MN-major operands, 3xFP16 split,
    fp32-grade) for a multinomial one, taken in row blocks and accumulated in float64.
Torch is the allocator and the NCCL binding; the elementwise steps between the two library calls of the
second way (loss terms, the -1 at each row's label) are torch tensor expressions.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _lib
from .sharding import allreduce_sum


class FitResult:
    def __init__(self, coef, intercept, classes, n_iter, loss):
        self.coef_ = coef
        self.intercept_ = intercept
        self.classes_ = classes
        self.n_iter_ = np.array([n_iter], dtype=np.int32)
        self.n_features_in_ = coef.shape[1]
        self.final_loss_ = loss

    def to_engine(self, **kw):
        from .estimator import B200LogisticRegression
        return B200LogisticRegression(self.coef_, self.intercept_, self.classes_, **kw)


def loss_grad_sums(X, y_idx, W32, b32, out):
    """One `lr_loss_grad` call on torch CUDA tensors; `out` (float64) receives the local sums."""
    import torch
    lib = _lib.load()
    n, d = X.shape
    stream = torch.cuda.current_stream(X.device).cuda_stream
    with torch.cuda.device(X.device):
        _lib.check(lib.lr_loss_grad(C.c_void_p(X.data_ptr()), C.c_void_p(y_idx.data_ptr()), n, X.stride(0),
                                    C.c_void_p(W32.data_ptr()), C.c_void_p(b32.data_ptr()), W32.shape[0], d,
                                    C.c_void_p(out.data_ptr()), C.c_void_p(stream)))
    return out


def fused_kernel_serves(rows: int, d: int, X) -> bool:
    """Shapes `lr_loss_grad` takes (include/lrscore.h): <= 8 score rows and d <= 128 with 16-byte aligned
    rows, or d <= 32 with any rows."""
    aligned = d % 4 == 0 and X.stride(0) % 4 == 0 and X.data_ptr() % 16 == 0
    return rows <= 8 and (d <= 32 or (d <= 128 and aligned))


class Evaluator:
    """[sum loss | sum r x^T | sum r] over this rank's rows at a point (W, b), for any model shape.

    Holds what does not change between the evaluations of a fit: the device model (`lr_model_update` swaps the
    coefficients in place), and for a multinomial model the gradient plan (`lr_grad_plan_create`: X scaled and
    split into fp16 halves once).  Follows SK/linear_model/_linear_loss.py:291-378: binary loss
    log(1+e^z) - y z with residual sigmoid(z) - y, multinomial logsumexp(z) - z_y with softmax(z) - onehot(y).

    Small models (`fused_kernel_serves`) take `lr_loss_grad`, one fused pass over X.  Every other shape: the
    forward pass -- scores, softmax, labels -- is the scoring engine on the current iterate (`lr_score`, whichever
    path the shape selects); the residual is formed in place in the probability block; the transposed product
    `residual^T X` is `lr_weighted_row_sum` for a binary model (a second streaming pass over X) and
    `lr_grad_gemm` for a multinomial one (tcgen05, MN-major operands, fp32-grade 3xFP16 split), per row block,
    blocks summed in float64.
    """

    def __init__(self, X, y_idx, rows: int, block_rows: int = 0, force_general: bool = False):
        import torch
        self.X, self.y_idx, self.rows = X, y_idx, int(rows)
        self.n, self.d = X.shape
        self.dev = X.device
        self.lib = _lib.load()
        self.fused = fused_kernel_serves(self.rows, self.d, X) and not force_general
        self.out = torch.empty(1 + self.rows * (self.d + 1), dtype=torch.float64, device=self.dev)
        self.est = None
        self.plan = C.c_void_p()
        if self.fused:
            self.W32 = torch.empty((self.rows, self.d), dtype=torch.float32, device=self.dev)
            self.b32 = torch.empty((self.rows,), dtype=torch.float32, device=self.dev)
            return
        if block_rows <= 0:                                       # ~1 GB of probabilities (or of X) per block
            block_rows = max(1024, min(self.n, (1 << 28) // max(self.rows, self.d // 8, 2)))
        self.block_rows = min(int(block_rows), self.n)
        if self.rows > 1:
            stream = torch.cuda.current_stream(self.dev).cuda_stream
            with torch.cuda.device(self.dev):
                _lib.check(self.lib.lr_grad_plan_create(C.c_void_p(X.data_ptr()), self.n, X.stride(0), self.d, self.rows,
                                                        self.block_rows, C.c_void_p(stream), C.byref(self.plan)))
            self.gblk = torch.empty((self.rows, self.d + 1), dtype=torch.float64, device=self.dev)
        else:
            self.gblk = torch.empty((self.d,), dtype=torch.float64, device=self.dev)

    def close(self):
        if self.est is not None:
            self.est.close()
            self.est = None
        if self.plan:
            self.lib.lr_grad_plan_destroy(self.plan)
            self.plan = C.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def sums(self, Wb64):
        """`Wb64`: rows x (d+1) float64, last column the intercept.  Returns the float64 device vector of local sums."""
        import torch
        rows, d, dev = self.rows, self.d, self.dev
        Wb64 = np.asarray(Wb64, dtype=np.float64).reshape(rows, d + 1)
        if self.fused:
            self.W32.copy_(torch.from_numpy(np.ascontiguousarray(Wb64[:, :d], dtype=np.float32)))
            self.b32.copy_(torch.from_numpy(np.ascontiguousarray(Wb64[:, d], dtype=np.float32)))
            return loss_grad_sums(self.X, self.y_idx, self.W32, self.b32, self.out)
        from .estimator import B200LogisticRegression
        if self.est is None:
            self.est = B200LogisticRegression(Wb64[:, :d], Wb64[:, d], np.arange(max(rows, 2)), devices=[dev.index])
        else:
            self.est.update(Wb64[:, :d], Wb64[:, d])
        est, lib = self.est, self.lib
        stream = torch.cuda.current_stream(dev).cuda_stream
        loss = torch.zeros((), dtype=torch.float64, device=dev)
        G = torch.zeros((rows, d + 1), dtype=torch.float64, device=dev)
        for lo in range(0, self.n, self.block_rows):
            Xc = self.X[lo:lo + self.block_rows]
            yc = self.y_idx[lo:lo + self.block_rows].long()
            if rows == 1:
                res = est._score_device(Xc, True, True, False, False)
                z = res["decision"].double()
                yv = yc.double()
                # log(1 + e^z) - y z, evaluated as max(z, 0) + log1p(e^-|z|) (SK/_loss/_loss.pyx.tp:728-751)
                loss += (torch.clamp(z, min=0) + torch.log1p(torch.exp(-z.abs())) - yv * z).sum()
                r = (res["proba"][:, 1] - yc.float()).contiguous()
                with torch.cuda.device(dev):                  # X^T r: the library's own streaming kernel
                    _lib.check(lib.lr_weighted_row_sum(C.c_void_p(Xc.data_ptr()), C.c_void_p(r.data_ptr()), Xc.shape[0],
                                                       Xc.stride(0), d, C.c_void_p(self.gblk.data_ptr()), C.c_void_p(stream)))
                G[0, :d] += self.gblk
                G[0, d] += r.double().sum()
            else:
                res = est._score_device(Xc, True, True, True, True)
                P, Z = res["proba"], res["decision"]
                lab = res["label"].long()
                # logsumexp from the winning class: lse = z_max - log p_max (p_max >= 1/C never underflows)
                lse = Z.gather(1, lab[:, None]).squeeze(1).double() - torch.log(res["maxprob"].double())
                loss += (lse - Z.gather(1, yc[:, None]).squeeze(1).double()).sum()
                P[torch.arange(P.shape[0], device=dev), yc] -= 1.0          # residual, in place
                with torch.cuda.device(dev):                  # R^T X and the column sums of R: the library's tcgen05 kernel
                    _lib.check(lib.lr_grad_gemm(self.plan, lo, P.shape[0], C.c_void_p(P.data_ptr()), P.stride(0), rows,
                                                C.c_void_p(self.gblk.data_ptr()), C.c_void_p(stream)))
                G += self.gblk
        if est.take_nonfinite():
            raise ValueError("Input X contains NaN or infinity.")
        out = self.out
        out[0] = loss
        out[1:1 + rows * d] = G[:, :d].reshape(-1)
        out[1 + rows * d:] = G[:, d]
        return out


def loss_grad_sums_general(X, y_idx, Wb64, out, block_rows: int = 0):
    """One evaluation through the general path (see `Evaluator`) even where `lr_loss_grad` would serve -- tests
    compare the two -- into `out` (float64)."""
    ev = Evaluator(X, y_idx, np.asarray(Wb64).shape[0], block_rows, force_general=True)
    try:
        out.copy_(ev.sums(Wb64))
    finally:
        ev.close()
    return out


def fit(X, y, classes=None, C_reg: float = 1.0, tol: float = 1e-4, max_iter: int = 100,
        device: Optional[int] = None) -> FitResult:
    """Fit on this rank's rows `X` (n_local, d) / labels `y`.  With torch.distributed initialised
    every rank passes its own shard (and the same `classes`); all ranks return the same model."""
    import torch
    import torch.distributed as dist
    from scipy.optimize import minimize

    if not torch.is_tensor(X):
        X = torch.as_tensor(np.ascontiguousarray(X, dtype=np.float32))
    if not X.is_cuda:
        X = X.to(torch.device("cuda", torch.cuda.current_device() if device is None else device))
    X = X.float()
    if X.stride(1) != 1:
        X = X.contiguous()
    dev = X.device
    y = np.asarray(y.cpu() if torch.is_tensor(y) else y)
    classes = np.unique(y) if classes is None else np.asarray(classes)
    n_classes = len(classes)
    if n_classes < 2:
        raise ValueError("This solver needs samples of at least 2 classes in the data, but the data "
                         f"contains only one class: {classes[0]!r}")
    y_idx = torch.as_tensor(np.searchsorted(classes, y).astype(np.int32), device=dev)
    n_local, d = X.shape
    rows = 1 if n_classes == 2 else n_classes
    n_total = torch.tensor([float(n_local)], dtype=torch.float64, device=dev)
    allreduce_sum(n_total)
    n = float(n_total.item())
    lam = 1.0 / (C_reg * n)                                   # SK/linear_model/_logistic.py:580
    evals = {"n": 0}
    ev = Evaluator(X, y_idx, rows)

    def fun(w):
        Wb = w.reshape(rows, d + 1)
        out = ev.sums(Wb)
        allreduce_sum(out)                                    # the design's only collective
        h = out.cpu().numpy()
        evals["n"] += 1
        loss = h[0] / n + 0.5 * lam * float(np.sum(Wb[:, :d] ** 2))
        gW = h[1:1 + rows * d].reshape(rows, d) / n + lam * Wb[:, :d]
        gb = h[1 + rows * d:] / n
        return loss, np.concatenate([gW, gb[:, None]], axis=1).ravel()

    res = minimize(fun, np.zeros(rows * (d + 1)), method="L-BFGS-B", jac=True,
                   options={"maxiter": max_iter, "maxls": 50, "gtol": tol,
                            "ftol": 64 * np.finfo(float).eps})
    ev.close()
    Wb = res.x.reshape(rows, d + 1)
    r = FitResult(Wb[:, :d].copy(), Wb[:, d].copy(), classes, int(res.nit), float(res.fun))
    r.n_evals_ = evals["n"]
    r.world_ = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    return r
