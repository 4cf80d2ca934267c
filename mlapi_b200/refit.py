"""Batch-gradient refit: `LogisticRegression().fit` (REF/Logistic Regression.ipynb:33-34) with the
loss/gradient pass on the GPU and, across ranks, ONE NCCL all-reduce per iteration.

scikit-learn minimises  (1/n) sum_i loss_i + 1/(2 C n) ||W||^2  (intercept unpenalised) with
scipy's L-BFGS-B (SK/linear_model/_logistic.py:580-598: maxiter=max_iter, maxls=50, gtol=tol,
ftol=64*eps, zero start).  The same optimiser runs here on every rank in float64; per iteration the
current (W, b) goes to the device (a few hundred bytes), `lr_loss_grad` reduces this rank's row
shard to one float64 vector [loss | grad_W | grad_b], the ranks sum that vector
(`sharding.allreduce_sum`: 132 B at d=32 binary), and every rank takes the identical step.
Torch is the allocator and the NCCL binding here; the arithmetic is liblrscore's.
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
    out = torch.empty(1 + rows * (d + 1), dtype=torch.float64, device=dev)
    W32 = torch.empty((rows, d), dtype=torch.float32, device=dev)
    b32 = torch.empty((rows,), dtype=torch.float32, device=dev)
    evals = {"n": 0}

    def fun(w):
        Wb = w.reshape(rows, d + 1)
        W32.copy_(torch.from_numpy(np.ascontiguousarray(Wb[:, :d], dtype=np.float32)))
        b32.copy_(torch.from_numpy(np.ascontiguousarray(Wb[:, d], dtype=np.float32)))
        loss_grad_sums(X, y_idx, W32, b32, out)
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
    Wb = res.x.reshape(rows, d + 1)
    r = FitResult(Wb[:, :d].copy(), Wb[:, d].copy(), classes, int(res.nit), float(res.fun))
    r.n_evals_ = evals["n"]
    r.world_ = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    return r
