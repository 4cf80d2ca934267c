"""CPU oracle for the /predict scoring path -- TEST INFRASTRUCTURE ONLY.

This file is a plain-numpy float64 restatement of the arithmetic that the
reference runs when `main.py` calls `loaded_model.predict(...)` and
`loaded_model.predict_proba(...)` (REF/main.py:21-22).  The reference itself
holds no arithmetic: it lives in the third-party dependency scikit-learn
(pinned `scikit-learn==0.24.1`, REF/requirements.txt:11; 1.9.0 is what this
image has, same function for a default `LogisticRegression()`).  Every function
below cites the scikit-learn 1.9.0 file:line it restates
(SK = site-packages/sklearn).

Who may import this: `tests/`, `__graft_entry__.smoke()` and `bench.py`'s
`cpu_baseline` / `--impl reference` legs, and only as the checker or the timed
CPU baseline.  Nothing under `mlapi_b200/` imports it; the product path has no
CPU fallback.

Parity status: the reference ships no golden vectors or tests for this path
(SURVEY.md section 8c), so parity is pinned against OUTPUTS OF THE REFERENCE'S
OWN IMPLEMENTATION RUN HERE: `tests/golden/make_golden.py` imports scikit-learn
in this container, runs `decision_function` / `predict_proba` / `predict`
(and `LinearModelLoss.loss_gradient` for the refit row) and commits the
float64 results under `tests/golden/`; `tests/test_oracle.py` checks every
function below against those files.
"""
from __future__ import annotations

import numpy as np

__all__ = [
    "decision_function",
    "expit",
    "softmax",
    "predict_proba",
    "predict_index",
    "predict",
    "max_proba",
    "log1pexp",
    "loss_gradient",
    "fit_lbfgs",
]


def _as_2d_coef(coef: np.ndarray, intercept: np.ndarray):
    coef = np.asarray(coef, dtype=np.float64)
    intercept = np.asarray(intercept, dtype=np.float64).reshape(-1)
    if coef.ndim == 1:
        coef = coef[None, :]
    return coef, intercept


def decision_function(X, coef, intercept) -> np.ndarray:
    """scores = X @ coef_.T + intercept_; (n,) if one score column else (n, C).

    SK/linear_model/_base.py:390-396 (`safe_sparse_dot(X, self.coef_.T) +
    self.intercept_`, then `reshape(scores, (-1,))` when `scores.shape[1] == 1`).
    float32 X is promoted against the float64 coef_.
    """
    coef, intercept = _as_2d_coef(coef, intercept)
    X = np.asarray(X, dtype=np.float64)
    scores = X @ coef.T + intercept
    if scores.shape[1] == 1:
        return scores.reshape(-1)
    return scores


def expit(z: np.ndarray) -> np.ndarray:
    """Logistic sigmoid 1/(1+exp(-z)), overflow-free.

    SK/linear_model/_base.py:437 calls `scipy.special.expit` (via
    SK/utils/_array_api.py:631-636).  scipy evaluates it branch-wise so that
    exp() never overflows; the same two branches are written out here.
    """
    z = np.asarray(z, dtype=np.float64)
    out = np.empty_like(z)
    pos = z >= 0
    out[pos] = 1.0 / (1.0 + np.exp(-z[pos]))
    ez = np.exp(z[~pos])
    out[~pos] = ez / (1.0 + ez)
    return out


def softmax(scores: np.ndarray) -> np.ndarray:
    """Row softmax: subtract the row max, exp, divide by the row sum.

    SK/utils/extmath.py:1012-1024.
    """
    s = np.array(scores, dtype=np.float64, copy=True)
    s -= s.max(axis=1, keepdims=True)
    np.exp(s, out=s)
    s /= s.sum(axis=1, keepdims=True)
    return s


def predict_proba(X, coef, intercept) -> np.ndarray:
    """(n, 2) = stack([1-p, p]) for <=2 classes, else (n, C) softmax.

    Branch: SK/linear_model/_logistic.py:1620-1625; binary arm
    SK/linear_model/_base.py:429-440 (`_predict_proba_lr`: expit in place, then
    `np.vstack([1 - prob, prob]).T`).
    """
    scores = decision_function(X, coef, intercept)
    if scores.ndim == 1:
        p = expit(scores)
        return np.stack([1.0 - p, p], axis=1)
    return softmax(scores)


def predict_index(X, coef, intercept) -> np.ndarray:
    """Index into classes_: `scores > 0` (binary, tie -> 0) or first argmax.

    SK/linear_model/_base.py:415-418.
    """
    scores = decision_function(X, coef, intercept)
    if scores.ndim == 1:
        return (scores > 0).astype(np.int64)
    return np.argmax(scores, axis=1).astype(np.int64)


def predict(X, coef, intercept, classes) -> np.ndarray:
    """classes_.take(indices): SK/linear_model/_base.py:423."""
    return np.asarray(classes).take(predict_index(X, coef, intercept), axis=0)


def max_proba(X, coef, intercept) -> np.ndarray:
    """Per-row max class probability: the `.max()` of REF/main.py:22 applied row-wise."""
    return predict_proba(X, coef, intercept).max(axis=1)


# --------------------------------------------------------------------------------------
# Refit row: the objective scikit-learn's lbfgs solver minimises.
# --------------------------------------------------------------------------------------

def log1pexp(z: np.ndarray) -> np.ndarray:
    """log(1+exp(z)) with the 4-branch stable evaluation of SK/_loss/_loss.pyx.tp:728-751
    (thresholds -37, -2, 18, 33.3)."""
    z = np.asarray(z, dtype=np.float64)
    out = np.empty_like(z)
    a = z <= -37
    b = (z > -37) & (z <= -2)
    c = (z > -2) & (z <= 18)
    d = (z > 18) & (z <= 33.3)
    e = z > 33.3
    out[a] = np.exp(z[a])
    out[b] = np.log1p(np.exp(z[b]))
    out[c] = np.log(1.0 + np.exp(z[c]))
    out[d] = z[d] + np.exp(-z[d])
    out[e] = z[e]
    return out


def loss_gradient(W, b, X, y_index, l2_reg_strength):
    """Loss and gradient of the averaged log-loss + L2 penalty (intercept unpenalised).

    SK/linear_model/_linear_loss.py:291-378 with `l2_reg_strength = 1/(C*n)`
    (SK/linear_model/_logistic.py:580) and sample_weight = 1/n.
    Binary: W (1,d), pointwise loss log1pexp(z) - y*z, gradient sigmoid(z) - y
    (SK/_loss/_loss.pyx.tp:728-751).  Multinomial: W (C,d), loss logsumexp(z) - z_y,
    gradient softmax(z) - onehot(y) (SK/_loss/_loss.pyx.tp:269,1202).
    Returns (loss, grad_W, grad_b) in float64.
    """
    W, b = _as_2d_coef(W, b)
    X = np.asarray(X, dtype=np.float64)
    y_index = np.asarray(y_index).astype(np.int64)
    n = X.shape[0]
    z = X @ W.T + b
    if W.shape[0] == 1:
        z = z[:, 0]
        y = y_index.astype(np.float64)
        loss = np.sum(log1pexp(z) - y * z) / n
        r = (expit(z) - y)[:, None] / n
    else:
        zmax = z.max(axis=1, keepdims=True)
        lse = np.log(np.exp(z - zmax).sum(axis=1)) + zmax[:, 0]
        loss = np.sum(lse - z[np.arange(n), y_index]) / n
        r = softmax(z)
        r[np.arange(n), y_index] -= 1.0
        r /= n
    loss += 0.5 * l2_reg_strength * np.sum(W * W)
    grad_W = r.T @ X + l2_reg_strength * W
    grad_b = r.sum(axis=0)
    return float(loss), grad_W, grad_b


def fit_lbfgs(X, y_index, n_classes, C=1.0, tol=1e-4, max_iter=100):
    """Default `LogisticRegression().fit` (REF/Logistic Regression.ipynb:33-34):
    L-BFGS-B on `loss_gradient`, settings from SK/linear_model/_logistic.py:584-598
    (maxiter=max_iter, maxls=50, gtol=tol, ftol=64*eps), zero init.
    Returns (coef, intercept, n_iter)."""
    from scipy.optimize import minimize

    X = np.asarray(X, dtype=np.float64)
    n, d = X.shape
    rows = 1 if n_classes <= 2 else n_classes
    lam = 1.0 / (C * n)

    def fun(w):
        Wb = w.reshape(rows, d + 1)
        loss, gW, gb = loss_gradient(Wb[:, :d], Wb[:, d], X, y_index, lam)
        return loss, np.concatenate([gW, gb[:, None]], axis=1).ravel()

    res = minimize(
        fun,
        np.zeros(rows * (d + 1)),
        method="L-BFGS-B",
        jac=True,
        options={"maxiter": max_iter, "maxls": 50, "gtol": tol,
                 "ftol": 64 * np.finfo(float).eps},
    )
    Wb = res.x.reshape(rows, d + 1)
    return Wb[:, :d].copy(), Wb[:, d].copy(), int(res.nit)
