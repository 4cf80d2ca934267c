"""sklearn-compatible wrapper over liblrscore: the Python operator boundary of the hot path.

`REF/main.py:19-22` unpickles a `sklearn.linear_model.LogisticRegression` and calls `predict` and
`predict_proba` on it.  `B200LogisticRegression` is the drop-in for that object: same method names,
argument meaning, return shapes and error behaviour (SK/linear_model/_base.py:366-440,
SK/linear_model/_logistic.py:1595-1650), built from the same pickle, with the arithmetic done by
the sm_100a kernels behind include/lrscore.h.  There is no CPU fallback.

Host arrays (lists, numpy) go through `lr_score_host`; CUDA torch tensors go through `lr_score`
without leaving the device and come back as torch tensors.  With several devices the rows are split
into contiguous blocks, one model replica and one host thread per GPU, no collective
(SURVEY.md section 8e).
"""
from __future__ import annotations

import ctypes as C
import pickle
import threading
from concurrent.futures import ThreadPoolExecutor
from typing import Optional, Sequence

import numpy as np

from . import _lib
from .sharding import shard_bounds


def _is_torch_tensor(x) -> bool:
    return type(x).__module__.split(".")[0] == "torch" and hasattr(x, "data_ptr")


class PinnedArray:
    """A float32/int32 numpy array living in page-locked memory from lr_host_alloc."""

    def __init__(self, shape, dtype=np.float32):
        self.shape = tuple(int(s) for s in np.atleast_1d(shape))
        self.dtype = np.dtype(dtype)
        nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        self._ptr = C.c_void_p()
        _lib.check(_lib.load().lr_host_alloc(C.byref(self._ptr), max(nbytes, 1)))
        buf = (C.c_char * max(nbytes, 1)).from_address(self._ptr.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def free(self):
        if self._ptr is not None and self._ptr.value:
            self.array = None
            _lib.load().lr_host_free(self._ptr)
            self._ptr = None

    def __del__(self):  # pragma: no cover - best effort
        try:
            self.free()
        except Exception:
            pass


class _Replica:
    """One lr_model on one device."""

    def __init__(self, coef64: np.ndarray, intercept64: np.ndarray, device: int):
        self.lib = _lib.load()
        self.device = int(device)
        self.handle = C.c_void_p()
        _lib.check(self.lib.lr_model_create(
            coef64.ctypes.data_as(C.c_void_p), intercept64.ctypes.data_as(C.c_void_p),
            coef64.shape[0], coef64.shape[1], self.device, C.byref(self.handle)))

    def close(self):
        if self.handle:
            self.lib.lr_model_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class B200LogisticRegression:
    """Duck-type of a fitted `sklearn.linear_model.LogisticRegression` for scoring.

    Parameters
    ----------
    coef, intercept, classes : the estimator's `coef_` (n_scores, n_features), `intercept_`
        (n_scores,), `classes_`.
    devices : CUDA device indices to replicate the model on (default: device 0).
    out_dtype : dtype of returned host arrays.  sklearn returns float64 for a float64 `coef_`
        (SURVEY.md section 3.2); the engine computes in fp32, so float64 here is a widening copy
        kept for drop-in fidelity; pass np.float32 to skip it.
    """

    def __init__(self, coef, intercept, classes, devices: Optional[Sequence[int]] = None,
                 out_dtype=np.float64):
        coef64 = np.ascontiguousarray(np.atleast_2d(np.asarray(coef, dtype=np.float64)))
        intercept64 = np.ascontiguousarray(np.asarray(intercept, dtype=np.float64).reshape(-1))
        if coef64.ndim != 2 or intercept64.shape[0] != coef64.shape[0]:
            raise ValueError(f"coef_ {coef64.shape} and intercept_ {intercept64.shape} do not match")
        classes = np.asarray(classes)
        n_classes = max(coef64.shape[0], 2)
        if classes.shape[0] != n_classes:
            raise ValueError(f"classes_ has {classes.shape[0]} entries, coef_ implies {n_classes}")
        self.coef_ = coef64
        self.intercept_ = intercept64
        self.classes_ = classes
        self.n_features_in_ = coef64.shape[1]
        self.out_dtype = np.dtype(out_dtype)
        self.devices = [0] if devices is None else [int(d) for d in devices]
        self._replicas = [_Replica(coef64, intercept64, d) for d in self.devices]
        self._pool = ThreadPoolExecutor(max_workers=len(self._replicas)) if len(self._replicas) > 1 else None
        self._lock = threading.Lock()

    # ---- construction ---------------------------------------------------------------------------
    @classmethod
    def from_sklearn(cls, est, **kw) -> "B200LogisticRegression":
        """From a fitted sklearn estimator.  `coef_` out of lbfgs is F-ordered:
        it is copied to C order here."""
        for attr in ("coef_", "intercept_", "classes_"):
            if not hasattr(est, attr):
                from sklearn.exceptions import NotFittedError
                raise NotFittedError(
                    f"This {type(est).__name__} instance is not fitted yet. Call 'fit' with "
                    "appropriate arguments before using this estimator.")
        # The engine computes what a default LogisticRegression() does: sigmoid for a binary model, softmax
        # otherwise.  Pickles from the scikit-learn the reference pins (0.24.1, REF/requirements.txt:11) can
        # ask for something else -- one-vs-rest normalised sigmoids (multi_class='ovr', or 'auto' with
        # solver='liblinear') for > 2 classes, softmax([-z, z]) for a binary multi_class='multinomial' model
        # (SK 0.24 _logistic.py predict_proba) -- and would load and answer with different probabilities.
        multi_class = getattr(est, "multi_class", "auto")
        solver = getattr(est, "solver", "lbfgs")
        n_classes = len(est.classes_)
        if multi_class == "deprecated":
            multi_class = "auto"
        ovr = multi_class == "ovr" or (multi_class == "auto" and solver == "liblinear")
        if n_classes > 2 and ovr:
            raise NotImplementedError(
                "one-vs-rest LogisticRegression (multi_class='ovr' or solver='liblinear') with more than two "
                "classes normalises per-class sigmoids; this engine implements the softmax (multinomial) arm only")
        if n_classes <= 2 and multi_class == "multinomial":
            raise NotImplementedError(
                "a binary multi_class='multinomial' model scores with softmax([-z, z]); this engine implements "
                "the sigmoid arm for binary models")
        return cls(est.coef_, est.intercept_, est.classes_, **kw)

    def update(self, coef, intercept):
        """New coefficients of the same shape, in place on every replica (`lr_model_update`): what the refit
        driver does once per lbfgs evaluation instead of building a new model."""
        coef64 = np.ascontiguousarray(np.atleast_2d(np.asarray(coef, dtype=np.float64)))
        intercept64 = np.ascontiguousarray(np.asarray(intercept, dtype=np.float64).reshape(-1))
        if coef64.shape != self.coef_.shape or intercept64.shape != self.intercept_.shape:
            raise ValueError(f"update keeps the shape: coef_ {self.coef_.shape}, got {coef64.shape}")
        with self._lock:
            for r in self._replicas:
                _lib.check(r.lib.lr_model_update(r.handle, coef64.ctypes.data_as(C.c_void_p),
                                                 intercept64.ctypes.data_as(C.c_void_p)))
        self.coef_ = coef64
        self.intercept_ = intercept64

    def close(self):
        for r in self._replicas:
            r.close()
        if self._pool:
            self._pool.shutdown(wait=False)

    # ---- input validation (SK/linear_model/_base.py:385-389 semantics) ---------------------------
    def _validate_host(self, X) -> np.ndarray:
        Xa = np.asarray(X)
        if Xa.dtype == object or Xa.dtype.kind not in "fiub":
            Xa = np.asarray(X, dtype=np.float64)     # raises ValueError on non-numeric content
        if Xa.ndim == 1 or Xa.ndim == 0:
            raise ValueError(
                f"Expected 2D array, got {'scalar' if Xa.ndim == 0 else '1D'} array instead:\narray={Xa}.\n"
                "Reshape your data either using array.reshape(-1, 1) if your data has a single feature "
                "or array.reshape(1, -1) if it contains a single sample.")
        if Xa.ndim != 2:
            raise ValueError(f"Found array with dim {Xa.ndim}. LogisticRegression expected <= 2.")
        if Xa.shape[0] == 0:
            raise ValueError(
                f"Found array with 0 sample(s) (shape={Xa.shape}) while a minimum of 1 is required "
                "by LogisticRegression.")
        if Xa.shape[1] != self.n_features_in_:
            raise ValueError(
                f"X has {Xa.shape[1]} features, but LogisticRegression is expecting "
                f"{self.n_features_in_} features as input.")
        # float64 rows (numpy's default, what sklearn users hold) stay as they are: the library rounds them to fp32 on
        # their way into its pinned staging buffers (lr_score_host_f64) -- an fp32 copy made here by numpy, on one
        # thread, took ten times as long as scoring it
        if Xa.dtype == np.float64:
            return np.ascontiguousarray(Xa)
        return np.ascontiguousarray(Xa, dtype=np.float32)

    # ---- the one scoring call ----------------------------------------------------------------
    def score_rows(self, X, proba=False, decision=False, label=False, maxprob=False, out_f64=False) -> dict:
        """Single pass over X producing any subset of {proba, decision, label, maxprob}.  Host arrays: float32
        results, or float64 ones (`out_f64`, scikit-learn's output type) widened by the library's own host threads."""
        if _is_torch_tensor(X) and X.is_cuda:
            return self._score_device(X, proba, decision, label, maxprob)
        Xf = self._validate_host(X)
        n = Xf.shape[0]
        ns = self.coef_.shape[0]
        ft = np.float64 if out_f64 else np.float32
        out = {
            "proba": np.empty((n, max(ns, 2)), ft) if proba else None,
            "decision": np.empty((n,) if ns == 1 else (n, ns), ft) if decision else None,
            "label": np.empty((n,), np.int32) if label else None,
            "maxprob": np.empty((n,), ft) if maxprob else None,
        }
        d = self.n_features_in_

        def run(rep: _Replica, lo: int, hi: int):
            sl = lambda a: None if a is None else a[lo:hi]  # noqa: E731
            return rep.lib.lr_score_host_ex(rep.handle, _ptr(Xf[lo:hi]), int(Xf.dtype == np.float64), hi - lo, d,
                                            _ptr(sl(out["proba"])), _ptr(sl(out["decision"])), _ptr(sl(out["label"])),
                                            _ptr(sl(out["maxprob"])), int(out_f64)), _lib.last_error()

        with self._lock:
            G = len(self._replicas)
            if G == 1 or n < 2 * G:
                results = [run(self._replicas[0], 0, n)]
            else:
                bounds = shard_bounds(n, G)                     # contiguous row blocks, no collective
                results = list(self._pool.map(lambda g: run(self._replicas[g], *bounds[g]), range(G)))
        for code, msg in results:
            if code == _lib.LR_ERR_NONFINITE:
                raise ValueError("Input X contains NaN or infinity (or a value too large for float32).")
            if code != _lib.LR_OK:
                raise _lib.EngineError(code, msg)
        return out

    def _score_device(self, X, proba, decision, label, maxprob) -> dict:
        import torch

        if X.dim() != 2:
            raise ValueError(f"Expected 2D tensor, got {X.dim()}D")
        if X.shape[0] == 0:
            raise ValueError("Found tensor with 0 sample(s) while a minimum of 1 is required by LogisticRegression.")
        if X.shape[1] != self.n_features_in_:
            raise ValueError(f"X has {X.shape[1]} features, but LogisticRegression is expecting "
                             f"{self.n_features_in_} features as input.")
        if X.dtype != torch.float32:
            X = X.float()
        if X.stride(1) != 1 and X.shape[1] > 1:
            X = X.contiguous()
        dev = X.device.index if X.device.index is not None else torch.cuda.current_device()
        rep = next((r for r in self._replicas if r.device == dev), None)
        if rep is None:
            raise ValueError(f"X is on cuda:{dev} but the model is replicated on {self.devices}")
        n, ns = X.shape[0], self.coef_.shape[0]
        ldx = X.stride(0) if n > 1 else max(X.stride(0), X.shape[1])
        mk = lambda shape, dt: torch.empty(shape, dtype=dt, device=X.device)  # noqa: E731
        out = {
            "proba": mk((n, max(ns, 2)), torch.float32) if proba else None,
            "decision": mk((n,) if ns == 1 else (n, ns), torch.float32) if decision else None,
            "label": mk((n,), torch.int32) if label else None,
            "maxprob": mk((n,), torch.float32) if maxprob else None,
        }
        p = lambda t: None if t is None else C.c_void_p(t.data_ptr())  # noqa: E731
        stream = torch.cuda.current_stream(X.device).cuda_stream
        _lib.check(rep.lib.lr_score(rep.handle, C.c_void_p(X.data_ptr()), n, ldx, p(out["proba"]),
                                    p(out["decision"]), p(out["label"]), p(out["maxprob"]),
                                    C.c_void_p(stream)))
        return out

    def take_nonfinite(self, device_index: int = 0) -> int:
        """Number of rows that scored to NaN/inf on the device path since the last call (syncs)."""
        rep = self._replicas[device_index]
        v = C.c_uint64()
        _lib.check(rep.lib.lr_model_take_nonfinite(rep.handle, None, C.byref(v)))
        return int(v.value)

    def take_rescored(self, device_index: int = 0) -> int:
        """Rows the tensor-core path re-scored in fp32 since the last call (diagnostic; syncs)."""
        rep = self._replicas[device_index]
        v = C.c_uint64()
        _lib.check(rep.lib.lr_model_take_rescored(rep.handle, None, C.byref(v)))
        return int(v.value)

    def take_fallback(self, device_index: int = 0) -> int:
        """Rows the tensor-core path scored in 3xTF32 because the batch's dynamic range ruled out 3xFP16,
        since the last call (diagnostic; syncs)."""
        rep = self._replicas[device_index]
        v = C.c_uint64()
        _lib.check(rep.lib.lr_model_take_fallback(rep.handle, None, C.byref(v)))
        return int(v.value)

    # ---- sklearn surface -----------------------------------------------------------------------
    def _cast(self, a):
        return a if _is_torch_tensor(a) or a.dtype == self.out_dtype else a.astype(self.out_dtype)

    def decision_function(self, X):
        """SK/linear_model/_base.py:366-396: (n,) for a binary model, (n, C) otherwise."""
        return self._cast(self.score_rows(X, decision=True, out_f64=self.out_dtype == np.float64)["decision"])

    def predict_proba(self, X):
        """SK/linear_model/_logistic.py:1595-1625."""
        return self._cast(self.score_rows(X, proba=True, out_f64=self.out_dtype == np.float64)["proba"])

    def predict_log_proba(self, X):
        """SK/linear_model/_logistic.py:1627-1650: log of predict_proba."""
        P = self.predict_proba(X)
        return P.log() if _is_torch_tensor(P) else np.log(P)

    def predict(self, X):
        """SK/linear_model/_base.py:398-427: classes_[index]."""
        idx = self.score_rows(X, label=True)["label"]
        if _is_torch_tensor(idx):
            return idx
        return self.classes_.take(idx, axis=0)

    def predict_with_max_proba(self, X):
        """`predict(X)` and `predict_proba(X).max(axis=1)` -- the pair REF/main.py:21-22 computes
        with two passes -- from one pass over X."""
        out = self.score_rows(X, label=True, maxprob=True, out_f64=self.out_dtype == np.float64)
        if _is_torch_tensor(out["label"]):
            return out["label"], out["maxprob"]
        return self.classes_.take(out["label"], axis=0), self._cast(out["maxprob"])

    def score(self, X, y):
        """Mean accuracy (REF/Logistic Regression.ipynb:42)."""
        pred = self.predict(X)
        return float(np.mean(np.asarray(pred) == np.asarray(y)))

    def set_path(self, path: int):
        for r in self._replicas:
            _lib.check(r.lib.lr_model_set_path(r.handle, path))

    def __repr__(self):
        return (f"B200LogisticRegression(n_features={self.n_features_in_}, "
                f"n_classes={len(self.classes_)}, devices={self.devices})")


def load_model(path: str = "LRClassifier.pkl", **kw) -> B200LogisticRegression:
    """Model loader: the same pickle `REF/main.py:19` opens (needs scikit-learn importable to
    unpickle), loaded ONCE and moved to device-resident fp32 buffers."""
    with open(path, "rb") as fh:
        est = pickle.load(fh)
    return B200LogisticRegression.from_sklearn(est, **kw)
