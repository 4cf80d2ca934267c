#!/usr/bin/env python
"""BASELINE cfg 0, first half: one /predict-sized call (1 row, d=4, C=3).  Times what
REF/main.py:19-22 does per request (unpickle + predict + predict_proba().max(), scikit-learn on the
host) next to the engine's equivalents: the estimator wrapper and the bare C-ABI call."""
import ctypes as C
import io
import json
import os
import pickle
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import synth  # noqa: E402


def timeit(fn, n=2000, warm=200):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    ts = np.array(ts) * 1e6
    return {"p50_us": float(np.percentile(ts, 50)), "p99_us": float(np.percentile(ts, 99)), "mean_us": float(ts.mean())}


def main():
    import warnings
    warnings.simplefilter("ignore")
    import __graft_entry__
    __graft_entry__.build()
    from mlapi_b200 import B200LogisticRegression, _lib
    from mlapi_b200.estimator import PinnedArray

    sk, _ = synth.fit_notebook_model()
    blob = pickle.dumps(sk)
    row = [[5.1, 3.5, 1.4, 0.2]]

    def reference_request():
        m = pickle.load(io.BytesIO(blob))                 # REF/main.py:19 (from memory, kinder than disk)
        return m.predict(row)[0], m.predict_proba(row).max()

    clf = B200LogisticRegression.from_sklearn(sk)

    def wrapper_request():
        lab, p = clf.predict_with_max_proba(row)
        return lab[0], p[0]

    L = _lib.load()
    rep = clf._replicas[0]
    x = PinnedArray((1, 4)); x.array[...] = row
    lab = PinnedArray((1,), np.int32); mp = PinnedArray((1,), np.float32)

    def abi_request():
        L.lr_score_host(rep.handle, x.array.ctypes.data_as(C.c_void_p), 1, 4, None, None,
                        lab.array.ctypes.data_as(C.c_void_p), mp.array.ctypes.data_as(C.c_void_p))
        return lab.array[0], mp.array[0]

    out = {"reference_sklearn_per_request": timeit(reference_request),
           "b200_estimator_wrapper": timeit(wrapper_request),
           "b200_c_abi_lr_score_host": timeit(abi_request),
           "answers": [str(reference_request()), str(wrapper_request()), str(abi_request())],
           "host_cores": len(os.sched_getaffinity(0))}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
