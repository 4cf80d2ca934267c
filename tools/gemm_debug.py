"""First-contact check of the tcgen05 GEMM path on a GPU: small cases against the oracle with the
error printed, instead of a bare pass/fail.  Usage: python tools/gemm_debug.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import synth  # noqa: E402
from oracle import lr_oracle as orc  # noqa: E402
from mlapi_b200 import B200LogisticRegression, _lib  # noqa: E402

cases = [(32, 100, 64), (32, 256, 128), (64, 512, 300), (1024, 1000, 48), (1024, 1000, 1000),
         (36, 17, 5), (100, 2048, 257), (1024, 1000, 40_000)]
if len(sys.argv) > 1:
    cases = cases[: int(sys.argv[1])]
for d, c, n in cases:
    W, b = synth.make_model(d, c, 6)
    X = synth.make_X(n, d, 5)
    clf = B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(_lib.PATH_GEMM)
    out = clf.score_rows(X, proba=True, decision=True, label=True, maxprob=True)
    z = orc.decision_function(X, W, b)
    P = orc.predict_proba(X, W, b)
    idx = orc.predict_index(X, W, b)
    ez = np.abs(out["decision"] - z).max()
    ep = np.abs(out["proba"] - P).max()
    lab = (out["label"] != idx).sum()
    emp = np.abs(out["maxprob"] - P.max(1)).max()
    print(f"d={d} C={c} n={n}: max|dz|={ez:.3e} max|dP|={ep:.3e} label mismatches={lab} max|dmaxprob|={emp:.3e}",
          flush=True)
    clf.close()
