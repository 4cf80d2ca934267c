"""How far the tensor-core path is from the float64 oracle, per logit scale.
Usage: python tools/gemm_precision.py [n_rows]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import synth  # noqa: E402
from oracle import lr_oracle as orc  # noqa: E402
from mlapi_b200 import B200LogisticRegression, _lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
d, c = 1024, 1000
X = synth.make_X(n, d, 5)
for scale in (1.0, 3.0, 8.0):
    W, b = synth.make_model(d, c, 6)
    W = W * scale
    z = orc.decision_function(X, W, b)
    P = orc.softmax(z)
    idx = z.argmax(1)
    top2 = np.partition(z, -2, axis=1)[:, -2:]
    margin = top2[:, 1] - top2[:, 0]
    for fix in ("1", "0"):
        os.environ["LRS_GEMM_FIXUP"] = fix
        clf = B200LogisticRegression(W, b, np.arange(c), out_dtype=np.float32)
        clf.set_path(_lib.PATH_GEMM)
        out = clf.score_rows(X, proba=True, decision=True, label=True, maxprob=True)
        resc = clf.take_rescored()
        bad = out["label"] != idx
        tol_viol = np.abs(out["proba"] - P) > (1e-5 + 1e-5 * P)
        print(f"scale={scale} fixup={fix}: |z|max={np.abs(z).max():.1f} pmax={P.max():.3f} "
              f"max|dz|={np.abs(out['decision'] - z).max():.2e} max|dP|={np.abs(out['proba'] - P).max():.2e} "
              f"P-tolerance violations={int(tol_viol.sum())} label mismatches: total={int(bad.sum())} "
              f"outside 1e-5 band={int((bad & (margin >= 1e-5)).sum())} rows with margin<1e-4: {int((margin < 1e-4).sum())} "
              f"rescored rows={resc}",
              flush=True)
        clf.close()
