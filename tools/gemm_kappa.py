"""Relative logit error of the tensor-core path (re-scoring off) per row length, layout and arithmetic:
the numbers behind GemmParams::kappa.  Prints max and 99.9th percentile of |dz| / max(1, |z|).
Usage: python tools/gemm_kappa.py [n_rows]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import synth  # noqa: E402
from oracle import lr_oracle as orc  # noqa: E402
from mlapi_b200 import B200LogisticRegression, _lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 2048
os.environ["LRS_GEMM_FIXUP"] = "0"
if "--short" in sys.argv:
    # the short-row kernel (dense device rows, >= 4096 of them)
    import torch
    for d, c in ((16, 26), (32, 20), (48, 16), (54, 20), (64, 32), (64, 17)):
        m = 1 << 17
        X = synth.make_X(m, d, 5)
        W, b = synth.make_model(d, c, 6)
        z = orc.decision_function(X, W, b)
        clf = B200LogisticRegression(W, b, np.arange(c), out_dtype=np.float32)
        out = clf.score_rows(torch.from_numpy(X).cuda(), decision=True)
        torch.cuda.synchronize()
        rel = np.abs(out["decision"].cpu().numpy() - z) / np.maximum(1.0, np.abs(z))
        print(f"short rows d={d} C={c}: max {rel.max():.2e}  p99.9 {np.quantile(rel, 0.999):.2e}  mean {rel.mean():.2e}", flush=True)
        clf.close()
    sys.exit(0)
for d, c in ((256, 1000), (1024, 1000), (4096, 1000), (16384, 1000), (4096, 64), (16384, 64)):
    X = synth.make_X(n, d, 5)
    W, b = synth.make_model(d, c, 6)
    z = orc.decision_function(X, W, b)
    for layout in (0, 1, 2):
        for arith, f16 in (("fp16", "2"), ("tf32", "0")):
            os.environ["LRS_GEMM_LAYOUT"] = str(layout)
            os.environ["LRS_GEMM_F16"] = f16
            clf = B200LogisticRegression(W, b, np.arange(c), out_dtype=np.float32)
            clf.set_path(_lib.PATH_GEMM)
            out = clf.score_rows(X, decision=True)
            rel = np.abs(out["decision"] - z) / np.maximum(1.0, np.abs(z))
            print(f"d={d} C={c} layout={layout} {arith}: max {rel.max():.2e}  p99.9 {np.quantile(rel, 0.999):.2e}  "
                  f"mean {rel.mean():.2e}  (x d/1024: max {rel.max() * 1024 / d:.2e})", flush=True)
            clf.close()
