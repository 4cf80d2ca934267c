"""Re-scoring rate of the short-row kernel per shape (debug)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlapi_b200 import B200LogisticRegression
for d, c in ((64, 32), (64, 17), (16, 26), (32, 20), (64, 24), (64, 25), (64, 28)):
    rng = np.random.default_rng(d * 7919 + c)
    W = rng.standard_normal((c, d)) / np.sqrt(d)
    b = rng.standard_normal(c)
    clf = B200LogisticRegression(W, b, list(range(c)))
    n = 1 << 20
    X = torch.randn((n, d), device="cuda")
    out = clf.score_rows(X, proba=True, label=True, maxprob=True)
    torch.cuda.synchronize()
    r = clf.take_rescored()
    mp = out["maxprob"]
    print(d, c, "rescored", r, f"{r / n:.4f}", "fallback", clf.take_fallback(), "maxprob quantiles",
          [round(float(v), 3) for v in torch.quantile(mp[:100000], torch.tensor([0.1, 0.5, 0.9], device="cuda"))], flush=True)
    clf.close()
