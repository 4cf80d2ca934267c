"""Host-call latency of the tensor-core path for request-sized batches (lr_score_host, 1000 classes x d=1024)."""
import os, sys, time, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlapi_b200 import B200LogisticRegression
rng = np.random.default_rng(0)
d, c = 1024, 1000
clf = B200LogisticRegression(rng.standard_normal((c, d)) / 32, rng.standard_normal(c), np.arange(c), out_dtype=np.float32)
out = []
for n in (1, 64, 1024):
    X = rng.standard_normal((n, d), dtype=np.float32)
    for _ in range(20):
        clf.predict_with_max_proba(X)
    ts = []
    for _ in range(200):
        t0 = time.perf_counter(); clf.predict_with_max_proba(X); ts.append((time.perf_counter() - t0) * 1e6)
    ts.sort()
    rec = {"rows": n, "p50_us": round(ts[100], 1), "p99_us": round(ts[197], 1)}
    print(json.dumps(rec)); out.append(rec)
clf.close()
