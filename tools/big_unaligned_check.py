"""Full-size runs of the unaligned-row kernels (flat-tile ring at d=30, 100M rows; 4-byte-load wide kernel at
d=100001, 60k rows): rows sampled across the whole batch against a float64 dot product."""
import ctypes as C, os, sys, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlapi_b200 import B200LogisticRegression, _lib
L = _lib.load()
st = torch.cuda.current_stream().cuda_stream
for d, n in ((30, 100_000_000), (100_001, 60_000), (54, 40_000_000)):
    c = 7 if d == 54 else 1
    X = torch.empty((n, d), device="cuda")
    step = max(1, (1 << 27) // d)
    for lo in range(0, n, step):
        X[lo:lo + step].normal_()
    rng = np.random.default_rng(d)
    W = rng.standard_normal((c, d)) / np.sqrt(d); b = rng.standard_normal(c)
    clf = B200LogisticRegression(W, b, list(range(max(c, 2))))
    P = torch.empty((n, max(c, 2)), device="cuda")
    _lib.check(L.lr_score(clf._replicas[0].handle, C.c_void_p(X.data_ptr()), n, d, C.c_void_p(P.data_ptr()), None, None, None, C.c_void_p(st)))
    torch.cuda.synchronize()
    idx = torch.tensor(sorted(set([0, 1, n - 1, n - 2] + list(rng.integers(0, n, 2000)))), device="cuda")
    Xs = X[idx].double().cpu().numpy(); Ps = P[idx].cpu().numpy()
    z = Xs @ W.T + b
    if c == 1:
        p1 = 1 / (1 + np.exp(-z[:, 0])); ref = np.stack([1 - p1, p1], 1)
    else:
        e = np.exp(z - z.max(1, keepdims=True)); ref = e / e.sum(1, keepdims=True)
    err = np.abs(Ps - ref).max()
    s = float(P.sum(dtype=torch.float64).item())
    print(json.dumps({"d": d, "C": c, "rows": n, "max_abs_dP_sampled": float(err), "sum_of_P_minus_rows": s - n, "nonfinite": clf.take_nonfinite()}), flush=True)
    assert err < 1e-5 and abs(s - n) < 1e-3 * n
    clf.close(); del X, P
print("OK")
