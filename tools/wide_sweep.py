"""Wide-kernel HBM throughput against row length at a fixed 24 GB footprint (is the d=100k figure
limited by the access pattern -- 8 row streams per CTA, 400 KB apart?)."""
import ctypes as C, os, sys, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlapi_b200 import B200LogisticRegression, _lib
L = _lib.load()
total = 6 * (1 << 30)          # floats
buf = torch.empty(total, device="cuda")
for lo in range(0, total, 1 << 27):
    buf[lo:lo + (1 << 27)].normal_()
for d in (1024, 4096, 8192, 32768, 100_000, 400_000):
    n = total // d
    X = buf[: n * d].view(n, d)
    W = np.random.default_rng(0).standard_normal((1, d)) / np.sqrt(d)
    clf = B200LogisticRegression(W, [0.1], [0, 1])
    clf.set_path(_lib.PATH_WIDE)
    rep = clf._replicas[0]
    P = torch.empty((n, 2), device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    def step():
        _lib.check(L.lr_score(rep.handle, C.c_void_p(X.data_ptr()), n, d, C.c_void_p(P.data_ptr()), None, None, None, C.c_void_p(st)))
    for _ in range(3): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): step()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(json.dumps({"d": d, "rows": n, "ms": ms, "GBps": (n * d * 4 + n * 8) / ms / 1e6}), flush=True)
    clf.close()
