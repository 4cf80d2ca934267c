"""lr_weighted_row_sum (X^T r) bandwidth at a few shapes, 16 GB of X."""
import ctypes as C, os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlapi_b200 import _lib
L = _lib.load()
total = 4 << 30
buf = torch.empty(total, device="cuda")
for lo in range(0, total, 1 << 27):
    buf[lo:lo + (1 << 27)].normal_()
st = torch.cuda.current_stream().cuda_stream
for d in (1024, 4096, 20000, 100000, 100001):
    n = total // d
    X = buf[: n * d].view(n, d)
    r = torch.randn(n, device="cuda")
    out = torch.empty(d, dtype=torch.float64, device="cuda")
    f = lambda: _lib.check(L.lr_weighted_row_sum(C.c_void_p(X.data_ptr()), C.c_void_p(r.data_ptr()), n, d, d, C.c_void_p(out.data_ptr()), C.c_void_p(st)))
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(json.dumps({"d": d, "rows": n, "ms": round(ms, 3), "GBps": round(n * d * 4 / ms / 1e6, 1)}), flush=True)
