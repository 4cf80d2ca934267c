"""Every kernel family once, on small shapes, for `compute-sanitizer --tool memcheck`.
Usage (GPU box): compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import synth  # noqa: E402
from oracle import lr_oracle as orc  # noqa: E402
from mlapi_b200 import B200LogisticRegression, _lib  # noqa: E402

ALL = dict(proba=True, decision=True, label=True, maxprob=True)


def check(name, clf, X, W, b):
    out = clf.score_rows(X, **ALL)
    err = np.abs(out["proba"] - orc.predict_proba(X, W, b)).max()
    print(f"{name}: max|dP| = {err:.2e}", flush=True)
    assert err < 1e-5
    clf.close()


rng = np.random.default_rng(0)
for d, ns, n, path in ((32, 1, 1000, _lib.PATH_NARROW), (4, 3, 333, _lib.PATH_NARROW), (7, 2, 100, _lib.PATH_NARROW),
                       (200, 5, 77, _lib.PATH_NARROW), (5000, 1, 37, _lib.PATH_WIDE), (33, 20, 65, _lib.PATH_GENERIC),
                       (64, 300, 200, _lib.PATH_GEMM), (1024, 1000, 300, _lib.PATH_GEMM), (40, 1500, 130, _lib.PATH_GEMM)):
    W, b = synth.make_model(d, ns, d + ns)
    X = rng.standard_normal((n, d), dtype=np.float32)
    clf = B200LogisticRegression(W, b, np.arange(max(ns, 2)), out_dtype=np.float32)
    clf.set_path(path)
    check(f"path {_lib.PATH_NAMES[path]} d={d} C={ns} n={n}", clf, X, W, b)

# this round's kernels: 5..16 classes (W in shared memory / through L1), flat tiles (unaligned and long aligned
# rows, >= 4096 rows), small-N tensor form, unaligned rows on the tensor path
for d, ns, n in ((40, 10, 500), (61, 7, 300), (2600, 12, 200), (33, 1, 5000), (5, 1, 4100), (257, 3, 4100), (512, 1, 4100),
                 (1001, 4, 4100), (64, 32, 1000), (200, 100, 700), (131, 100, 500), (1022, 300, 260)):
    W, b = synth.make_model(d, ns, d + ns)
    X = rng.standard_normal((n, d), dtype=np.float32)
    clf = B200LogisticRegression(W, b, np.arange(max(ns, 2)), out_dtype=np.float32)
    check(f"auto d={d} C={ns} n={n}", clf, X, W, b)

# late round 2: 33..64 classes in registers, class counts that leave through the store tiles, and device rows TMA
# cannot address (the kernel's cp.async loader: 4- and 8-byte copies, both arithmetics)
for d, ns, n in ((96, 50, 700), (300, 64, 300), (96, 101, 400), (64, 255, 300), (128, 1001, 200), (64, 150, 129)):
    W, b = synth.make_model(d, ns, d + ns)
    X = rng.standard_normal((n, d), dtype=np.float32)
    clf = B200LogisticRegression(W, b, np.arange(ns), out_dtype=np.float32)
    check(f"auto d={d} C={ns} n={n}", clf, X, W, b)
import torch  # noqa: E402
for f16 in ("1", "0"):
    os.environ["LRS_GEMM_F16"] = f16
    for d, pitch, off, ns, n in ((64, 67, 1, 300, 300), (1022, 1022, 0, 100, 200), (617, 617, 0, 26, 150), (70, 72, 2, 40, 129)):
        W, b = synth.make_model(d, ns, d + ns)
        buf = rng.standard_normal(n * pitch + 8, dtype=np.float32)
        Xh = buf[off:off + n * pitch].reshape(n, pitch)[:, :d]
        Xt = torch.from_numpy(buf).cuda()[off:off + n * pitch].view(n, pitch)[:, :d]
        clf = B200LogisticRegression(W, b, np.arange(ns))
        out = clf.score_rows(Xt, proba=True, label=True)
        torch.cuda.synchronize()
        err = np.abs(out["proba"].cpu().numpy() - orc.predict_proba(Xh, W, b)).max()
        lab = clf.score_rows(Xt, label=True, maxprob=True)["label"].cpu().numpy()
        print(f"device rows d={d} pitch={pitch} off={off} C={ns} f16={f16}: max|dP| = {err:.2e}", flush=True)
        assert err < 1e-5 and lab.shape == (n,)
        clf.close()
del os.environ["LRS_GEMM_F16"]

os.environ["LRS_NARROW_BULK"] = "1"
os.environ["LRS_NARROW_BULK_MIN_ROWS"] = "1"
W, b = synth.make_model(32, 1, 3)
X = rng.standard_normal((3000, 32), dtype=np.float32)
check("narrow bulk ring", B200LogisticRegression(W, b, [0, 1], out_dtype=np.float32), X, W, b)

for env in ({"LRS_GEMM_F16": "0"}, {"LRS_GEMM_PAIR": "0"}, {"LRS_GEMM_SPLIT": "1"}, {"LRS_GEMM_PAIR": "0", "LRS_GEMM_MC": "1"}):
    os.environ.update(env)
    W, b = synth.make_model(128, 700, 9)
    X = rng.standard_normal((500, 128), dtype=np.float32)
    clf = B200LogisticRegression(W * 5, b, np.arange(700), out_dtype=np.float32)
    check(f"gemm {env}", clf, X, W * 5, b)
    for k in env:
        del os.environ[k]

from mlapi_b200 import refit  # noqa: E402
X = torch.randn(5000, 32, device="cuda")
y = (torch.rand(5000, device="cuda") < 0.5).int()
Wt = torch.randn(1, 32, device="cuda") * 0.1
bt = torch.zeros(1, device="cuda")
out = torch.empty(1 + 33, dtype=torch.float64, device="cuda")
refit.loss_grad_sums(X, y, Wt, bt, out)
L = _lib.load()
out2 = torch.empty_like(out)
rc = L.lr_loss_grad_static(C.c_void_p(X.data_ptr()), C.c_void_p(y.data_ptr()), 5000, 32, C.c_void_p(Wt.data_ptr()),
                           C.c_void_p(bt.data_ptr()), 1, 32, C.c_void_p(out2.data_ptr()), None)
torch.cuda.synchronize()
assert rc == 0 and torch.allclose(out, out2, rtol=1e-5, atol=1e-6)
print("refit + static refit ok", flush=True)
for n, d, rows in ((3000, 300, 1), (2000, 64, 20)):
    Xg = torch.randn(n, d, device="cuda")
    yg = torch.randint(0, max(rows, 2), (n,), device="cuda", dtype=torch.int32)
    Wb = np.random.default_rng(1).standard_normal((rows, d + 1)) * 0.1
    og = torch.empty(1 + rows * (d + 1), dtype=torch.float64, device="cuda")
    refit.loss_grad_sums_general(Xg, yg, Wb, og, block_rows=1024)
    torch.cuda.synchronize()
    assert torch.isfinite(og).all()
print("general refit (lr_weighted_row_sum, 3xTF32 GEMMs) ok", flush=True)
print("ALL OK")
