"""Share of rows the tensor-core path hands to its fp32 re-scoring pass, and the step time, per shape and layout
(random model, N(0,1) rows: tools/shape_sweep.py's data).  Usage: python tools/rescoring_rate.py [d:C ...]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlapi_b200 import B200LogisticRegression, _lib  # noqa: E402

shapes = [tuple(int(v) for v in a.split(":")) for a in sys.argv[1:]] or [(4096, 64), (1024, 300), (1024, 128), (768, 256),
                                                                         (2048, 2000), (1024, 1000), (512, 100)]
L = _lib.load()
st = torch.cuda.current_stream().cuda_stream
for d, c in shapes:
    n = min((1 << 30) // d, int(3e9 // (4 * c)))
    X = torch.randn((n, d), device="cuda")
    rng = np.random.default_rng(d * 7919 + c)
    W = rng.standard_normal((c, d)) / np.sqrt(d)
    b = rng.standard_normal(c)
    for split in ("0", "1"):
        os.environ["LRS_GEMM_SPLIT"] = split
        clf = B200LogisticRegression(W, b, list(range(c)))
        rep = clf._replicas[0]
        P = torch.empty((n, c), device="cuda")
        lab = torch.empty(n, dtype=torch.int32, device="cuda")

        def step():
            _lib.check(L.lr_score(rep.handle, C.c_void_p(X.data_ptr()), n, d, C.c_void_p(P.data_ptr()), None,
                                  C.c_void_p(lab.data_ptr()), None, C.c_void_p(st)))
        step()
        torch.cuda.synchronize()
        clf.take_rescored()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            step()
        e1.record()
        torch.cuda.synchronize()
        r = clf.take_rescored() / 3
        print(f"d={d} C={c} layout={'accurate' if split == '1' else 'fast'}: {e0.elapsed_time(e1) / 3:.3f} ms, "
              f"re-scored {r / n:.3%} of {n} rows", flush=True)
        clf.close()
        del P, lab
    del X
