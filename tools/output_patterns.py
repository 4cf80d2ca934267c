"""Step time of lr_score per output pattern (which of P, decision values, labels, max probabilities the caller
takes), across the kernel families: looks for patterns that cost more than the full P matrix.
Usage: python tools/output_patterns.py [--gb 2]"""
import argparse
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlapi_b200 import B200LogisticRegression, _lib  # noqa: E402

SHAPES = [(32, 1), (4, 3), (64, 10), (1024, 12), (54, 20), (64, 32), (300, 50), (1024, 128), (512, 300), (1024, 1000),
          (100000, 1), (64, 5000)]
PATTERNS = ["proba", "label", "label,maxprob", "decision", "proba,decision,label,maxprob"]
NAMES = {1: "narrow", 2: "wide", 3: "gemm", 4: "generic"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gb", type=float, default=2.0)
    args = ap.parse_args()
    L = _lib.load()
    total = int(args.gb * (1 << 30) / 4)
    buf = torch.randn(total, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    print("| d | C | path | " + " | ".join(PATTERNS) + " |")
    print("|---|---|---|" + "---|" * len(PATTERNS))
    for d, c in SHAPES:
        pc = max(c, 2)
        n = min(total // d, int(2e9 // (4 * pc)))
        X = buf[: n * d].view(n, d)
        rng = np.random.default_rng(d * 7919 + c)
        clf = B200LogisticRegression(rng.standard_normal((c, d)) / np.sqrt(d), rng.standard_normal(c), list(range(pc)))
        rep = clf._replicas[0]
        P = torch.empty((n, pc), device="cuda")
        Z = torch.empty((n, c), device="cuda")
        lab = torch.empty(n, dtype=torch.int32, device="cuda")
        mp = torch.empty(n, device="cuda")
        path = C.c_int32(0)
        _lib.check(L.lr_model_select_path(rep.handle, C.c_void_p(X.data_ptr()), n, d, C.byref(path)))
        cells = []
        for pat in PATTERNS:
            want = pat.split(",")
            ptr = lambda t, k: C.c_void_p(t.data_ptr()) if k in want else None  # noqa: E731

            def step():
                _lib.check(L.lr_score(rep.handle, C.c_void_p(X.data_ptr()), n, d, ptr(P, "proba"), ptr(Z, "decision"),
                                      ptr(lab, "label"), ptr(mp, "maxprob"), C.c_void_p(st)))
            for _ in range(2):
                step()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(4):
                step()
            e1.record()
            torch.cuda.synchronize()
            cells.append(f"{e0.elapsed_time(e1) / 4:.3f}")
        print(f"| {d} | {c} | {NAMES.get(path.value)} | " + " | ".join(cells) + f" |  ({n} rows, ms)", flush=True)
        del P, Z, lab, mp
        clf.close()


if __name__ == "__main__":
    main()
