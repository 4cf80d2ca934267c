"""Throughput of lr_score across model shapes (d, C): which kernel family the dispatcher picks and how
close to the HBM / tensor roofline it lands.  Looks for cliffs between the benchmarked configs.

Every shape gets about 4 GB of X (resident, >> L2), P is written (the predict_proba contract) and the
label too.  Algorithmic bytes per row = 4 d + 4 max(C, 2) + 4; flops per row = 2 d C.

    python tools/shape_sweep.py [--gb 4] [--out gpurun_out/shape_sweep.json]
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlapi_b200 import B200LogisticRegression, _lib  # noqa: E402

PATH_NAMES = {1: "narrow", 2: "wide", 3: "gemm", 4: "generic"}

BINARY_D = [4, 5, 8, 16, 30, 32, 33, 64, 100, 101, 128, 256, 257, 512, 1000, 1001, 1024, 2048, 2049, 4096, 5001, 20000,
            100000, 100001]
MULTI = [(4, 3), (16, 4), (30, 3), (33, 4), (256, 3), (1000, 3), (54, 7), (64, 7), (784, 6), (64, 10), (61, 10), (784, 10),
         (785, 10), (3072, 10), (4096, 10), (8000, 10), (8000, 3), (1024, 12), (128, 16), (784, 16), (32, 20), (54, 20), (16, 26), (64, 32), (256, 32), (64, 50), (300, 50), (96, 96), (512, 100), (512, 101),
         (1024, 128), (4096, 64), (768, 255), (768, 256), (1024, 300), (1024, 1000), (1024, 1001), (1022, 1000), (128, 1000), (2048, 2000)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gb", type=float, default=4.0)
    ap.add_argument("--out", default="")
    ap.add_argument("--only", default="", help="comma-separated d:C pairs")
    args = ap.parse_args()
    L = _lib.load()
    total = int(args.gb * (1 << 30) / 4)
    buf = torch.empty(total, device="cuda")
    for lo in range(0, total, 1 << 27):
        buf[lo:lo + (1 << 27)].normal_()
    shapes = [(d, 1) for d in BINARY_D] + MULTI
    if args.only:
        shapes = [tuple(int(v) for v in s.split(":")) for s in args.only.split(",")]
    st = torch.cuda.current_stream().cuda_stream
    out = []
    for d, c in shapes:
        pc = max(c, 2)
        n = min(total // d, int(6e9 // (4 * pc)))          # keep P under 6 GB
        X = buf[: n * d].view(n, d)
        rng = np.random.default_rng(d * 7919 + c)
        W = rng.standard_normal((c, d)) / np.sqrt(d)
        b = rng.standard_normal(c)
        clf = B200LogisticRegression(W, b, list(range(pc)))
        rep = clf._replicas[0]
        P = torch.empty((n, pc), device="cuda")
        lab = torch.empty(n, dtype=torch.int32, device="cuda")
        path = C.c_int32(0)
        _lib.check(L.lr_model_select_path(rep.handle, C.c_void_p(X.data_ptr()), n, d, C.byref(path)))

        def step():
            _lib.check(L.lr_score(rep.handle, C.c_void_p(X.data_ptr()), n, d, C.c_void_p(P.data_ptr()), None,
                                  C.c_void_p(lab.data_ptr()), None, C.c_void_p(st)))

        for _ in range(2):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record()
        for _ in range(reps):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        rec = {"d": d, "C": c, "rows": n, "path": PATH_NAMES.get(path.value, path.value), "ms": round(ms, 4),
               "GBps": round(n * (4 * d + 4 * pc + 4) / ms / 1e6, 1),
               "TFLOPs": round(2.0 * n * d * c / ms / 1e9, 2), "rows_per_s": round(n / ms * 1e3)}
        print(json.dumps(rec), flush=True)
        out.append(rec)
        del P, lab
        clf.close()
    if args.out:
        with open(args.out, "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
