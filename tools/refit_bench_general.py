#!/usr/bin/env python
"""BASELINE cfg 4, refit at the cfg-2 and cfg-3 shapes (reduced n): one loss+gradient evaluation through
`refit.Evaluator` -- forward pass by the scoring engine, residual^T X by lr_weighted_row_sum (binary) or
lr_grad_gemm (multinomial, tcgen05) -- and how the time splits between the two.

    python tools/refit_bench_general.py [--out gpurun_out/refit_general.json]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(fn, reps):
    import torch
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    import __graft_entry__
    __graft_entry__.build()
    from mlapi_b200 import B200LogisticRegression, refit
    dev = torch.device("cuda", 0)
    results = []
    for name, n, d, rows in (("cfg2 shape: binary d=100000", 40_000, 100_000, 1),
                             ("cfg3 shape: d=1024, 1000 classes", 1_000_000, 1024, 1000)):
        X = torch.empty((n, d), device=dev)
        for lo in range(0, n, 1 << 14):
            X[lo:lo + (1 << 14)].normal_()
        y = torch.randint(0, 2 if rows == 1 else rows, (n,), device=dev, dtype=torch.int32)
        rng = np.random.default_rng(1)
        Wb = np.concatenate([rng.standard_normal((rows, d)) / np.sqrt(d), rng.standard_normal((rows, 1))], axis=1)
        ev = refit.Evaluator(X, y, rows, force_general=True)
        ms_pass = timed(lambda: ev.sums(Wb), 3)
        ev.close()
        # the forward pass alone, same blocks
        est = B200LogisticRegression(Wb[:, :d], Wb[:, d], np.arange(max(rows, 2)), devices=[0])
        blk = max(1024, min(n, (1 << 28) // max(rows, d // 8, 2)))

        def fwd():
            for lo in range(0, n, blk):
                est._score_device(X[lo:lo + blk], True, True, rows > 1, rows > 1)
        ms_fwd = timed(fwd, 3)
        est.close()
        rec = {"workload": f"refit loss+gradient evaluation, {name}, {n} rows, 1 GPU", "ms_per_evaluation": round(ms_pass, 3),
               "ms_forward_pass_only": round(ms_fwd, 3), "rows_per_s": round(n / ms_pass * 1e3),
               "allreduce_bytes": 8 * (1 + rows * (d + 1)),
               "note": "includes lr_model_update (the float64 iterate -> device copies) each evaluation; the plan "
                       "(X split into fp16 halves) is made once per fit and not timed"}
        print(json.dumps(rec), flush=True)
        results.append(rec)
        del X, y
        torch.cuda.empty_cache()
    if args.out:
        with open(args.out, "w") as f:
            json.dump(results, f, indent=1)


if __name__ == "__main__":
    main()
