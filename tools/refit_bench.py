#!/usr/bin/env python
"""BASELINE cfg 4 (second half): batch-gradient refit at the cfg-1 shape.  Times the fused
loss+gradient pass (lr_loss_grad, one pass over X per L-BFGS evaluation) plus the gradient
all-reduce, and runs a short fit to convergence.

    python tools/refit_bench.py [--rows 100000000] [--iters 20]
    python -m torch.distributed.run --nproc-per-node N ... tools/refit_bench.py    (row shards, NCCL all-reduce)
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=100_000_000)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--fit-rows", type=int, default=4_000_000)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import __graft_entry__
    from mlapi_b200 import refit, sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        __graft_entry__.build()
    if world > 1:
        dist.barrier()
    dev = torch.device("cuda", local)
    d, n = 32, args.rows
    torch.manual_seed(10 + rank)
    X = torch.empty((n, d), device=dev)
    for lo in range(0, n, 1 << 21):
        X[lo:lo + (1 << 21)].normal_()
    gen = torch.Generator(device=dev)
    gen.manual_seed(7)                                    # the same generating model on every rank
    w_true = torch.randn(d, device=dev, generator=gen) / d ** 0.5
    y = (torch.rand(n, device=dev) < torch.sigmoid(X @ w_true * 3 + 0.3)).int()
    W = torch.zeros((1, d), device=dev)
    b = torch.zeros((1,), device=dev)
    out = torch.empty(1 + d + 1, dtype=torch.float64, device=dev)

    def step():
        refit.loss_grad_sums(X, y, W, b, out)
        sharding.allreduce_sum(out)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.iters):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = sharding.max_over_ranks(e0.elapsed_time(e1) / args.iters, dev)

    # a real fit on a bounded sample (host L-BFGS + one device pass and one all-reduce per evaluation)
    nf = min(n, args.fit_rows)
    t0 = time.perf_counter()
    res = refit.fit(X[:nf], y[:nf].cpu().numpy(), classes=np.array([0, 1]))
    fit_s = time.perf_counter() - t0
    if rank == 0:
        bytes_per_row = d * 4 + 4
        print(json.dumps({
            "workload": f"refit loss+gradient pass, binary d={d}, {n} rows per GPU, {world} GPU(s)",
            "ms_per_pass": ms, "rows_per_s": world * n / (ms * 1e-3),
            "hbm_gbs_per_gpu": bytes_per_row * n / (ms * 1e-3) / 1e9,
            "allreduce_bytes": int(out.numel() * 8), "collective": "nccl all_reduce(sum) float64" if world > 1 else "none",
            "fit": {"rows_per_gpu": nf, "n_iter": int(res.n_iter_[0]), "n_evals": res.n_evals_, "seconds": fit_s,
                    "final_loss": res.final_loss_, "coef_cosine_with_truth": float(
                        np.dot(res.coef_[0], w_true.cpu().numpy()) /
                        (np.linalg.norm(res.coef_[0]) * float(w_true.norm())))},
        }))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
