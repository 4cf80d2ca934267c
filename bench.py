#!/usr/bin/env python
"""bench.py -- predict_proba throughput of the B200 scoring engine (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg1|cfg0|cfg2|cfg3] [--impl reference]

A "step" is one pass of the hot path (lr_score: decision function + sigmoid/softmax, proba out) over
one resident batch of synthetic rows.  Default workload = BASELINE.json configs[1]: binary LR, d=32,
100M fp32 rows per GPU (12.8 GB, far larger than the 126 MB L2, so no flush is needed between
steps).  Rows are independent, so N GPUs = N row shards, one process per GPU, no collective on the
data path (`"scaling": "weak"`: every rank scores its own 100M rows).

Prints ONE JSON line (rank 0).  `value` is device-resident throughput; `e2e` is the same metric
through lr_score_host with pinned HOST buffers (H2D of X and D2H of P inside the timed region);
`roofline` is the dominant kernel against the measured HBM peak; `cpu_baseline` is scikit-learn's own
predict_proba (the reference's implementation of the path) on this box's host cores.
`--impl reference` times only that CPU implementation and prints the same line shape.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "predict_proba_rows_per_sec"
UNIT = "rows/s"

WORKLOADS = {
    # name: (description, n_rows per GPU, d, n_scores, algorithmic bytes/row, flops/row)
    "cfg1": ("binary LR d=32, 100M synthetic fp32 rows per GPU, fused GEMV+sigmoid (BASELINE configs[1])",
             100_000_000, 32, 1, 32 * 4 + 2 * 4, 2 * 32),
    "cfg0": ("notebook iris model d=4 C=3, 10k-row synthetic batch (BASELINE configs[0])",
             10_000, 4, 3, 4 * 4 + 3 * 4, 2 * 4 * 3),
    "cfg2": ("wide binary LR d=100000, resident chunk of 80k rows per GPU (2M rows = 25 chunk passes; "
             "BASELINE configs[2])", 80_000, 100_000, 1, 100_000 * 4 + 8, 2 * 100_000),
    "cfg3": ("multinomial LR d=1024 C=1000, resident chunk of 4M rows per GPU (20M rows = 5 chunk "
             "passes; BASELINE configs[3])", 4_000_000, 1024, 1000, 1024 * 4 + 1000 * 4, 2 * 1024 * 1000),
}
CPU_SAMPLE_ROWS = {"cfg1": 4_000_000, "cfg0": 10_000, "cfg2": 4_000, "cfg3": 40_000}


def make_model(d, ns, seed):
    rng = np.random.default_rng(seed)
    return rng.standard_normal((ns, d)) / np.sqrt(d), rng.standard_normal(ns)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            p = json.load(fh)
        return dict(hbm=float(p["hbm_gbs"]), bf16=float(p.get("bf16_tflops", 1590.0)),
                    bf16_sustained=float(p.get("bf16_tflops_sustained", 1400.0)), src="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            pass

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_reference(workload, threads=None, reps=None, budget_s=20.0):
    """scikit-learn's LogisticRegression.predict_proba -- the reference's own implementation of the
    path (REF/main.py:22) -- on a bounded sample of the workload, on the host cores."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import synth
    from threadpoolctl import threadpool_info

    _, _, d, ns, _, _ = WORKLOADS[workload]
    n = CPU_SAMPLE_ROWS[workload]
    W, b = make_model(d, ns, 2)
    clf = synth.sklearn_model(W, b)
    X = np.random.default_rng(1).standard_normal((n, d), dtype=np.float32)
    cores = len(os.sched_getaffinity(0))
    times = []
    from threadpoolctl import threadpool_limits
    with threadpool_limits(limits=cores):        # torchrun exports OMP_NUM_THREADS=1: undo it here
        clf.predict_proba(X[: max(1, n // 8)])
        t_end = time.perf_counter() + budget_s
        while len(times) < (reps or 3) or (reps is None and time.perf_counter() < t_end and len(times) < 10):
            t0 = time.perf_counter()
            clf.predict_proba(X)
            times.append(time.perf_counter() - t0)
    blas = [f"{i.get('internal_api')}:{i.get('num_threads')}" for i in threadpool_info()]
    return dict(rows=n, times=times, best=min(times), median=float(np.median(times)), cores=cores,
                sample=f"{n} rows x d={d} fp32 (same distribution as the GPU batch), sklearn "
                       f"{__import__('sklearn').__version__} predict_proba, {len(times)} reps, median; "
                       f"BLAS threads {','.join(blas) or 'n/a'}; os.cpu_count={os.cpu_count()}")


def bind_to_gpu_numa_node(gpu_index):
    """Pin this rank's host threads to the CPUs next to its GPU before any pinned allocation, so the
    e2e staging buffers are placed on the local NUMA node and every rank uses its own memory
    controller / PCIe root (with 8 ranks on one socket's memory the H2D copies contend)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        n_words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        cpus = {w * 64 + b for w, word in enumerate(mask) for b in range(64) if (word >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return 0


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    desc, n, d, ns, bpr, fpr = WORKLOADS[args.workload]
    rows = CPU_SAMPLE_ROWS[args.workload]
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import synth
    W, b = make_model(d, ns, 2)
    clf = synth.sklearn_model(W, b)
    X = np.random.default_rng(1).standard_normal((rows, d), dtype=np.float32)
    cores = len(os.sched_getaffinity(0))
    from threadpoolctl import threadpool_limits
    with threadpool_limits(limits=cores):        # all host threads, whatever the launcher exported
        for _ in range(args.warmup):
            clf.predict_proba(X)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            clf.predict_proba(X)
        dt = (time.perf_counter() - t0) / args.steps
    value = rows / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "n_features": d, "n_scores": ns,
                   "note": "reference = scikit-learn LogisticRegression.predict_proba on host cores; "
                           f"each step scores a bounded sample of {rows} rows"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": f"{rows} rows x d={d} fp32 per step, sklearn "
                                   f"{__import__('sklearn').__version__}, default BLAS threading"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="cfg1", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=0, help="override rows per GPU")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--outputs", default="proba", help="comma list of proba,decision,label,maxprob")
    ap.add_argument("--logit-scale", type=float, default=1.0,
                    help="multiply W by this (confident models stress the tensor path's re-scoring pass)")
    args = ap.parse_args()
    if args.impl == "b200":
        args.warmup = max(args.warmup, 3)          # timing rule: at least 3 untimed steps
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist

    import __graft_entry__
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    numa_cpus = bind_to_gpu_numa_node(local) if world > 1 else 0
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # keep stdout to the one JSON line
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        __graft_entry__.build()
    if world > 1:
        dist.barrier()
    from mlapi_b200 import _lib
    from mlapi_b200.estimator import B200LogisticRegression, PinnedArray
    L = _lib.load()

    desc, n, d, ns, bytes_per_row, flops_per_row = WORKLOADS[args.workload]
    if args.rows:
        n = args.rows
    n_out = max(ns, 2)
    W, b = make_model(d, ns, 2)
    W = W * args.logit_scale
    clf = B200LogisticRegression(W, b, np.arange(n_out), devices=[local])
    rep = clf._replicas[0]
    dev = torch.device("cuda", local)
    torch.manual_seed(1 + rank)
    X = torch.empty((n, d), dtype=torch.float32, device=dev)
    step_rows = max(1, (256 << 20) // (4 * d))
    for lo in range(0, n, step_rows):          # chunked fill keeps the RNG's temporaries small
        X[lo:lo + step_rows].normal_()
    want = set(args.outputs.split(","))
    P = torch.empty((n, n_out), dtype=torch.float32, device=dev) if "proba" in want else None
    Z = torch.empty((n,) if ns == 1 else (n, ns), dtype=torch.float32, device=dev) if "decision" in want else None
    Lb = torch.empty((n,), dtype=torch.int32, device=dev) if "label" in want else None
    Mp = torch.empty((n,), dtype=torch.float32, device=dev) if "maxprob" in want else None
    ptr = lambda t: None if t is None else C.c_void_p(t.data_ptr())  # noqa: E731
    stream = torch.cuda.current_stream(dev)

    def step():
        _lib.check(L.lr_score(rep.handle, C.c_void_p(X.data_ptr()), n, d, ptr(P), ptr(Z), ptr(Lb), ptr(Mp),
                              C.c_void_p(stream.cuda_stream)))

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    path = C.c_int32()
    _lib.check(L.lr_model_select_path(rep.handle, C.c_void_p(X.data_ptr()), n, d, C.byref(path)))
    for _ in range(args.warmup):
        step()
    sync_all()
    sampler = ClockSampler(local) if rank == 0 else None
    launches0 = L.lr_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    ev1.record(stream)
    sync_all()
    launches = L.lr_launch_count() - launches0
    ms_local = ev0.elapsed_time(ev1) / args.steps
    t = torch.tensor([ms_local], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    clocks = sampler.stop() if sampler else None
    assert clf.take_nonfinite() == 0 or os.environ.get("LRS_GEMM_DEBUG_SKIP")
    value = world * n / (ms * 1e-3)

    # ---- end to end through lr_score_host with pinned host buffers --------------------------------
    e2e = None
    if not args.no_e2e:
        import psutil
        n_e = n
        need = n_e * (d + n_out) * 4
        avail = psutil.virtual_memory().available // max(world, 1)
        while need > 0.5 * avail and n_e > 1:
            n_e //= 2
            need = n_e * (d + n_out) * 4
        xh = PinnedArray((n_e, d), np.float32)
        ph = PinnedArray((n_e, n_out), np.float32)
        th = torch.from_numpy(xh.array)
        th.copy_(X[:n_e])                         # D2H fill, outside the timed region
        torch.cuda.synchronize(dev)

        def e2e_step():
            _lib.check(L.lr_score_host(rep.handle, xh.array.ctypes.data_as(C.c_void_p), n_e, d,
                                       ph.array.ctypes.data_as(C.c_void_p), None, None, None))

        e2e_step()
        sync_all()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            e2e_step()                            # returns when P is in host memory
        dt_local = (time.perf_counter() - t0) / args.e2e_steps
        te = torch.tensor([dt_local], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        dt = float(te.item())
        # spot-check: host result equals the device-resident result for the same rows
        same = bool(np.array_equal(ph.array[:1000], P[:1000].cpu().numpy())) if P is not None else None
        e2e = {"value": world * n_e / dt, "unit": UNIT, "h2d_bytes_per_step": int(n_e * d * 4),
               "d2h_bytes_per_step": int(n_e * n_out * 4), "rows_per_step_per_gpu": n_e,
               "steps": args.e2e_steps, "ms_per_step": dt * 1e3,
               "api": "lr_score_host (C ABI), pinned host X and P, 3-stream chunked H2D/kernel/D2H",
               "host_cpus_bound_per_rank": numa_cpus,
               "matches_device_result": same}
        xh.free(); ph.free()

    if rank == 0:
        peaks = measured_peaks()
        extra_out = 4 * (("decision" in want) * ns + ("label" in want) + ("maxprob" in want))
        if "proba" not in want:
            extra_out -= 4 * n_out
        alg_bytes = (bytes_per_row + extra_out) * n
        tensor_bound = args.workload == "cfg3"
        if tensor_bound:
            achieved = flops_per_row * n / (ms * 1e-3) / 1e12
            f16 = os.environ.get("LRS_GEMM_F16", "1") != "0"
            peak = peaks["bf16"] if f16 else peaks["bf16"] / 2.0
            sustained = peaks["bf16_sustained"] if f16 else peaks["bf16_sustained"] / 2.0
            roof = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                    "frac": achieved / peak, "traffic": None,
                    "peak_source": (f"{peaks['src']} 16-bit tensor burst peak (cuBLAS bf16; the kernel issues fp16 MMAs)"
                                    if f16 else f"half of {peaks['src']} bf16 burst peak (TF32 rate)") +
                                   "; algorithmic flops 2*d*C per row counted once, the kernel executes 3 MMAs per "
                                   "product (hi.hi + lo.hi + hi.lo), so frac cannot exceed 1/3",
                    "executed_tflops": 3.0 * achieved,
                    "frac_of_sustained": achieved / sustained,
                    "sustained_note": "steps are tens of ms of tensor work under the 1 kW cap (sw_power_cap active): "
                                      "the sustained figure is the applicable ceiling",
                    "hbm_gbs": alg_bytes / (ms * 1e-3) / 1e9}
        else:
            achieved = alg_bytes / (ms * 1e-3) / 1e9
            roof = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm"], "unit": "GB/s",
                    "frac": achieved / peaks["hbm"], "traffic": None,
                    "peak_source": f"{peaks['src']} HBM copy bandwidth (MEASURED_PEAKS.json)" if peaks["src"] == "measured"
                    else "fallback 6650 GB/s (B200_PROFILING.md)",
                    "frac_of_nominal_8000": achieved / 8000.0,
                    "algorithmic_bytes_per_row": bytes_per_row + extra_out}
        tr = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tr):
            with open(tr) as fh:
                t_all = json.load(fh)
            ent = t_all.get(args.workload)
            if ent and ent.get("rows"):
                roof["traffic"] = ent["dram_bytes"] / ent["rows"] * n     # per launch, scaled by rows
                roof["traffic_source"] = ent.get("source")
        cpu = None
        if world == 1 and not args.no_cpu:
            c = cpu_reference(args.workload)
            cpu = {"value": c["rows"] / c["median"], "unit": UNIT, "cores": c["cores"], "kind": "reference",
                   "sample": c["sample"], "best_value": c["rows"] / c["best"]}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"workload": desc, "rows_per_gpu": n, "n_features": d, "n_scores": ns,
                       "outputs": sorted(want), "logit_scale": args.logit_scale, "kernel_path": _lib.PATH_NAMES.get(path.value, "?"),
                       "parallelism": f"row-shard x{world}, no collective",
                       "l2": "inputs larger than L2 (no flush needed)" if n * d * 4 > (512 << 20)
                       else "inputs fit in L2: latency figure, not a bandwidth figure"},
            "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
