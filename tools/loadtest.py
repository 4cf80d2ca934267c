#!/usr/bin/env python
"""BASELINE cfg 4: POST /predict under load (requests/s, p50/p99) against the B200 app, and -- when a
path to the reference's main.py is given and exists -- against the unmodified reference app.

    python tools/loadtest.py [--seconds 8] [--clients 4] [--concurrency 64] [--reference /root/reference/main.py]

Closed-loop load: `clients` processes x `concurrency` in-flight requests each over keep-alive
connections (httpx).  The server is uvicorn with one worker, as in REF/README.md:16.  Also reports
the micro-batcher alone (no HTTP) to separate engine capacity from Python/HTTP overhead.
"""
import argparse
import asyncio
import json
import multiprocessing as mp
import os
import pickle
import shutil
import socket
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def wait_port(port, timeout=120):
    t0 = time.time()
    while time.time() - t0 < timeout:
        try:
            socket.create_connection(("127.0.0.1", port), timeout=0.5).close()
            return True
        except OSError:
            time.sleep(0.2)
    return False


def client_proc(port, concurrency, seconds, seed, q):
    import httpx

    rng = np.random.default_rng(seed)
    bodies = [dict(sepal_length=float(a), sepal_width=float(b), petal_length=float(c), petal_width=float(d))
              for a, b, c, d in rng.uniform([4, 2, 1, 0.1], [8, 4.5, 7, 2.5], size=(512, 4))]
    lat = []

    async def worker(client, t_end, k):
        i = k
        while time.perf_counter() < t_end:
            t0 = time.perf_counter()
            r = await client.post("/predict", json=bodies[i % len(bodies)])
            lat.append(time.perf_counter() - t0)
            assert r.status_code == 200
            i += 7

    async def main():
        limits = httpx.Limits(max_connections=concurrency, max_keepalive_connections=concurrency)
        async with httpx.AsyncClient(base_url=f"http://127.0.0.1:{port}", limits=limits, timeout=30) as client:
            await client.post("/predict", json=bodies[0])
            t_end = time.perf_counter() + seconds
            await asyncio.gather(*[worker(client, t_end, k) for k in range(concurrency)])

    asyncio.run(main())
    q.put(lat)


def fast_client_proc(port, concurrency, seconds, seed, q):
    """Raw-socket keep-alive client: pre-built request bytes, just enough HTTP parsing to find the end of the
    response.  httpx spends ~0.5 ms of client CPU per request, which made the load generator the bottleneck."""
    rng = np.random.default_rng(seed)
    reqs = []
    for a, b, c, d in rng.uniform([4, 2, 1, 0.1], [8, 4.5, 7, 2.5], size=(512, 4)):
        body = json.dumps(dict(sepal_length=float(a), sepal_width=float(b), petal_length=float(c), petal_width=float(d))).encode()
        reqs.append(b"POST /predict HTTP/1.1\r\nHost: 127.0.0.1\r\nContent-Type: application/json\r\nContent-Length: "
                    + str(len(body)).encode() + b"\r\n\r\n" + body)
    lat = []
    errors = [0]

    async def worker(t_end, k):
        reader, writer = await asyncio.open_connection("127.0.0.1", port)
        i = k
        while time.perf_counter() < t_end:
            t0 = time.perf_counter()
            writer.write(reqs[i % len(reqs)])
            head = await reader.readuntil(b"\r\n\r\n")
            n = 0
            for line in head.split(b"\r\n"):
                if line[:15].lower() == b"content-length:":
                    n = int(line[15:])
            if n:
                await reader.readexactly(n)
            if not head.startswith(b"HTTP/1.1 200"):
                errors[0] += 1
            lat.append(time.perf_counter() - t0)
            i += 7
        writer.close()

    async def main():
        t_end = time.perf_counter() + seconds
        await asyncio.gather(*[worker(t_end, k) for k in range(concurrency)])

    try:
        import uvloop
        uvloop.install()
    except ImportError:
        pass
    asyncio.run(main())
    assert errors[0] == 0, f"{errors[0]} non-200 answers"
    q.put(lat)


def run_load(port, clients, concurrency, seconds, fast=False):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    target = fast_client_proc if fast else client_proc
    procs = [ctx.Process(target=target, args=(port, concurrency, seconds, 100 + i, q)) for i in range(clients)]
    t0 = time.perf_counter()
    [p.start() for p in procs]
    lats = []
    for _ in procs:
        lats += q.get(timeout=seconds + 120)
    wall = time.perf_counter() - t0
    [p.join(timeout=30) for p in procs]
    lat = np.array(lats)
    return {"requests": int(lat.size), "req_per_s": float(lat.size / seconds), "p50_ms": float(np.percentile(lat, 50) * 1e3),
            "p99_ms": float(np.percentile(lat, 99) * 1e3), "mean_ms": float(lat.mean() * 1e3), "wall_s": wall,
            "clients": clients, "concurrency_per_client": concurrency, "client": "raw sockets" if fast else "httpx"}


def serve(module_app, cwd, env, port, workers=1):
    cmd = [sys.executable, "-m", "uvicorn", module_app, "--port", str(port), "--log-level", "warning"]
    if workers > 1:
        cmd += ["--workers", str(workers)]
    return subprocess.Popen(cmd, cwd=cwd, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)


def batcher_only(model_path, n=200_000):
    from mlapi_b200 import load_model
    from mlapi_b200.batcher import MicroBatcher, engine_score_fns
    from mlapi_b200.estimator import PinnedArray
    clf = load_model(model_path)
    pins = []

    def alloc(shape, dtype):
        pins.append(PinnedArray(shape, dtype))
        return pins[-1].array

    mb = MicroBatcher(engine_score_fns(clf), n_features=4, max_batch=4096, max_wait_us=200, alloc=alloc)
    rows = [tuple(map(float, r)) for r in np.random.default_rng(0).uniform(0, 8, size=(n, 4))]

    async def burst():
        return await asyncio.gather(*[mb.submit(r) for r in rows])

    asyncio.run(burst())
    t0 = time.perf_counter()
    asyncio.run(burst())
    dt = time.perf_counter() - t0
    st = mb.stats()
    mb.close()
    return {"requests": n, "req_per_s": n / dt, "mean_batch": st["mean_batch"], "note": "asyncio futures -> MicroBatcher -> lr_score_host, no HTTP"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=8.0)
    ap.add_argument("--clients", type=int, default=4)
    ap.add_argument("--concurrency", type=int, default=64)
    ap.add_argument("--reference", default="", help="path of the unmodified reference main.py (where it exists)")
    ap.add_argument("--reference-restated", action="store_true",
                    help="serve tools/ref_handler_app.py: the reference handler's per-request work, restated")
    ap.add_argument("--skip-b200", action="store_true")
    ap.add_argument("--workers", type=int, default=1, help="uvicorn worker processes (the reference's README uses 1)")
    ap.add_argument("--engine", action="store_true",
                    help="one scoring engine process + --workers HTTP fronts (mlapi_b200.serve) instead of a model per worker")
    ap.add_argument("--devices", default="0")
    ap.add_argument("--fast-client", action="store_true")
    ap.add_argument("--no-batcher-only", action="store_true")
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    import synth
    clf, _ = synth.fit_notebook_model()
    tmp = tempfile.mkdtemp()
    model_path = os.path.join(tmp, "LRClassifier.pkl")
    with open(model_path, "wb") as fh:
        pickle.dump(clf, fh)
    result = {"host_cores": len(os.sched_getaffinity(0)), "seconds": args.seconds}
    load = lambda port, c, k, s: run_load(port, c, k, s, fast=args.fast_client)  # noqa: E731

    if not args.skip_b200:
        port = free_port()
        procs = []
        try:
            if args.engine:
                from mlapi_b200 import serve as lr_serve
                engine, fronts, env = lr_serve.start(model_path, args.devices, args.workers, port)
                procs = [fronts, engine]
            else:
                env = dict(os.environ, LR_MODEL_PATH=model_path, LR_DEVICES=args.devices, PYTHONPATH=ROOT)
                procs = [serve("mlapi_b200.app:app", ROOT, env, port, args.workers)]
            assert wait_port(port, 300), "B200 app did not start"
            time.sleep(3.0 if args.workers > 1 else 0.5)
            load(port, 1, 8, 1.0)                                   # every worker has connected / warmed up
            import httpx
            before = httpx.get(f"http://127.0.0.1:{port}/batcher_stats").json()
            result["b200_app"] = load(port, args.clients, args.concurrency, args.seconds)
            result["b200_app"]["uvicorn_workers"] = args.workers
            result["b200_app"]["devices"] = args.devices
            result["b200_app"]["layout"] = ("one engine process (all devices, one micro-batcher) + HTTP front workers"
                                            if args.engine else "a model and a micro-batcher in every worker")
            time.sleep(1.5)                                         # the engine rewrites its statistics once a second
            after = httpx.get(f"http://127.0.0.1:{port}/batcher_stats").json()
            result["b200_app"]["batcher_before_load"] = before
            result["b200_app"]["batcher"] = after
            if "requests" in after and "requests" in before and after.get("batches", 0) > before.get("batches", 0):
                result["b200_app"]["mean_batch_under_load"] = ((after["requests"] - before["requests"]) /
                                                               (after["batches"] - before["batches"]))
            # latency at light load: few requests in flight, so queueing does not hide the service time
            result["b200_app_light_load"] = load(port, 2, 4, min(args.seconds, 5.0))
        finally:
            for p in procs:
                p.terminate()
            for p in procs:
                p.wait(timeout=30)
        if not args.no_batcher_only:
            result["b200_batcher_only"] = batcher_only(model_path)

    ref_runs = []
    if args.reference and os.path.exists(args.reference):
        shutil.copy(args.reference, os.path.join(tmp, "main.py"))
        ref_runs.append(("reference_app", "main:app", tmp,
                         "unmodified REF/main.py (per-request unpickle, blocking handler), uvicorn 1 worker"))
    if args.reference_restated:
        shutil.copy(os.path.join(ROOT, "tools", "ref_handler_app.py"), os.path.join(tmp, "ref_handler_app.py"))
        ref_runs.append(("reference_restated_app", "ref_handler_app:app", tmp,
                         "tools/ref_handler_app.py: the per-request work of REF/main.py:16-27 restated (per-request "
                         "unpickle, predict + predict_proba on the event loop), uvicorn 1 worker"))
    for key, target, cwd, note in ref_runs:
        port = free_port()
        srv = serve(target, cwd, dict(os.environ), port, 1)
        try:
            assert wait_port(port), f"{key} did not start"
            time.sleep(0.5)
            load(port, 1, 2, 1.5)                                   # first requests import scikit-learn's estimator code
            result[key] = load(port, args.clients, args.concurrency, args.seconds)
            result[key]["uvicorn_workers"] = 1
            result[key + "_light_load"] = load(port, 2, 4, min(args.seconds, 5.0))
            result[key]["note"] = note
        finally:
            srv.terminate()
            srv.wait(timeout=20)
    print(json.dumps(result, indent=1))
    if args.out:
        with open(args.out, "w") as fh:
            json.dump(result, fh, indent=1)


if __name__ == "__main__":
    main()
