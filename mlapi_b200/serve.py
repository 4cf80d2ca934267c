"""Start the scoring engine process and N uvicorn front workers in one go.

    python -m mlapi_b200.serve --model LRClassifier.pkl --devices 0,1,2,3,4,5,6,7 --workers 8 --port 8000

The reference's run line is `uvicorn main:app` (REF/README.md:16): one process that does everything.  Here the
GPUs belong to one engine process (`mlapi_b200.engine_link`: one model replica per device behind ONE
micro-batcher) and the HTTP side scales by worker processes that only parse requests and forward rows, so
the batches the devices see are filled by all of the traffic.
"""
from __future__ import annotations

import argparse
import os
import signal
import subprocess
import sys
import tempfile
import time


def start(model: str, devices: str, workers: int, port: int, host: str = "127.0.0.1", max_batch: int = 4096,
          max_wait_us: float = 200.0, log_level: str = "warning"):
    """Returns (engine Popen, uvicorn Popen, env) -- used by tools/loadtest.py as well."""
    tmp = tempfile.mkdtemp(prefix="lr_engine_")
    sock = os.path.join(tmp, "engine.sock")
    stats = os.path.join(tmp, "stats.json")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, LR_ENGINE_SOCKET=sock, LR_ENGINE_STATS=stats,
               PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", ""))
    engine = subprocess.Popen([sys.executable, "-m", "mlapi_b200.engine_link", "--model", model, "--socket", sock,
                               "--devices", devices, "--max-batch", str(max_batch), "--max-wait-us", str(max_wait_us),
                               "--stats", stats], env=env, cwd=root)
    t0 = time.time()
    while not os.path.exists(sock):
        if engine.poll() is not None:
            raise RuntimeError("scoring engine exited during start-up")
        if time.time() - t0 > 300:
            engine.terminate()
            raise RuntimeError("scoring engine did not come up")
        time.sleep(0.1)
    cmd = [sys.executable, "-m", "uvicorn", "mlapi_b200.app:app", "--host", host, "--port", str(port), "--log-level",
           log_level]
    if workers > 1:
        cmd += ["--workers", str(workers)]
    fronts = subprocess.Popen(cmd, env=env, cwd=root)
    return engine, fronts, env


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default=os.environ.get("LR_MODEL_PATH", "LRClassifier.pkl"))
    ap.add_argument("--devices", default=os.environ.get("LR_DEVICES", "0"))
    ap.add_argument("--workers", type=int, default=4)
    ap.add_argument("--port", type=int, default=8000)
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--max-batch", type=int, default=4096)
    ap.add_argument("--max-wait-us", type=float, default=200.0)
    a = ap.parse_args()
    engine, fronts, _ = start(a.model, a.devices, a.workers, a.port, a.host, a.max_batch, a.max_wait_us, "info")

    def stop(*_):
        for p in (fronts, engine):
            if p.poll() is None:
                p.terminate()
    signal.signal(signal.SIGINT, stop)
    signal.signal(signal.SIGTERM, stop)
    try:
        while engine.poll() is None and fronts.poll() is None:
            time.sleep(0.5)
    finally:
        stop()


if __name__ == "__main__":
    main()
