"""FastAPI application with the reference's surface (REF/main.py:8-39), served by the B200 engine.

    uvicorn mlapi_b200.app:app            # the reference's run line is `uvicorn main:app` (REF/README.md:16)

Same routes, same request/response schema:
  POST /predict  body {"sepal_length": float, "sepal_width": float, "petal_length": float,
                       "petal_width": float}  ->  {"prediction": <label>, "probability": <max proba>}
  POST /files/   multipart `file` + form `token` (CSV echo; not on the scoring path, unaccelerated)
What changes underneath: the pickle is loaded ONCE at start-up into device-resident fp32 buffers
(the reference unpickles on every request, REF/main.py:19), and requests are coalesced by
`MicroBatcher` into device batches instead of being scored one by one on the event loop.

Environment: LR_MODEL_PATH (default 'LRClassifier.pkl', the reference's hard-coded path),
LR_DEVICES (comma list, default '0'), LR_MAX_BATCH (4096), LR_MAX_WAIT_US (200).
LR_ENGINE_SOCKET: when set, this process is an HTTP front only -- it owns no model and forwards its rows to
the scoring engine process listening there (`python -m mlapi_b200.engine_link`, or `python -m mlapi_b200.serve`
which starts both), so that several uvicorn workers fill the same device batches (engine_link.py).
"""
from __future__ import annotations

import math
import os
import threading
from contextlib import asynccontextmanager
from io import StringIO
from typing import Optional

from fastapi import FastAPI, File, Form
from fastapi.responses import JSONResponse
from pydantic import BaseModel

from .batcher import MicroBatcher, engine_score_fns


class IrisSpecies(BaseModel):
    """Request schema, field for field as REF/main.py:10-14."""
    sepal_length: float
    sepal_width: float
    petal_length: float
    petal_width: float


def create_app(model=None, model_path: Optional[str] = None, devices=None, max_batch: Optional[int] = None,
               max_wait_us: Optional[float] = None, score_fns=None, classes=None,
               engine_socket: Optional[str] = None) -> FastAPI:
    """Build the app.  `model` may be a ready `B200LogisticRegression` (tests); otherwise the pickle
    at `model_path` is loaded at start-up.  `score_fns` (+ `classes`) lets host-logic tests inject a scorer.
    `engine_socket` (or LR_ENGINE_SOCKET): be a front of the engine process listening there instead."""
    state = {}
    if engine_socket is None and model is None and score_fns is None:
        engine_socket = os.environ.get("LR_ENGINE_SOCKET") or None

    @asynccontextmanager
    async def lifespan(app: FastAPI):
        if engine_socket:
            from .engine_link import EngineClient
            client = await EngineClient(engine_socket).connect()
            state["client"] = client
            state["classes"] = client.classes
            state["submit"] = client.submit
            yield
            await client.close()
            return
        clf = model
        if clf is None and score_fns is None:
            from .estimator import load_model
            path = model_path or os.environ.get("LR_MODEL_PATH", "LRClassifier.pkl")
            devs = devices if devices is not None else [int(x) for x in os.environ.get("LR_DEVICES", "0").split(",")]
            clf = load_model(path, devices=devs)
        fns = score_fns if score_fns is not None else engine_score_fns(clf)
        alloc = None
        if score_fns is None:
            from .estimator import PinnedArray
            pins = []

            def alloc(shape, dtype):           # pinned staging rows: the H2D copy is a plain DMA
                pa = PinnedArray(shape, dtype)
                pins.append(pa)
                return pa.array
            state["pins"] = pins
        state["classes"] = list(clf.classes_) if clf is not None else list(classes)
        state["bulk_fn"] = (engine_score_fns(clf)[0] if score_fns is None else score_fns[0])
        state["bulk_lock"] = threading.Lock()    # the closure owns one set of output buffers
        state["batcher"] = MicroBatcher(
            fns, n_features=4,
            max_batch=max_batch or int(os.environ.get("LR_MAX_BATCH", "4096")),
            max_wait_us=max_wait_us if max_wait_us is not None else float(os.environ.get("LR_MAX_WAIT_US", "200")),
            alloc=alloc)
        app.state.batcher = state["batcher"]
        state["submit"] = state["batcher"].submit
        yield
        state["batcher"].close()

    app = FastAPI(lifespan=lifespan)

    @app.post("/predict")
    async def predict_species(iris: IrisSpecies):
        # REF/main.py:18-22: dict -> [[sl, sw, pl, pw]] -> predict + predict_proba.max()
        row = (iris.sepal_length, iris.sepal_width, iris.petal_length, iris.petal_width)
        # finite as float32, which is what the engine computes in: 1e39 is a finite Python float and inf on the device
        if not all(math.isfinite(v) and abs(v) <= 3.4028234663852886e38 for v in row):
            # sklearn raises ValueError here (SK/linear_model/_base.py:385-389) -> HTTP 500, as in the reference
            raise ValueError("Input X contains NaN or infinity.")
        label_idx, prob = await state["submit"](row)
        cls = state["classes"][label_idx]
        # same body as the reference's dict return (REF/main.py:24-27), without FastAPI's generic encoder walk
        return JSONResponse({"prediction": cls.item() if hasattr(cls, "item") else cls, "probability": prob})

    @app.post("/files/")
    async def create_file(file: bytes = File(...), token: str = Form(...)):
        # REF/main.py:29-39.  Not on the scoring path.  The reference returns the DataFrame object
        # itself, which current FastAPI cannot serialise (HTTP 500); the frame is returned as columns.
        import pandas as pd
        df = pd.read_csv(StringIO(str(file, "utf-8")))
        return {"file": df.to_dict(orient="list"), "token": token}

    @app.post("/files/predict")
    async def predict_file(file: bytes = File(...), token: str = Form(...)):
        """Bulk scoring of an uploaded CSV (the natural use of the reference's /files/ upload,
        REF/main.py:29-39): columns sepal_length, sepal_width, petal_length, petal_width by name, or
        the first four columns when there is no such header.  One device batch, not the micro-batcher."""
        import numpy as np
        import pandas as pd
        df = pd.read_csv(StringIO(str(file, "utf-8")))
        names = ["sepal_length", "sepal_width", "petal_length", "petal_width"]
        if not set(names) <= set(df.columns):
            df = pd.read_csv(StringIO(str(file, "utf-8")), header=None)
            if df.shape[1] < 4:
                raise ValueError(f"X has {df.shape[1]} features, but LogisticRegression is expecting 4 features as input.")
            X = df.iloc[:, :4].to_numpy(dtype=np.float32)
        else:
            X = df[names].to_numpy(dtype=np.float32)
        if not np.isfinite(X).all():
            raise ValueError("Input X contains NaN or infinity.")
        n = X.shape[0]
        if "client" in state:                 # front process: the rows travel to the engine as one frame
            import asyncio
            res = await asyncio.gather(*[state["submit"](tuple(map(float, r))) for r in X])
            labels = [state["classes"][li] for li, _ in res]
            return {"rows": n, "predictions": labels, "probabilities": [p for _, p in res], "token": token}
        fn = state["bulk_fn"]

        def score_all():                      # blocking device calls: off the event loop (/predict keeps flowing)
            labels, probs = [], []
            with state["bulk_lock"]:
                for lo in range(0, n, 1 << 20):
                    chunk = np.ascontiguousarray(X[lo:lo + (1 << 20)])
                    li, pi = fn(chunk, chunk.shape[0])
                    labels += [state["classes"][int(i)] for i in li[:chunk.shape[0]]]
                    probs += [float(v) for v in pi[:chunk.shape[0]]]
            return [c.item() if hasattr(c, "item") else c for c in labels], probs

        import asyncio
        labels, probs = await asyncio.get_running_loop().run_in_executor(None, score_all)
        return {"rows": n, "predictions": labels, "probabilities": probs, "token": token}

    @app.get("/batcher_stats")
    async def batcher_stats():
        if "client" in state:
            out = {"front": state["client"].stats()}
            path = os.environ.get("LR_ENGINE_STATS", "")
            if path and os.path.exists(path):
                import json
                with open(path) as fh:
                    out.update(json.load(fh))
            return out
        return state["batcher"].stats()

    return app


def __getattr__(name):
    # `uvicorn mlapi_b200.app:app` builds the default app lazily (needs the pickle and a GPU)
    if name == "app":
        return create_app()
    raise AttributeError(name)
