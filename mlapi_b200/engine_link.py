"""One scoring process for many HTTP processes.

The reference serves `/predict` from one uvicorn worker that scores on its event loop
(REF/main.py:16-27, REF/README.md:16).  One Python process parses a few thousand requests per second, so
an app that wants to fill device batches runs several HTTP processes -- and if each of them owns a model
and a micro-batcher of its own (`uvicorn --workers N`), no batch ever sees more than 1/N of the traffic.
Here the GPUs belong to ONE process, the engine: it holds the model replicas (one per device) behind one
`MicroBatcher`.  HTTP front processes keep one Unix-domain stream connection each to the engine and
forward rows over it; what a front collected during one turn of its event loop travels as one frame and
enters the batcher as one block, so the per-request cost on the engine side is a few bytes of framing.

Wire format (little endian), front -> engine:   u32 k | k x (u32 request id, 4 x f32 features)
                             engine -> front:   u32 k | k x (u32 request id, i32 label index, f32 max probability)
A NaN probability marks a row that scored to a non-finite value (that request fails alone).
On connect the engine first sends  u32 n | n bytes  of JSON: {"classes": [...], "n_features": d}.
"""
from __future__ import annotations

import asyncio
import json
import os
import struct
from typing import Dict, List, Optional

import numpy as np

_HDR = struct.Struct("<I")


def _req_dtype(d: int):
    return np.dtype([("id", "<u4"), ("x", "<f4", (d,))])


_RES = np.dtype([("id", "<u4"), ("label", "<i4"), ("prob", "<f4")])


class EngineServer:
    """Accepts front connections on a Unix socket and feeds their frames to a `MicroBatcher`."""

    def __init__(self, batcher, classes, n_features: int, path: str):
        self.batcher = batcher
        self.path = path
        self.n_features = int(n_features)
        self._req = _req_dtype(self.n_features)
        hello = json.dumps({"classes": [c.item() if hasattr(c, "item") else c for c in classes],
                            "n_features": self.n_features}).encode()
        self._hello = _HDR.pack(len(hello)) + hello
        self._server: Optional[asyncio.AbstractServer] = None
        self.frames = 0

    async def start(self):
        if os.path.exists(self.path):
            os.unlink(self.path)
        self._server = await asyncio.start_unix_server(self._serve, path=self.path)
        return self

    async def close(self):
        if self._server is not None:
            self._server.close()
            await self._server.wait_closed()
        if os.path.exists(self.path):
            os.unlink(self.path)

    async def _serve(self, reader: asyncio.StreamReader, writer: asyncio.StreamWriter):
        writer.write(self._hello)
        pending = set()
        try:
            while True:
                try:
                    (k,) = _HDR.unpack(await reader.readexactly(4))
                    body = await reader.readexactly(k * self._req.itemsize)
                except asyncio.IncompleteReadError:
                    return
                self.frames += 1
                t = asyncio.ensure_future(self._score_frame(np.frombuffer(body, dtype=self._req), writer))
                pending.add(t)
                t.add_done_callback(pending.discard)
        finally:
            for t in pending:
                t.cancel()
            writer.close()

    async def _score_frame(self, req: np.ndarray, writer: asyncio.StreamWriter):
        out = np.empty(req.shape[0], dtype=_RES)
        out["id"] = req["id"]
        mb = self.batcher.max_batch
        for lo in range(0, req.shape[0], mb):                    # a frame larger than a device batch is cut
            try:
                lab, prob = await self.batcher.submit_block(req["x"][lo:lo + mb])
            except Exception:  # noqa: BLE001 - an engine failure fails the rows of this frame, not the connection
                lab, prob = 0, np.nan
            out["label"][lo:lo + mb] = lab
            out["prob"][lo:lo + mb] = prob
        writer.write(_HDR.pack(out.shape[0]) + out.tobytes())


class EngineClient:
    """The front side: `await submit(row)` like `MicroBatcher.submit`, answered by the engine process."""

    def __init__(self, path: str):
        self.path = path
        self.classes: List = []
        self.n_features = 0
        self._reader: Optional[asyncio.StreamReader] = None
        self._writer: Optional[asyncio.StreamWriter] = None
        self._waiting: Dict[int, asyncio.Future] = {}
        self._next_id = 0
        self._out_ids: List[int] = []
        self._out_rows: List = []
        self._flush_scheduled = False
        self._reader_task: Optional[asyncio.Task] = None
        self._req = None
        self.frames_sent = 0
        self.rows_sent = 0

    async def connect(self, timeout: float = 60.0):
        deadline = asyncio.get_running_loop().time() + timeout
        while True:
            try:
                self._reader, self._writer = await asyncio.open_unix_connection(self.path)
                break
            except (FileNotFoundError, ConnectionRefusedError):
                if asyncio.get_running_loop().time() > deadline:
                    raise
                await asyncio.sleep(0.1)
        (n,) = _HDR.unpack(await self._reader.readexactly(4))
        hello = json.loads(await self._reader.readexactly(n))
        self.classes = hello["classes"]
        self.n_features = int(hello["n_features"])
        self._req = _req_dtype(self.n_features)
        self._reader_task = asyncio.ensure_future(self._read_loop())
        return self

    async def close(self):
        if self._reader_task is not None:
            self._reader_task.cancel()
        if self._writer is not None:
            self._writer.close()

    async def submit(self, row):
        if len(row) != self.n_features:
            raise ValueError(f"X has {len(row)} features, but LogisticRegression is expecting "
                             f"{self.n_features} features as input.")
        loop = asyncio.get_running_loop()
        fut = loop.create_future()
        rid = self._next_id
        self._next_id = (rid + 1) & 0xFFFFFFFF
        self._waiting[rid] = fut
        self._out_ids.append(rid)
        self._out_rows.append(row)
        if not self._flush_scheduled:          # everything submitted during this turn of the loop leaves as one frame
            self._flush_scheduled = True
            loop.call_soon(self._flush)
        return await fut

    def _flush(self):
        self._flush_scheduled = False
        k = len(self._out_ids)
        if not k:
            return
        frame = np.empty(k, dtype=self._req)
        frame["id"] = self._out_ids
        frame["x"] = self._out_rows
        self._out_ids, self._out_rows = [], []
        self._writer.write(_HDR.pack(k) + frame.tobytes())
        self.frames_sent += 1
        self.rows_sent += k

    async def _read_loop(self):
        try:
            while True:
                (k,) = _HDR.unpack(await self._reader.readexactly(4))
                res = np.frombuffer(await self._reader.readexactly(k * _RES.itemsize), dtype=_RES)
                ids, labels, probs = res["id"].tolist(), res["label"].tolist(), res["prob"].tolist()
                for rid, lab, prob in zip(ids, labels, probs):
                    fut = self._waiting.pop(rid, None)
                    if fut is None or fut.done():
                        continue
                    if prob != prob:       # NaN
                        fut.set_exception(ValueError("Input X contains NaN or infinity (or a value too large for float32)."))
                    else:
                        fut.set_result((lab, prob))
        except (asyncio.IncompleteReadError, asyncio.CancelledError, ConnectionError):
            err = ConnectionError("scoring engine went away")
            for fut in self._waiting.values():
                if not fut.done():
                    fut.set_exception(err)
            self._waiting.clear()

    def stats(self) -> dict:
        return {"frames_sent": self.frames_sent, "rows_sent": self.rows_sent,
                "mean_rows_per_frame": self.rows_sent / self.frames_sent if self.frames_sent else 0.0}


def run_engine(model_path: str, socket_path: str, devices, max_batch: int = 4096, max_wait_us: float = 200.0,
               stats_path: str = ""):
    """Entry point of the engine process: load the pickle once, one model replica per device behind one
    MicroBatcher, serve fronts until killed.  `stats_path`: where the batcher statistics are rewritten every second
    (the fronts' /batcher_stats reads them)."""
    from .batcher import MicroBatcher, engine_score_fns
    from .estimator import PinnedArray, load_model

    clf = load_model(model_path, devices=list(devices))
    pins = []

    def alloc(shape, dtype):
        pins.append(PinnedArray(shape, dtype))
        return pins[-1].array

    mb = MicroBatcher(engine_score_fns(clf), n_features=clf.n_features_in_, max_batch=max_batch, max_wait_us=max_wait_us,
                      alloc=alloc)

    async def main():
        srv = await EngineServer(mb, clf.classes_, clf.n_features_in_, socket_path).start()
        try:
            while True:
                await asyncio.sleep(1.0)
                if stats_path:
                    st = dict(mb.stats(), frames=srv.frames, devices=list(devices))
                    tmp = stats_path + ".tmp"
                    with open(tmp, "w") as fh:
                        json.dump(st, fh)
                    os.replace(tmp, stats_path)
        finally:
            await srv.close()

    try:
        import uvloop
        uvloop.install()
    except ImportError:
        pass
    try:
        asyncio.run(main())
    finally:
        mb.close()
        clf.close()


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser(description="scoring engine process for mlapi_b200.app fronts (LR_ENGINE_SOCKET)")
    ap.add_argument("--model", default=os.environ.get("LR_MODEL_PATH", "LRClassifier.pkl"))
    ap.add_argument("--socket", default=os.environ.get("LR_ENGINE_SOCKET", "/tmp/lr_engine.sock"))
    ap.add_argument("--devices", default=os.environ.get("LR_DEVICES", "0"))
    ap.add_argument("--max-batch", type=int, default=int(os.environ.get("LR_MAX_BATCH", "4096")))
    ap.add_argument("--max-wait-us", type=float, default=float(os.environ.get("LR_MAX_WAIT_US", "200")))
    ap.add_argument("--stats", default="")
    a = ap.parse_args()
    run_engine(a.model, a.socket, [int(x) for x in a.devices.split(",")], a.max_batch, a.max_wait_us, a.stats)
