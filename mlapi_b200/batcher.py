"""Request micro-batcher: coalesces single-row /predict calls into device batches.

The reference scores one row per request on the event loop, re-loading the pickle each time
(REF/main.py:17-22).  Here requests are queued; one worker thread per model replica (= per GPU)
drains the queue into a pinned staging matrix and makes ONE `lr_score_host` call per batch (label +
max probability in a single pass, i.e. REF/main.py:21-22 fused), then resolves the waiting futures.
A batch closes when it reaches `max_batch` rows or when its oldest request has waited
`max_wait_us`.  Replicas pull from the same queue, so multi-GPU serving is "replicas only": no
collective, natural load balance.
"""
from __future__ import annotations

import asyncio
import threading
import time
from collections import Counter, deque
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np

ScoreFn = Callable[[np.ndarray, int], Tuple[np.ndarray, np.ndarray]]
"""(staging rows [max_batch, d] float32, n valid rows) -> (label index [n] int, max proba [n] float)"""


class MicroBatcher:
    def __init__(self, score_fns: Sequence[ScoreFn], n_features: int, max_batch: int = 4096,
                 max_wait_us: float = 200.0, alloc: Optional[Callable] = None):
        if max_batch < 1:
            raise ValueError("max_batch must be >= 1")
        self.n_features = int(n_features)
        self.max_batch = int(max_batch)
        self.max_wait = float(max_wait_us) * 1e-6
        # one deque guarded by one condition variable: producers append and notify, each worker takes a
        # batch under the lock (an item is either still queued or owned by exactly one worker)
        self._dq: deque = deque()
        self._cv = threading.Condition()
        self._stop = threading.Event()
        self._stats_lock = threading.Lock()
        self.batch_sizes: Counter = Counter()
        self.requests = 0
        self._alloc = alloc or (lambda shape, dtype: np.empty(shape, dtype))
        self._threads: List[threading.Thread] = []
        for i, fn in enumerate(score_fns):
            t = threading.Thread(target=self._worker, args=(fn,), name=f"lr-batcher-{i}", daemon=True)
            t.start()
            self._threads.append(t)

    # ---- producer side (event loop) ----------------------------------------------------------------
    async def submit(self, row: Sequence[float]) -> Tuple[int, float]:
        """Queue one row; resolves to (label index, max class probability)."""
        if len(row) != self.n_features:
            raise ValueError(f"X has {len(row)} features, but LogisticRegression is expecting "
                             f"{self.n_features} features as input.")
        loop = asyncio.get_running_loop()
        fut = loop.create_future()
        self._put((row, fut, loop))
        return await fut

    def submit_blocking(self, row: Sequence[float], timeout: float = 30.0) -> Tuple[int, float]:
        """Thread-side variant for callers without an event loop."""
        done = threading.Event()
        box = {}

        class _Fut:
            def set_result(self, v):
                box["v"] = v
                done.set()

            def set_exception(self, e):
                box["e"] = e
                done.set()

            def done(self):
                return done.is_set()

        self._put((row, _Fut(), None))
        if not done.wait(timeout):
            raise TimeoutError("micro-batcher did not answer")
        if "e" in box:
            raise box["e"]
        return box["v"]

    async def submit_block(self, rows: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """Queue k rows at once (k <= max_batch; an HTTP front process forwards what it collected in one event-loop
        turn); resolves to (label indices [k], max probabilities [k]).  A row that scored to a non-finite value
        carries a NaN probability: the caller fails that request alone."""
        rows = np.ascontiguousarray(rows, dtype=np.float32)
        if rows.ndim != 2 or rows.shape[1] != self.n_features:
            raise ValueError(f"X has {rows.shape[-1]} features, but LogisticRegression is expecting "
                             f"{self.n_features} features as input.")
        if rows.shape[0] > self.max_batch:
            raise ValueError(f"block of {rows.shape[0]} rows exceeds max_batch {self.max_batch}")
        loop = asyncio.get_running_loop()
        fut = loop.create_future()
        self._put((rows, fut, loop))
        return await fut

    def _put(self, item):
        with self._cv:
            self._dq.append(item)
            self._cv.notify()

    # ---- consumer side (one thread per replica) ----------------------------------------------------
    def _take_batch(self):
        """Block for the first request, then keep collecting until the batch is full or its oldest
        request has waited max_wait.  Returns [] when the batcher is closing."""
        with self._cv:
            while not self._dq:
                if self._stop.is_set():
                    return []
                self._cv.wait(0.05)
            items = [self._dq.popleft()]
            rows = self._rows_of(items[0])
            deadline = time.perf_counter() + self.max_wait
            while rows < self.max_batch:
                if self._dq:
                    k = self._rows_of(self._dq[0])
                    if rows + k > self.max_batch:
                        break                # a block is never split: it opens the next batch
                    items.append(self._dq.popleft())
                    rows += k
                    continue
                remaining = deadline - time.perf_counter()
                if remaining <= 0 or self._stop.is_set():
                    break
                self._cv.wait(remaining)
            if self._dq:
                self._cv.notify()            # more work is queued: wake another replica
            return items

    @staticmethod
    def _rows_of(item) -> int:
        return item[0].shape[0] if isinstance(item[0], np.ndarray) else 1

    def _worker(self, score_fn: ScoreFn):
        stage = self._alloc((self.max_batch, self.n_features), np.float32)
        while True:
            items = self._take_batch()
            if not items:
                return
            n = 0
            try:
                spans = []
                for row, _, _ in items:
                    k = self._rows_of((row,))
                    stage[n:n + k] = row
                    spans.append((n, k, isinstance(row, np.ndarray)))
                    n += k
                labels, probs = score_fn(stage, n)
                # A row that scored to NaN / infinity (sklearn: ValueError, SK/linear_model/_base.py:385-389) fails
                # alone: the other requests that happened to share its batch get their answers.
                results = []
                for lo, k, block in spans:
                    if block:
                        results.append((np.array(labels[lo:lo + k], dtype=np.int32), np.array(probs[lo:lo + k], dtype=np.float32)))
                    elif np.isfinite(probs[lo]):
                        results.append((int(labels[lo]), float(probs[lo])))
                    else:
                        results.append(ValueError("Input X contains NaN or infinity (or a value too large for float32)."))
                err = None
            except Exception as e:  # noqa: BLE001 - forwarded to every waiter
                results, err = None, e
                n = sum(self._rows_of(it) for it in items)
            with self._stats_lock:
                self.batch_sizes[n] += 1
                self.requests += n
            self._resolve(items, results, err)

    @staticmethod
    def _resolve(items, results, err):
        by_loop = {}
        for i, (_, fut, loop) in enumerate(items):
            by_loop.setdefault(loop, []).append((fut, None if results is None else results[i]))

        def finish(pairs):
            for fut, res in pairs:
                if fut.done():
                    continue
                if err is not None:
                    fut.set_exception(err)
                elif isinstance(res, Exception):
                    fut.set_exception(res)
                else:
                    fut.set_result(res)

        for loop, pairs in by_loop.items():
            if loop is None:
                finish(pairs)
            else:
                loop.call_soon_threadsafe(finish, pairs)     # one wake-up per batch, not per request

    def close(self):
        self._stop.set()
        with self._cv:
            self._cv.notify_all()
        for t in self._threads:
            t.join(timeout=2.0)

    def stats(self) -> dict:
        sizes = sorted(self.batch_sizes.items())
        total = sum(c for _, c in sizes)
        # histogram in powers of two: {upper bound: batches}
        hist = {}
        for size, count in sizes:
            ub = 1 << max(0, int(size - 1).bit_length())
            hist[ub] = hist.get(ub, 0) + count
        return {"requests": self.requests, "batches": total,
                "mean_batch": (self.requests / total) if total else 0.0,
                "max_batch_seen": sizes[-1][0] if sizes else 0,
                "batch_size_histogram": {str(k): v for k, v in sorted(hist.items())}}


def engine_score_fns(clf) -> List[ScoreFn]:
    """One scoring closure per replica of a `B200LogisticRegression`: label + max proba through
    `lr_score_host`, outputs in preallocated arrays."""
    import ctypes as C

    from . import _lib

    fns = []
    for rep in clf._replicas:
        def make(rep=rep):
            labels = None
            probs = None
            pins = []

            def fn(stage: np.ndarray, n: int):
                nonlocal labels, probs
                if labels is None or labels.shape[0] < stage.shape[0]:
                    # pinned outputs: request-sized batches are then scored in place (zero-copy path
                    # of lr_score_host), larger ones get asynchronous D2H copies
                    from .estimator import PinnedArray
                    pins[:] = [PinnedArray((stage.shape[0],), np.int32), PinnedArray((stage.shape[0],), np.float32)]
                    labels, probs = pins[0].array, pins[1].array
                code = rep.lib.lr_score_host(rep.handle, stage.ctypes.data_as(C.c_void_p), n, stage.shape[1],
                                             None, None, labels.ctypes.data_as(C.c_void_p),
                                             probs.ctypes.data_as(C.c_void_p))
                # LR_ERR_NONFINITE: the outputs are written all the same, the offending rows hold a non-finite
                # maximum probability -- the batcher fails those requests only
                if code != _lib.LR_ERR_NONFINITE:
                    _lib.check(code)
                return labels, probs

            return fn
        fns.append(make())
    return fns
