"""The FastAPI surface and the micro-batcher.

CPU part: host logic only -- the scorer injected into the batcher is the oracle (this is a test; the
product path has no CPU scorer), responses are compared with what the unmodified reference app
answered for the same requests (tests/golden/http_golden.json, made by make_http_golden.py).
GPU part: the same requests through the real engine.
"""
import asyncio
import json
import os
import threading

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import lr_oracle as orc

HTTP = json.load(open(os.path.join(GOLDEN, "http_golden.json")))


def _oracle_score_fn(golden):
    g = golden("iris_model.npz")
    coef, intercept = g["coef"].copy(), g["intercept"].copy()     # NpzFile is not thread-safe
    calls = []

    def fn(stage, n):
        calls.append(n)
        X = stage[:n]
        return orc.predict_index(X, coef, intercept), orc.max_proba(X, coef, intercept)

    return fn, calls, list(g["classes"])


def _check_against_reference(client):
    for item in HTTP["responses"]:
        r = client.post("/predict", json=item["request"])
        assert r.status_code == item["status"], item
        if item["status"] == 200:
            body = r.json()
            assert set(body) == {"prediction", "probability"}
            assert body["prediction"] == item["json"]["prediction"]
            assert abs(body["probability"] - item["json"]["probability"]) <= 1e-5 + 1e-5 * item["json"]["probability"]
        else:
            # pydantic's 422 body: same error locations and types as the reference app
            got = sorted((tuple(e["loc"]), e["type"]) for e in r.json()["detail"])
            ref = sorted((tuple(e["loc"]), e["type"]) for e in item["json"]["detail"])
            assert got == ref


def test_routes_match_reference(golden):
    from mlapi_b200.app import create_app
    fn, _, classes = _oracle_score_fn(golden)
    app = create_app(score_fns=[fn], classes=classes)
    mine = {(r.path, tuple(sorted(r.methods))) for r in app.routes if hasattr(r, "methods")}
    for path, methods in HTTP["routes"]:
        assert (path, tuple(methods)) in mine, path


def test_predict_matches_reference_app_with_stub_scorer(golden):
    from fastapi.testclient import TestClient
    from mlapi_b200.app import create_app
    fn, calls, classes = _oracle_score_fn(golden)
    with TestClient(create_app(score_fns=[fn], classes=classes, max_wait_us=50)) as client:
        _check_against_reference(client)
        # NaN reaches the handler (Python's JSON parser accepts it): like sklearn's validation in the
        # reference, it surfaces as an unhandled ValueError (HTTP 500 under uvicorn), never a 200
        with pytest.raises(ValueError, match="NaN or infinity"):
            client.post("/predict", content='{"sepal_length": NaN, "sepal_width": 1, "petal_length": 1, "petal_width": 1}',
                        headers={"content-type": "application/json"})
        r = client.post("/files/", files={"file": ("a.csv", b"a,b\n1,2\n3,4\n")}, data={"token": "t"})
        assert r.status_code == 200 and r.json() == {"file": {"a": [1, 3], "b": [2, 4]}, "token": "t"}
        csv = "sepal_width,sepal_length,petal_length,petal_width\n3.5,5.1,1.4,0.2\n2.9,6.0,4.5,1.5\n"
        r = client.post("/files/predict", files={"file": ("a.csv", csv.encode())}, data={"token": "t"})
        body = r.json()
        assert r.status_code == 200 and body["rows"] == 2 and body["token"] == "t"
        assert body["predictions"] == ["Iris-setosa", "Iris-versicolor"]
        np.testing.assert_allclose(body["probabilities"], [0.9793734699646953, 0.7493304124804981], atol=1e-5)
        r = client.post("/files/predict", files={"file": ("b.csv", b"5.1,3.5,1.4,0.2\n6.0,2.9,4.5,1.5\n6.3,3.3,6.0,2.5\n")},
                        data={"token": "u"})
        assert r.json()["predictions"] == ["Iris-setosa", "Iris-versicolor", "Iris-virginica"]
    assert sum(calls) >= 8


def test_batcher_coalesces_and_preserves_order(golden):
    from mlapi_b200.batcher import MicroBatcher
    g = golden("iris_model.npz")
    X = g["X"].astype(np.float32)
    fn, calls, _ = _oracle_score_fn(golden)
    mb = MicroBatcher([fn], n_features=4, max_batch=64, max_wait_us=20_000)

    async def burst():
        return await asyncio.gather(*[mb.submit(tuple(map(float, row))) for row in X])

    res = asyncio.run(burst())
    ref_idx = orc.predict_index(X, g["coef"], g["intercept"])
    ref_p = orc.max_proba(X, g["coef"], g["intercept"])
    assert [r[0] for r in res] == list(ref_idx)
    np.testing.assert_allclose([r[1] for r in res], ref_p, rtol=1e-6)
    assert len(calls) < 150 and max(calls) <= 64 and sum(calls) == 150      # coalesced, capped
    with pytest.raises(ValueError, match="expecting 4 features"):
        asyncio.run(mb.submit((1.0, 2.0)))
    assert mb.stats()["requests"] == 150
    mb.close()


def test_batcher_forwards_engine_errors_and_serves_threads(golden):
    from mlapi_b200.batcher import MicroBatcher

    def bad(stage, n):
        raise ValueError("Input X contains NaN or infinity.")

    mb = MicroBatcher([bad], n_features=4, max_batch=8, max_wait_us=100)
    with pytest.raises(ValueError, match="NaN or infinity"):
        mb.submit_blocking((1.0, 2.0, 3.0, 4.0))
    mb.close()
    fn, calls, _ = _oracle_score_fn(golden)
    mb = MicroBatcher([fn, fn], n_features=4, max_batch=16, max_wait_us=500)     # two "replicas"
    out = []
    ts = [threading.Thread(target=lambda: out.append(mb.submit_blocking((5.1, 3.5, 1.4, 0.2)))) for _ in range(40)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert len(out) == 40 and all(o[0] == 0 and abs(o[1] - 0.9793734699646953) < 1e-6 for o in out)
    mb.close()


def test_batcher_never_drops_a_request_under_contention():
    """Regression: with two replicas draining one queue every request must come back (an earlier
    queue.SimpleQueue-with-timeouts version lost about one request in a thousand)."""
    from mlapi_b200.batcher import MicroBatcher

    def fn(stage, n):
        return np.zeros(n, int), stage[:n, 0].astype(np.float64)

    for rep in range(60):
        mb = MicroBatcher([fn, fn], n_features=4, max_batch=16, max_wait_us=300)
        out = []
        ts = [threading.Thread(target=lambda k=k: out.append(mb.submit_blocking((float(k), 0.0, 0.0, 0.0), timeout=5)))
              for k in range(40)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        assert sorted(o[1] for o in out) == [float(k) for k in range(40)], rep
        assert mb.stats()["requests"] == 40
        mb.close()


@pytest.mark.gpu
def test_predict_matches_reference_app_on_gpu(golden, tmp_path):
    """End to end: the same pickle file the reference opens, our app, the real engine."""
    import pickle
    import __graft_entry__
    __graft_entry__.build()
    import synth
    from fastapi.testclient import TestClient
    from mlapi_b200.app import create_app
    clf, _ = synth.fit_notebook_model()
    path = tmp_path / "LRClassifier.pkl"
    with open(path, "wb") as fh:
        pickle.dump(clf, fh)
    with TestClient(create_app(model_path=str(path), max_wait_us=50)) as client:
        _check_against_reference(client)
        stats = client.get("/batcher_stats").json()
        assert stats["requests"] >= 8


@pytest.mark.gpu
def test_batcher_under_concurrent_load_on_gpu(golden):
    import __graft_entry__
    __graft_entry__.build()
    from mlapi_b200 import B200LogisticRegression
    from mlapi_b200.batcher import MicroBatcher, engine_score_fns
    g = golden("iris_batch10k.npz")
    m = golden("iris_model.npz")
    rng = np.random.default_rng(0)
    X = (rng.random((10_000, 4), dtype=np.float32) * 8.0).astype(np.float32)
    clf = B200LogisticRegression(m["coef"], m["intercept"], m["classes"])
    mb = MicroBatcher(engine_score_fns(clf), n_features=4, max_batch=1024, max_wait_us=2000)

    async def burst():
        return await asyncio.gather(*[mb.submit(tuple(map(float, row))) for row in X])

    res = asyncio.run(burst())
    lab = np.array([r[0] for r in res])
    p = np.array([r[1] for r in res])
    top2 = np.sort(g["decision"], axis=1)[:, -2:]
    safe = (top2[:, 1] - top2[:, 0]) >= 1e-5
    assert np.array_equal(lab[safe], g["label_idx"][safe])
    np.testing.assert_allclose(p, g["proba"].max(axis=1), atol=1e-5, rtol=1e-5)
    assert mb.stats()["mean_batch"] > 4
    mb.close()
    clf.close()


def test_one_bad_row_fails_alone_in_its_batch():
    """A request whose row scores to a non-finite value must not take the rest of its micro-batch down
    (the reference scores every request on its own, REF/main.py:20-22)."""
    from mlapi_b200.batcher import MicroBatcher

    def fn(stage, n):                           # stand-in engine: the probability is the first feature
        return np.zeros(n, int), stage[:n, 0].astype(np.float32)

    mb = MicroBatcher([fn], n_features=4, max_batch=64, max_wait_us=20_000)

    async def burst():
        rows = [(0.5, 0.0, 0.0, 0.0)] * 10 + [(float("inf"), 0.0, 0.0, 0.0)] + [(0.25, 0.0, 0.0, 0.0)] * 10
        return await asyncio.gather(*[mb.submit(r) for r in rows], return_exceptions=True)

    res = asyncio.run(burst())
    assert isinstance(res[10], ValueError) and "NaN or infinity" in str(res[10])
    assert all(r == (0, 0.5) for r in res[:10]) and all(r == (0, 0.25) for r in res[11:])
    assert max(mb.batch_sizes) > 1              # they did share a batch
    mb.close()


def test_predict_rejects_values_beyond_float32(golden):
    """1e39 is a finite Python float but infinite in the float32 staging row: that request fails by itself."""
    from fastapi.testclient import TestClient
    from mlapi_b200.app import create_app
    fn, calls, classes = _oracle_score_fn(golden)
    with TestClient(create_app(score_fns=[fn], classes=classes, max_wait_us=50)) as client:
        with pytest.raises(ValueError, match="NaN or infinity"):
            client.post("/predict", json={"sepal_length": 1e39, "sepal_width": 1, "petal_length": 1, "petal_width": 1})
        r = client.post("/predict", json={"sepal_length": 5.1, "sepal_width": 3.5, "petal_length": 1.4, "petal_width": 0.2})
        assert r.status_code == 200 and r.json()["prediction"] == "Iris-setosa"


def test_from_sklearn_rejects_pickles_the_engine_does_not_implement():
    """scikit-learn 0.24 (the reference's pin) pickles may ask for one-vs-rest probabilities or a binary
    softmax: loading them must fail, not answer with different numbers."""
    from mlapi_b200.estimator import B200LogisticRegression

    class Est:
        pass

    e = Est()
    e.coef_, e.intercept_, e.classes_ = np.zeros((3, 4)), np.zeros(3), np.arange(3)
    e.multi_class, e.solver = "ovr", "lbfgs"
    with pytest.raises(NotImplementedError, match="one-vs-rest"):
        B200LogisticRegression.from_sklearn(e)
    e.multi_class, e.solver = "auto", "liblinear"
    with pytest.raises(NotImplementedError, match="one-vs-rest"):
        B200LogisticRegression.from_sklearn(e)
    b = Est()
    b.coef_, b.intercept_, b.classes_ = np.zeros((1, 4)), np.zeros(1), np.arange(2)
    b.multi_class, b.solver = "multinomial", "lbfgs"
    with pytest.raises(NotImplementedError, match="softmax"):
        B200LogisticRegression.from_sklearn(b)


@pytest.mark.gpu
def test_overflowing_row_fails_alone_on_gpu(golden):
    """Finite inputs whose score overflows float32 sit in one device batch with good rows."""
    import __graft_entry__
    __graft_entry__.build()
    from mlapi_b200 import B200LogisticRegression
    from mlapi_b200.batcher import MicroBatcher, engine_score_fns
    m = golden("iris_model.npz")
    clf = B200LogisticRegression(m["coef"] * 1e3, m["intercept"], m["classes"])
    mb = MicroBatcher(engine_score_fns(clf), n_features=4, max_batch=64, max_wait_us=20_000)

    async def burst():
        rows = [(5.1, 3.5, 1.4, 0.2)] * 8 + [(3e38, -3e38, 3e38, -3e38)] + [(6.0, 2.9, 4.5, 1.5)] * 8
        return await asyncio.gather(*[mb.submit(r) for r in rows], return_exceptions=True)

    res = asyncio.run(burst())
    assert isinstance(res[8], ValueError)
    assert all(isinstance(r, tuple) and np.isfinite(r[1]) for r in res[:8] + res[9:])
    mb.close()
    clf.close()


def test_engine_link_fronts_share_one_batcher(golden, tmp_path):
    """Two front connections forward rows over the Unix socket to one engine-side MicroBatcher: answers match the
    oracle, come back to the right request, a non-finite row fails alone, and the engine saw blocks, not rows."""
    from mlapi_b200.batcher import MicroBatcher
    from mlapi_b200.engine_link import EngineClient, EngineServer
    g = golden("iris_model.npz")
    X = g["X"].astype(np.float32)
    fn, calls, classes = _oracle_score_fn(golden)
    sock = str(tmp_path / "e.sock")

    async def main():
        mb = MicroBatcher([fn], n_features=4, max_batch=64, max_wait_us=5_000)
        srv = await EngineServer(mb, classes, 4, sock).start()
        a = await EngineClient(sock).connect()
        b = await EngineClient(sock).connect()
        assert a.classes == list(classes) and a.n_features == 4
        rows = [tuple(map(float, r)) for r in X]
        res = await asyncio.gather(*[(a if i % 2 else b).submit(r) for i, r in enumerate(rows)])
        bad = await asyncio.gather(a.submit((float("nan"), 1.0, 1.0, 1.0)), b.submit(rows[0]), return_exceptions=True)
        with pytest.raises(ValueError, match="expecting 4 features"):
            await a.submit((1.0, 2.0))
        stats = (a.stats(), b.stats(), srv.frames)
        await a.close()
        await b.close()
        await srv.close()
        mb.close()
        return res, bad, stats

    res, bad, (sa, sb, frames) = asyncio.run(main())
    assert [r[0] for r in res] == list(orc.predict_index(X, g["coef"], g["intercept"]))
    np.testing.assert_allclose([r[1] for r in res], orc.max_proba(X, g["coef"], g["intercept"]), rtol=1e-6)
    assert isinstance(bad[0], ValueError) and bad[1] == res[0]
    assert sa["rows_sent"] + sb["rows_sent"] == 152 and frames <= 8          # 150 rows travelled in a handful of frames
    assert max(calls) > 1                                                    # and were scored in shared batches


def test_app_as_front_of_an_engine(golden, tmp_path):
    """The FastAPI app with `engine_socket`: same route, schema and answers, no model in the HTTP process."""
    import threading as th
    from fastapi.testclient import TestClient
    from mlapi_b200.app import create_app
    from mlapi_b200.batcher import MicroBatcher
    from mlapi_b200.engine_link import EngineServer
    fn, calls, classes = _oracle_score_fn(golden)
    sock = str(tmp_path / "e.sock")
    ready, stop = th.Event(), th.Event()

    def engine():
        async def main():
            mb = MicroBatcher([fn], n_features=4, max_batch=64, max_wait_us=50)
            srv = await EngineServer(mb, classes, 4, sock).start()
            ready.set()
            while not stop.is_set():
                await asyncio.sleep(0.02)
            await srv.close()
            mb.close()
        asyncio.run(main())

    t = th.Thread(target=engine, daemon=True)
    t.start()
    assert ready.wait(10)
    try:
        with TestClient(create_app(engine_socket=sock)) as client:
            _check_against_reference(client)
            csv = "sepal_width,sepal_length,petal_length,petal_width\n3.5,5.1,1.4,0.2\n2.9,6.0,4.5,1.5\n"
            r = client.post("/files/predict", files={"file": ("a.csv", csv.encode())}, data={"token": "t"})
            assert r.json()["predictions"] == ["Iris-setosa", "Iris-versicolor"]
            assert client.get("/batcher_stats").json()["front"]["rows_sent"] >= 8
    finally:
        stop.set()
        t.join(5)
