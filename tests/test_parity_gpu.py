"""GPU parity tests: the CUDA path, called through the C ABI (ctypes -> liblrscore.so), against
the numpy oracle on seeded inputs and against the committed sklearn golden vectors."""
import ctypes as C

import numpy as np
import pytest

import parity
import synth
from oracle import lr_oracle as orc

pytestmark = pytest.mark.gpu

ALL = dict(proba=True, decision=True, label=True, maxprob=True)


@pytest.fixture(scope="module")
def eng():
    import __graft_entry__
    __graft_entry__.build()
    import mlapi_b200
    from mlapi_b200 import _lib
    return mlapi_b200, _lib


def _model(eng, W, b, classes=None, **kw):
    mlapi, _ = eng
    W = np.atleast_2d(W)
    n_classes = max(W.shape[0], 2)
    return mlapi.B200LogisticRegression(W, b, np.arange(n_classes) if classes is None else classes, **kw)


def _check(clf, X, W, b, what, **sel):
    out = clf.score_rows(X, **(sel or ALL))
    parity.assert_parity(out, orc.decision_function(X, W, b), orc.predict_proba(X, W, b),
                         orc.predict_index(X, W, b), what)
    return out


# ---- cfg 0: the notebook's model -----------------------------------------------------------------
def test_iris_against_golden(eng, golden):
    g = golden("iris_model.npz")
    clf = _model(eng, g["coef"], g["intercept"], g["classes"])
    out = clf.score_rows(g["X"], **ALL)
    parity.assert_parity(out, g["decision"], g["proba"], g["label_idx"], "iris/golden")
    assert list(clf.predict(g["X"][:3])) == ["Iris-setosa"] * 3
    # the two requests of SURVEY.md 8c
    lab, p = clf.predict_with_max_proba([[5.1, 3.5, 1.4, 0.2], [6.0, 2.9, 4.5, 1.5]])
    assert list(lab) == ["Iris-setosa", "Iris-versicolor"]
    np.testing.assert_allclose(p, [0.9793734699646953, 0.7493304124804981], atol=1e-5)
    # notebook accuracy (REF/Logistic Regression.ipynb:13) through the engine
    _, (Xtr, Xte, ytr, yte) = synth.fit_notebook_model()
    assert clf.score(Xte.astype(np.float64), yte) == 0.9666666666666667


def test_iris_batch10k_against_golden(eng, golden):
    g, m = golden("iris_batch10k.npz"), golden("iris_model.npz")
    rng = np.random.default_rng(0)
    X = (rng.random((10_000, 4), dtype=np.float32) * 8.0).astype(np.float32)
    assert synth.sha(X) == str(g["x_sha"])
    clf = _model(eng, m["coef"], m["intercept"], m["classes"])
    out = clf.score_rows(X, **ALL)
    parity.assert_parity(out, g["decision"], g["proba"], g["label_idx"], "iris10k/golden")


# ---- cfg 1..3 at oracle size, against golden and oracle ------------------------------------------
@pytest.mark.parametrize("name", list(synth.CONFIGS))
def test_synthetic_configs_against_golden(eng, golden, name):
    g = golden(name + ".npz")
    X, W, b = synth.make_config(name, int(g["n"]))
    assert synth.sha(X) == str(g["x_sha"])
    clf = _model(eng, W, b)
    out = clf.score_rows(X, **ALL)
    parity.assert_parity(out, g["decision"], g["proba"], g["label_idx"], name + "/golden",
                         z_atol=parity.Z_ATOL_TENSOR if W.shape[0] > 16 else parity.Z_ATOL)


def test_adversarial_rows(eng, golden):
    g = golden("adversarial.npz")
    for pre in ("bin_", "mul_"):
        clf = _model(eng, g[pre + "W"], g[pre + "b"])
        out = clf.score_rows(g[pre + "X"], **ALL)
        parity.assert_parity(out, g[pre + "decision"], g[pre + "proba"], g[pre + "label_idx"], pre)
        if pre == "bin_":
            # exact ties -> classes_[0]; saturation -> exactly 0 / 1
            assert list(out["label"][:2]) == [0, 0]
            assert out["proba"][2, 1] == 1.0 and out["proba"][2, 0] == 0.0
            assert out["proba"][3, 0] == 1.0 and out["proba"][3, 1] == 0.0
        else:
            assert list(out["label"][:3]) == [0, 0, 2]            # first maximum wins


# ---- shape sweep: every narrow lane mapping, ragged n, every path -----------------------------------
@pytest.mark.parametrize("d", [1, 2, 3, 4, 5, 7, 8, 12, 16, 20, 31, 32, 33, 64, 100, 128, 200, 256, 700])
@pytest.mark.parametrize("ns", [1, 3])
def test_narrow_shapes(eng, d, ns):
    _, lib = eng
    rng = np.random.default_rng(1000 + d * 7 + ns)
    W, b = synth.make_model(d, ns, 2000 + d)
    clf = _model(eng, W, b)
    clf.set_path(lib.PATH_NARROW)
    for n in (1, 31, 257, 4099):
        X = rng.standard_normal((n, d), dtype=np.float32)
        _check(clf, X, W, b, f"narrow d={d} ns={ns} n={n}")


@pytest.mark.parametrize("d", [4, 8, 12, 32, 36, 64, 128])
@pytest.mark.parametrize("ns", [1, 3, 4])
def test_narrow_bulk_ring_matches_plain_loads(eng, d, ns, monkeypatch):
    """The cp.async.bulk ring variant (large dense batches) against the oracle, and bit-for-bit
    against the plain-load kernel: both sum each row in the same order."""
    _, lib = eng
    rng = np.random.default_rng(d * 13 + ns)
    W, b = synth.make_model(d, ns, 400 + d)
    monkeypatch.setenv("LRS_NARROW_BULK", "0")
    plain = _model(eng, W, b)
    monkeypatch.setenv("LRS_NARROW_BULK", "1")
    monkeypatch.setenv("LRS_NARROW_BULK_MIN_ROWS", "1")
    bulk = _model(eng, W, b)
    for n in (1, 255, 256, 257, 70_001):
        X = rng.standard_normal((n, d), dtype=np.float32)
        got = _check(bulk, X, W, b, f"bulk d={d} ns={ns} n={n}")
        ref = plain.score_rows(X, **ALL)
        for k in ref:
            assert np.array_equal(ref[k], got[k]), (k, d, ns, n)


@pytest.mark.parametrize("ns", [2, 4, 5, 8, 9, 16])
def test_narrow_class_buckets(eng, ns):
    _, lib = eng
    rng = np.random.default_rng(ns)
    for d in (4, 24, 96):
        W, b = synth.make_model(d, ns, 77 + ns)
        clf = _model(eng, W, b)
        clf.set_path(lib.PATH_NARROW)
        X = rng.standard_normal((1037, d), dtype=np.float32)
        _check(clf, X, W, b, f"narrow ns={ns} d={d}")


@pytest.mark.parametrize("ns", list(range(5, 17)))
def test_mid_class_counts(eng, ns):
    """5..16 classes: the 4-rows-by-NS register-tile kernel.  d picks the lanes-per-row variant
    (up to 256: 4 lanes, above: 8), 4-aligned and not; d=2600 pushes W past the 96 KB that
    fit shared memory for the larger ns (W then comes through L1)."""
    _, lib = eng
    rng = np.random.default_rng(ns * 101)
    for d in (8, 61, 64, 200, 256, 784, 1023, 2600):
        W, b = synth.make_model(d, ns, 900 + ns + d)
        clf = _model(eng, W, b)
        clf.set_path(lib.PATH_NARROW)
        for n in (1, 3, 130, 2051):
            X = rng.standard_normal((n, d), dtype=np.float32)
            _check(clf, X, W, b, f"mid ns={ns} d={d} n={n}")
        clf.close()


@pytest.mark.parametrize("ns", [17, 20, 24])
def test_17_to_24_classes_short_rows(eng, ns):
    """17..24 classes with rows of at most 56 floats (40 when d % 4 == 0) run on the
    register-tile kernel as 24 columns -- small batches, and rows that are not dense (a host batch with
    d % 4 != 0 is staged at a padded pitch); dense batches of 4096 rows and more take the tensor path's
    short-row kernel.  fp32 parity bar everywhere; on the register-tile kernel the same bits for a row in a
    small batch (direct loads) and a large one (ring)."""
    import ctypes as C
    _, lib = eng
    L = lib.load()
    rng = np.random.default_rng(ns * 7)
    dmax = 55                                             # 56 for unaligned rows, 40 for aligned ones
    for d in (4, 16, 33, 40, 54, 55):
        W, b = synth.make_model(d, ns, 40 + ns + d)
        clf = _model(eng, W, b)
        for n in (1, 130, 4099):
            X = rng.standard_normal((n, d), dtype=np.float32)
            out = _check(clf, X, W, b, f"mid32 ns={ns} d={d} n={n}")
        if d % 4 != 0:                                    # both batch sizes ran on the register-tile kernel
            for lo in range(0, n, 4000):
                ref = clf.score_rows(X[lo:lo + 4000], **ALL)
                for k in ref:
                    assert np.array_equal(ref[k], out[k][lo:lo + 4000]), (k, ns, d, lo)
        clf.set_path(lib.PATH_NARROW)                     # fails loudly if the path did not take the shape
        _check(clf, X[:64], W, b, f"mid32 forced narrow ns={ns} d={d}")
        clf.close()
    W, b = synth.make_model(dmax + 1, ns, 5)              # one feature more: tensor-core path
    clf = _model(eng, W, b)
    X = rng.standard_normal((300, dmax + 1), dtype=np.float32)
    out = clf.score_rows(X, **ALL)
    parity.assert_parity(out, orc.decision_function(X, W, b), orc.predict_proba(X, W, b),
                         orc.predict_index(X, W, b), f"ns={ns} d={dmax + 1} (gemm)", z_atol=parity.Z_ATOL_TENSOR)
    clf.close()


@pytest.mark.parametrize("ns", [1, 2, 3, 4, 5, 7, 10, 12, 16])
def test_unaligned_rows_go_through_flat_tiles(eng, ns):
    """Rows that are not 16-byte aligned (odd d, odd pitch, odd base) and >= 4096 of them: tiles of rows
    are bulk-copied to shared memory and read back with scalar loads (<= 4 classes: one float per lane and
    step; 5..16 classes: four, against 128-bit pieces of W).  Against the oracle, and bit-for-bit
    against the plain scalar-load kernel, which serves the same rows when they come in batches below 4096:
    a row's result must not depend on the size of the batch it arrives in."""
    import torch
    mlapi, lib = eng
    rng = np.random.default_rng(ns * 17)
    for d in (1, 3, 5, 7, 13, 30, 33, 61, 127, 255, 509) + ((1001, 2047) if ns == 1 else (1001,)):
        W, b = synth.make_model(d, ns, 300 + ns + d)
        clf = _model(eng, W, b)
        clf.set_path(lib.PATH_NARROW)
        for n in (4096, 4099, 20_011):
            X = rng.standard_normal((n, d), dtype=np.float32)
            out = _check(clf, X, W, b, f"flat ns={ns} d={d} n={n}")
            for lo in range(0, n, 4000):                        # same lanes, same order of FMAs: same bits
                ref = clf.score_rows(X[lo:lo + 4000], **ALL)
                for k in ref:
                    assert np.array_equal(ref[k], out[k][lo:lo + 4000]), (k, ns, d, n, lo)
        clf.close()
    # device rows with an odd pitch and a base that is only 4-byte aligned
    d, n = 30, 9001
    W, b = synth.make_model(d, ns, 41)
    clf = _model(eng, W, b)
    buf = rng.standard_normal((n, d + 7), dtype=np.float32)
    for off in (0, 1, 2, 3):
        Xh = buf[:, off:off + d]
        Xt = torch.from_numpy(buf).cuda()[:, off:off + d]
        out = clf.score_rows(Xt, **ALL)
        torch.cuda.synchronize()
        host = {k: v.cpu().numpy() for k, v in out.items()}
        parity.assert_parity(host, orc.decision_function(Xh, W, b), orc.predict_proba(Xh, W, b),
                             orc.predict_index(Xh, W, b), f"flat pitched ns={ns} off={off}")
    # NaN in the last column of the last row is seen (and nothing past the matrix is read)
    Xb = rng.standard_normal((5000, d), dtype=np.float32)
    Xb[-1, -1] = np.nan
    with pytest.raises(ValueError, match="NaN or infinity"):
        clf.predict_proba(Xb)
    clf.close()


@pytest.mark.parametrize("ns", [1, 3, 4])
def test_long_aligned_rows_go_through_flat_tiles(eng, ns):
    """Aligned rows of 1 KB and more (d >= 256, <= 4 classes) in batches of >= 4096 take the same ring with
    128-bit shared loads; same bits as the direct-load kernel that serves smaller batches."""
    _, lib = eng
    rng = np.random.default_rng(ns + 90)
    for d in (256, 260, 512, 1000, 2048):
        if ns == 1 and d >= 1024:
            continue                                            # binary d >= 1024 belongs to the wide path
        W, b = synth.make_model(d, ns, 80 + d)
        clf = _model(eng, W, b)
        clf.set_path(lib.PATH_NARROW)
        for n in (4096, 6007):
            X = rng.standard_normal((n, d), dtype=np.float32)
            out = _check(clf, X, W, b, f"flat aligned ns={ns} d={d} n={n}")
            for lo in range(0, n, 4000):
                ref = clf.score_rows(X[lo:lo + 4000], **ALL)
                for k in ref:
                    assert np.array_equal(ref[k], out[k][lo:lo + 4000]), (k, ns, d, n, lo)
        clf.close()


@pytest.mark.parametrize("d,ns", [(4100, 2), (8000, 3), (8001, 3), (6000, 4), (5001, 9), (8000, 10)])
def test_few_classes_long_rows(eng, d, ns):
    """2..4 classes with rows too long for W in shared memory run as a 6-column model on the register-tile
    kernel; 5..16 classes read W through L2 there.  Dispatcher's choice must be the narrow path."""
    import ctypes as C
    _, lib = eng
    L = lib.load()
    rng = np.random.default_rng(d + ns)
    W, b = synth.make_model(d, ns, 10 + d)
    clf = _model(eng, W, b)
    for n in (3, 700, 4500):
        X = rng.standard_normal((n, d), dtype=np.float32)
        _check(clf, X, W, b, f"long rows ns={ns} d={d} n={n}")
    clf.set_path(lib.PATH_NARROW)                         # fails loudly if the path did not take the shape
    _check(clf, X[:64], W, b, f"long rows forced narrow ns={ns} d={d}")
    clf.close()


@pytest.mark.parametrize("d", [4, 5, 1000, 1024, 4097, 10_000, 100_000])
def test_wide_shapes(eng, d):
    _, lib = eng
    rng = np.random.default_rng(d)
    W, b = synth.make_model(d, 1, 31 + d)
    clf = _model(eng, W, b)
    clf.set_path(lib.PATH_WIDE)
    for n in (1, 7, 301, 2500):
        if n * d > 40_000_000:
            continue
        X = rng.standard_normal((n, d), dtype=np.float32)
        _check(clf, X, W, b, f"wide d={d} n={n}")


@pytest.mark.parametrize("d", [2049, 4097, 5001, 100_001])
def test_wide_unaligned_device_rows(eng, d):
    """Binary models whose device rows are not 16-byte aligned (odd d, dense) and too long for the narrow
    path: the wide path's 4-byte-load kernel (was the generic kernel at 2.6-3.0 TB/s)."""
    import ctypes as C
    import torch
    _, lib = eng
    L = lib.load()
    rng = np.random.default_rng(d)
    W, b = synth.make_model(d, 1, 3 + d)
    clf = _model(eng, W, b)
    for n in (1, 9, 700, 3000):
        if n * d > 60_000_000:
            continue
        Xh = rng.standard_normal((n, d), dtype=np.float32)
        Xt = torch.from_numpy(Xh).cuda()
        if n == 700:
            path = C.c_int32(0)
            lib.check(L.lr_model_select_path(clf._replicas[0].handle, C.c_void_p(Xt.data_ptr()), n, d, C.byref(path)))
            assert path.value == (lib.PATH_NARROW if d <= 2048 else lib.PATH_WIDE)
        out = clf.score_rows(Xt, **ALL)
        torch.cuda.synchronize()
        host = {k: v.cpu().numpy() for k, v in out.items()}
        parity.assert_parity(host, orc.decision_function(Xh, W, b), orc.predict_proba(Xh, W, b),
                             orc.predict_index(Xh, W, b), f"wide unaligned d={d} n={n}")
    clf.close()


@pytest.mark.parametrize("d,ns", [(1, 1), (7, 1), (33, 2), (50, 20), (130, 37), (1024, 100), (3000, 1)])
def test_generic_shapes(eng, d, ns):
    _, lib = eng
    rng = np.random.default_rng(d * 31 + ns)
    W, b = synth.make_model(d, ns, 5 + d)
    clf = _model(eng, W, b)
    clf.set_path(lib.PATH_GENERIC)
    for n in (1, 33, 1001):
        X = rng.standard_normal((n, d), dtype=np.float32)
        _check(clf, X, W, b, f"generic d={d} ns={ns} n={n}")


def test_auto_dispatch_and_output_subsets(eng):
    _, lib = eng
    rng = np.random.default_rng(3)
    for d, ns in ((32, 1), (4, 3), (2048, 1), (64, 40)):
        W, b = synth.make_model(d, ns, d + ns)
        clf = _model(eng, W, b)
        X = rng.standard_normal((777, d), dtype=np.float32)
        for sel in (dict(proba=True), dict(label=True), dict(maxprob=True), dict(decision=True),
                    dict(label=True, maxprob=True)):
            out = _check(clf, X, W, b, f"auto d={d} ns={ns} {sel}", **sel)
            assert all((out[k] is not None) == bool(sel.get(k)) for k in out)


# ---- sklearn surface and error behaviour --------------------------------------
def test_estimator_surface_matches_sklearn(eng):
    X, W, b = synth.make_config("cfg1_d32_binary", 500)
    sk = synth.sklearn_model(W, b, classes=np.array(["neg", "pos"]))
    clf = eng[0].B200LogisticRegression.from_sklearn(sk)
    assert clf.predict_proba(X).dtype == np.float64 and clf.predict_proba(X).shape == (500, 2)
    assert clf.decision_function(X).shape == (500,)
    safe = np.abs(sk.decision_function(X)) > 1e-5
    assert np.array_equal(clf.predict(X)[safe], sk.predict(X)[safe])
    np.testing.assert_allclose(clf.predict_proba(X), sk.predict_proba(X), atol=1e-5, rtol=1e-5)
    np.testing.assert_allclose(clf.predict_log_proba(X), sk.predict_log_proba(X), atol=2e-5, rtol=1e-4)
    assert abs(clf.score(X, sk.predict(X)) - 1.0) < 0.01
    # list-of-lists, ints, float64, F-order are all accepted like sklearn does
    Xi = np.asfortranarray((X[:9] * 3).astype(np.int64))
    np.testing.assert_allclose(clf.predict_proba(Xi.tolist()), sk.predict_proba(Xi), atol=1e-5)
    np.testing.assert_allclose(clf.predict_proba(Xi), sk.predict_proba(Xi), atol=1e-5)


def test_estimator_errors_match_sklearn(eng):
    X, W, b = synth.make_config("cfg1_d32_binary", 8)
    clf = _model(eng, W, b)
    with pytest.raises(ValueError, match="X has 5 features, but LogisticRegression is expecting 32"):
        clf.predict(np.zeros((3, 5)))
    with pytest.raises(ValueError, match="Expected 2D array"):
        clf.predict(np.zeros(32))
    with pytest.raises(ValueError, match="0 sample"):
        clf.predict(np.zeros((0, 32)))
    for bad in (np.nan, np.inf, -np.inf):
        Xb = X.copy()
        Xb[5, 7] = bad
        with pytest.raises(ValueError, match="NaN or infinity"):
            clf.predict_proba(Xb)
    clf.predict_proba(X)   # the status word was reset: a clean batch passes again
    from sklearn.exceptions import NotFittedError
    from sklearn.linear_model import LogisticRegression
    with pytest.raises(NotFittedError):
        eng[0].B200LogisticRegression.from_sklearn(LogisticRegression())


# ---- device-resident path (torch tensors), row pitch, misalignment ---------------------------------
def test_device_path_pitch_and_alignment(eng):
    import torch
    _, lib = eng
    rng = np.random.default_rng(9)
    for d, ns in ((32, 1), (4, 3), (2000, 1), (12, 5)):
        W, b = synth.make_model(d, ns, 90 + d)
        clf = _model(eng, W, b)
        n = 1237
        buf = rng.standard_normal((n, d + 9), dtype=np.float32)
        for off, pitch_pad in ((0, 0), (0, 4), (1, 3), (3, 9 - 3)):
            Xh = buf[:, off:off + d]
            t_full = torch.from_numpy(buf).cuda()
            Xt = t_full[:, off:off + d]                          # row pitch d+9, base offset `off` floats
            out = clf.score_rows(Xt, **ALL)
            torch.cuda.synchronize()
            host = {k: v.cpu().numpy() for k, v in out.items()}
            parity.assert_parity(host, orc.decision_function(Xh, W, b), orc.predict_proba(Xh, W, b),
                                 orc.predict_index(Xh, W, b), f"device d={d} ns={ns} off={off}")
            assert clf.take_nonfinite() == 0
    # non-finite rows are counted on the device path
    X = torch.randn(100, 32, device="cuda")
    X[3, 3] = float("nan")
    X[50, 0] = float("inf")
    W, b = synth.make_model(32, 1, 1)
    clf = _model(eng, W, b)
    clf.score_rows(X, proba=True)
    assert clf.take_nonfinite() == 2 and clf.take_nonfinite() == 0


def test_host_chunking_is_bit_identical(eng):
    """lr_score_host streams chunks; chunk boundaries must not change a single bit."""
    _, lib = eng
    X, W, b = synth.make_config("cfg1_d32_binary", 50_000)
    clf = _model(eng, W, b)
    ref = clf.score_rows(X, **ALL)
    rep = clf._replicas[0]
    lib.check(rep.lib.lr_model_set_host_chunk_rows(rep.handle, 999))
    got = clf.score_rows(X, **ALL)
    for k in ref:
        assert np.array_equal(ref[k], got[k]), k
    # pinned input/outputs take the same route
    from mlapi_b200.estimator import PinnedArray
    px = PinnedArray(X.shape)
    px.array[...] = X
    got = clf.score_rows(px.array, **ALL)
    for k in ref:
        assert np.array_equal(ref[k], got[k]), k
    px.free()


@pytest.mark.parametrize("d,ns,n,chunk", [(32, 1, 300_000, 0), (40, 3, 120_001, 7_777), (256, 300, 12_000, 5_000)])
def test_pageable_host_arrays_go_through_pinned_bounce_buffers(eng, d, ns, n, chunk):
    """Ordinary numpy arrays of 8 MB and more: lr_score_host copies them chunk by chunk into pinned buffers of its
    own with a few host threads (and the results back out) instead of letting cudaMemcpyAsync crawl through
    pageable memory.  Same bits as the pinned route -- on every kernel family, with ragged last chunks, a row pitch
    that is not the row length, and with only some of the buffers pageable."""
    from mlapi_b200.estimator import PinnedArray
    _, lib = eng
    W, b = synth.make_model(d, ns, 31 + d)
    clf = _model(eng, W, b)
    rep = clf._replicas[0]
    if chunk:
        lib.check(rep.lib.lr_model_set_host_chunk_rows(rep.handle, chunk))
    X = synth.make_X(n, d, 77)
    assert X.nbytes >= 8 << 20
    px = PinnedArray(X.shape)
    px.array[...] = X
    ref = clf.score_rows(px.array, **ALL)                    # pinned in (outputs are numpy arrays: pageable)
    got = clf.score_rows(X, **ALL)                           # pageable in
    for k in ref:
        assert np.array_equal(ref[k], got[k]), k
    got64 = clf.score_rows(X.astype(np.float64), **ALL)      # float64 in: rounded to fp32 inside lr_score_host_f64
    for k in ref:
        assert np.array_equal(ref[k], got64[k]), k
    wide64 = clf.score_rows(X, out_f64=True, **ALL)          # float64 results: the same fp32 values, widened
    for k in ref:
        assert wide64[k].dtype == (np.int32 if k == "label" else np.float64) and np.array_equal(ref[k].astype(wide64[k].dtype), wide64[k]), k
    few64 = clf.score_rows(X[:3], out_f64=True, **ALL)
    assert few64["proba"].dtype == np.float64 and (ns > 16 or np.array_equal(ref["proba"][:3].astype(np.float64), few64["proba"]))
    few = clf.score_rows(X[:7].astype(np.float64), **ALL)    # ... and a handful of float64 rows
    for k in ref:
        assert np.array_equal(ref[k][:7], few[k]) or ns > 16, k      # (the tensor path scales each call from its own rows)
    z = orc.decision_function(X[:4000], W, b)
    parity.assert_parity({k: v[:4000] for k, v in got.items()}, z if ns > 1 else z.ravel(), orc.predict_proba(X[:4000], W, b),
                         orc.predict_index(X[:4000], W, b), f"pageable d={d} ns={ns}",
                         z_atol=parity.Z_ATOL_TENSOR if ns > 16 else parity.Z_ATOL)
    # straight through the C ABI: pageable X with a pitch wider than the row, everything pinned on the way out
    L = lib.load()
    wide = np.zeros((n, d + 5), np.float32)
    wide[:, :d] = X
    n_out = max(ns, 2)
    pp = PinnedArray((n, n_out))
    pl = PinnedArray((n,), np.int32)
    lib.check(L.lr_score_host(rep.handle, wide.ctypes.data_as(C.c_void_p), n, d + 5, pp.array.ctypes.data_as(C.c_void_p), None,
                              pl.array.ctypes.data_as(C.c_void_p), None))
    assert np.array_equal(pp.array, ref["proba"]) and np.array_equal(pl.array, ref["label"])
    for a in (px, pp, pl):
        a.free()
    clf.close()


def test_request_sized_pinned_batches_are_scored_in_place(eng, golden):
    """lr_score_host with page-locked X and outputs of a few KB takes the zero-copy path; the results
    must be the bits the staged path produces, for every narrow/wide/generic shape class."""
    _, lib = eng
    from mlapi_b200.estimator import PinnedArray
    L = lib.load()
    rng = np.random.default_rng(12)
    for d, ns, path in ((4, 3, None), (32, 1, None), (2000, 1, None), (16, 40, lib.PATH_GENERIC)):   # narrow, narrow, wide, generic
        W, b = synth.make_model(d, ns, 70 + d)
        clf = _model(eng, W, b)
        if path is not None:
            clf.set_path(path)                                   # (the generic kernel serves any shape when asked to)
        rep = clf._replicas[0]
        for n in (1, 7, 8):
            X = rng.standard_normal((n, d), dtype=np.float32)
            ref = clf.score_rows(X, **ALL)                       # pageable arrays: staged path
            px = PinnedArray((n, d)); px.array[...] = X
            pp = PinnedArray((n, max(ns, 2))); pz = PinnedArray((n,) if ns == 1 else (n, ns))
            pl = PinnedArray((n,), np.int32); pm = PinnedArray((n,))
            before = L.lr_launch_count()
            ptr = lambda a: a.array.ctypes.data_as(C.c_void_p)  # noqa: E731
            assert L.lr_score_host(rep.handle, ptr(px), n, d, ptr(pp), ptr(pz), ptr(pl), ptr(pm)) == 0, lib.last_error()
            assert L.lr_launch_count() == before + 1
            for got, key in ((pp, "proba"), (pz, "decision"), (pl, "label"), (pm, "maxprob")):
                assert np.array_equal(got.array, ref[key]), (d, ns, n, key)
            px.array[0, 0] = np.nan
            assert L.lr_score_host(rep.handle, ptr(px), n, d, ptr(pp), None, None, None) == lib.LR_ERR_NONFINITE
            for a in (px, pp, pz, pl, pm):
                a.free()
        clf.close()


def test_c_abi_direct_call(eng):
    """Straight ctypes, no wrapper: what INTEGRATION.md's stub does."""
    _, lib = eng
    L = lib.load()
    g = np.load(__import__("os").path.join(__import__("conftest").GOLDEN, "iris_model.npz"))
    coef = np.ascontiguousarray(g["coef"])
    h = C.c_void_p()
    assert L.lr_model_create(coef.ctypes.data_as(C.c_void_p), g["intercept"].ctypes.data_as(C.c_void_p),
                             3, 4, 0, C.byref(h)) == 0
    X = np.array([[5.1, 3.5, 1.4, 0.2]], dtype=np.float32)
    label = np.zeros(1, np.int32)
    maxp = np.zeros(1, np.float32)
    before = L.lr_launch_count()
    assert L.lr_score_host(h, X.ctypes.data_as(C.c_void_p), 1, 4, None, None,
                           label.ctypes.data_as(C.c_void_p), maxp.ctypes.data_as(C.c_void_p)) == 0
    assert L.lr_launch_count() == before + 1
    assert label[0] == 0 and abs(maxp[0] - 0.9793734699646953) < 1e-5
    assert L.lr_score_host(h, X.ctypes.data_as(C.c_void_p), 1, 3, None, None, None, None) == lib.LR_ERR_INVALID
    assert "ldx" in lib.last_error()
    L.lr_model_destroy(h)
