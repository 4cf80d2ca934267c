"""The tcgen05 (3xFP16 / 3xTF32 split) GEMM + fused-softmax path against the oracle, through the C ABI.
Shapes cover one and two 256-class blocks per CTA, clusters of 1/2/4 class slices, class counts
that are not multiples of 4 (scalar stores) or of 32 (ragged last chunk), K tails and ragged n."""
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


@pytest.mark.parametrize("d,c", [(4, 17), (32, 100), (36, 255), (64, 256), (100, 300), (128, 512),
                                 (1024, 1000), (256, 1001), (40, 1024), (64, 1500), (32, 2048)])
def test_gemm_shapes(eng, d, c):
    mlapi, lib = eng
    W, b = synth.make_model(d, c, 60 + c)
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    rng = np.random.default_rng(d + c)
    for n in (1, 127, 129, 1000):
        X = rng.standard_normal((n, d), dtype=np.float32)
        out = clf.score_rows(X, **ALL)
        parity.assert_parity(out, orc.decision_function(X, W, b), orc.predict_proba(X, W, b),
                             orc.predict_index(X, W, b), f"gemm d={d} C={c} n={n}", z_atol=parity.Z_ATOL_TENSOR)
    clf.close()


def test_gemm_confident_rows_and_ties(eng):
    """Large logits (saturating softmax) and exactly tied maxima across class slices."""
    mlapi, lib = eng
    d, c = 32, 1000
    W = np.zeros((c, d))
    W[3, 0] = 1.0
    W[700, 0] = 1.0            # classes 3 and 700 (different slices) tie exactly -> 3 wins
    W[999, 1] = 50.0
    b = np.zeros(c)
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    X = np.zeros((5, d), dtype=np.float32)
    X[0, 0] = 4.0
    X[1, 1] = 10.0             # logit 500 on class 999: one-hot
    X[2, 0] = -2.0
    X[3, :] = 0.0              # all tie -> class 0
    X[4, 0] = 30.0
    out = clf.score_rows(X, **ALL)
    parity.assert_parity(out, orc.decision_function(X, W, b), orc.predict_proba(X, W, b),
                         orc.predict_index(X, W, b), "gemm ties", z_atol=parity.Z_ATOL_TENSOR)
    assert list(out["label"]) == [3, 999, 0, 0, 3]
    assert out["proba"][1, 999] == 1.0
    clf.close()


def test_gemm_matches_generic_path_and_counts_nonfinite(eng):
    mlapi, lib = eng
    X, W, b = synth.make_config("cfg3_d1024_c1000", 777)
    clf = mlapi.B200LogisticRegression(W, b, np.arange(1000))
    auto = clf.score_rows(X, **ALL)          # dispatcher picks the GEMM path for this shape
    clf.set_path(lib.PATH_GENERIC)
    gen = clf.score_rows(X, **ALL)
    np.testing.assert_allclose(auto["proba"], gen["proba"], atol=1e-5)
    np.testing.assert_allclose(auto["decision"], gen["decision"], atol=parity.Z_ATOL_TENSOR)
    clf.set_path(lib.PATH_GEMM)
    Xb = X.copy()
    Xb[5, 100] = np.nan
    Xb[600, 3] = np.inf
    with pytest.raises(ValueError, match="NaN or infinity"):
        clf.predict_proba(Xb)
    clf.predict_proba(X)
    clf.close()


@pytest.mark.parametrize("scale", [1.0, 4.0, 10.0])
def test_gemm_confident_models_meet_the_bar(eng, scale):
    """Large logits amplify the tensor core's truncating accumulation; the fp32 re-scoring pass must
    bring probabilities and labels back inside the parity bar, with and without a P matrix."""
    mlapi, lib = eng
    d, c, n = 1024, 1000, 20_000
    W, b = synth.make_model(d, c, 6)
    W = W * scale
    X = synth.make_X(n, d, 5)
    z = orc.decision_function(X, W, b)
    P = orc.softmax(z)
    idx = z.argmax(1)
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    out = clf.score_rows(X, **ALL)
    rescored = clf.take_rescored()
    # logits of classes that carry no weight stay at tensor-core accuracy (~3.5e-6 |z|)
    parity.assert_parity(out, z, P, idx, f"gemm confident x{scale}", z_atol=4e-6 * np.abs(z).max() + 2e-5)
    out2 = clf.score_rows(X, label=True, maxprob=True)          # the /predict pattern: no P matrix
    parity.assert_parity(out2, z, P, idx, f"gemm confident x{scale} (label+maxprob)")
    if scale >= 4.0:
        assert rescored > 0
    clf.close()


def test_gemm_layout_adapts_to_confident_data(eng):
    """Host batches feed the re-scoring rate back: a confident model moves to the layout with the
    separate correction accumulator, a random one stays on the fast layout; results hold either way."""
    mlapi, lib = eng
    d, c, n = 1024, 1000, 8192
    X = synth.make_X(n, d, 5)
    for scale, expect_many in ((1.0, False), (6.0, True)):
        W, b = synth.make_model(d, c, 6)
        W = W * scale
        clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
        z = orc.decision_function(X, W, b)
        P = orc.softmax(z)
        first = clf.score_rows(X, proba=True, label=True)
        r1 = clf.take_rescored()
        second = clf.score_rows(X, proba=True, label=True)
        r2 = clf.take_rescored()
        for out in (first, second):
            parity.assert_parity(out, z, P, z.argmax(1), f"adaptive x{scale}")
        if expect_many:
            assert r1 > n // 5 and r2 < r1           # second batch ran on the accurate layout
        else:
            assert r1 <= n // 5
        clf.close()


def test_gemm_device_pitch(eng):
    import torch
    mlapi, lib = eng
    d, c, n = 64, 600, 500
    W, b = synth.make_model(d, c, 9)
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    buf = np.random.default_rng(2).standard_normal((n, d + 8), dtype=np.float32)
    Xt = torch.from_numpy(buf).cuda()[:, 4:4 + d]          # pitch d+8, 16-byte aligned offset
    out = clf.score_rows(Xt, **ALL)
    torch.cuda.synchronize()
    host = {k: v.cpu().numpy() for k, v in out.items()}
    Xh = buf[:, 4:4 + d]
    parity.assert_parity(host, orc.decision_function(Xh, W, b), orc.predict_proba(Xh, W, b),
                         orc.predict_index(Xh, W, b), "gemm pitched", z_atol=parity.Z_ATOL_TENSOR)
    assert clf.take_nonfinite() == 0
    clf.close()


@pytest.mark.parametrize("d,c", [(5, 17), (33, 40), (131, 100), (1022, 300), (250, 1000), (1023, 2048)])
def test_gemm_odd_feature_counts(eng, d, c):
    """n_features % 4 != 0 from host arrays: the staging buffers keep a 16-byte pitch, the tensor map's row length
    is the real one (columns past it read as zeros) and the batch stays on the tensor-core path."""
    mlapi, lib = eng
    W, b = synth.make_model(d, c, 500 + c)
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    rng = np.random.default_rng(d * 3 + c)
    for n in (1, 130, 1111):
        X = rng.standard_normal((n, d), dtype=np.float32)
        out = clf.score_rows(X, **ALL)                      # dispatcher: gemm
        parity.assert_parity(out, orc.decision_function(X, W, b), orc.predict_proba(X, W, b),
                             orc.predict_index(X, W, b), f"gemm odd d={d} C={c} n={n}", z_atol=parity.Z_ATOL_TENSOR)
    Xb = rng.standard_normal((300, d), dtype=np.float32)
    Xb[7, d - 1] = np.nan                                   # last (odd) column must be seen
    with pytest.raises(ValueError, match="NaN or infinity"):
        clf.predict_proba(Xb)
    clf.close()


@pytest.mark.parametrize("f16", ["1", "0"])
@pytest.mark.parametrize("d,pitch,off,c,n", [(64, 67, 0, 300, 700),      # odd pitch
                                             (64, 67, 1, 300, 700),      # odd pitch, base 4 bytes past a 16-byte boundary
                                             (64, 68, 2, 40, 5000),      # aligned pitch, base 8 bytes off (8-byte copies)
                                             (1022, 1022, 0, 100, 3000), # dense rows of an even feature count (8-byte copies)
                                             (617, 617, 0, 26, 2500),    # dense rows of an odd feature count (4-byte copies)
                                             (131, 135, 3, 1000, 900),   # several class blocks, K tail, ragged last tile
                                             (70, 70, 2, 17, 129)])
def test_gemm_device_rows_tma_cannot_address(eng, d, pitch, off, c, n, f16, monkeypatch):
    """Row pitch or base not a multiple of 16 bytes: no tensor map can describe X (and the innermost box coordinate
    must be 16-byte aligned too), so the kernel's X producer warps copy the rows with 4- or 8-byte cp.async -- in
    both arithmetics (3xFP16: linear raw tile; 3xTF32: the operand's swizzled layout), with a P matrix and with
    labels + max probability only, whose re-scoring kernels read the same unaligned rows."""
    import ctypes as C
    import torch
    mlapi, lib = eng
    monkeypatch.setenv("LRS_GEMM_F16", f16)
    L = lib.load()
    W, b = synth.make_model(d, c, 11 + d)
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    buf = np.random.default_rng(5 + d).standard_normal(n * pitch + 8, dtype=np.float32)
    Xh = buf[off:off + n * pitch].reshape(n, pitch)[:, :d]
    Xt = torch.from_numpy(buf).cuda()[off:off + n * pitch].view(n, pitch)[:, :d]
    path = C.c_int32(0)
    lib.check(L.lr_model_select_path(clf._replicas[0].handle, C.c_void_p(Xt.data_ptr()), n, pitch, C.byref(path)))
    # (a few rows of a small model go to the one-warp-per-row kernel on their own: forced here)
    assert path.value == (lib.PATH_GEMM if d * c > 4096 or c > 32 or n > 4096 else lib.PATH_GENERIC)
    clf.set_path(lib.PATH_GEMM)
    z = orc.decision_function(Xh, W, b)
    for pattern in (ALL, dict(label=True, maxprob=True)):
        out = clf.score_rows(Xt, **pattern)
        torch.cuda.synchronize()
        host = {k: v.cpu().numpy() for k, v in out.items() if v is not None}
        parity.assert_parity(host, z, orc.softmax(z), z.argmax(1), f"gemm unaligned d={d} pitch={pitch} off={off} f16={f16}",
                             z_atol=parity.Z_ATOL_TENSOR)
    clf.close()


@pytest.mark.parametrize("d,c", [(64, 17), (64, 32), (200, 33), (300, 50), (256, 64), (2048, 40), (512, 100), (96, 101), (512, 150), (96, 128), (128, 160), (64, 224), (64, 225)])
def test_gemm_small_class_counts(eng, d, c):
    """17..224 classes: the unpaired small-N form (N = classes rounded up to 32, two accumulator buffers
    up to N = 128); 225..256 keep the 256-wide tile.  Several row tiles per CTA so both buffers cycle."""
    mlapi, lib = eng
    W, b = synth.make_model(d, c, 700 + c)
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    rng = np.random.default_rng(d + 7 * c)
    for n in (1, 128, 129, 60_000):
        X = rng.standard_normal((n, d), dtype=np.float32)
        out = clf.score_rows(X, **ALL)
        parity.assert_parity(out, orc.decision_function(X, W, b), orc.predict_proba(X, W, b),
                             orc.predict_index(X, W, b), f"gemm small-N d={d} C={c} n={n}", z_atol=parity.Z_ATOL_TENSOR)
    clf.close()


@pytest.mark.parametrize("fixup", ["1", "0"])
def test_gemm_33_to_64_classes_ties_across_column_halves(eng, fixup, monkeypatch):
    """33..64 classes: two warps per row keep 32 columns each in registers and meet in shared memory.  Exact ties
    between a class of the lower half and one of the upper half (5 / 40), inside the upper half (35 / 45) and over
    all classes go to the lowest index -- with the fp32 re-scoring pass and without it (the tied logits come from
    identical W rows, so the tensor arithmetic gives identical values too)."""
    mlapi, lib = eng
    monkeypatch.setenv("LRS_GEMM_FIXUP", fixup)
    d, c = 96, 50
    rng = np.random.default_rng(50)
    W = rng.standard_normal((c, d)) / np.sqrt(d)
    b = rng.standard_normal(c)
    W[40], b[40] = W[5], b[5]
    W[45], b[45] = W[35], b[35]
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    n = 3000
    X = rng.standard_normal((n, d)).astype(np.float32)
    X[:700] += (6.0 * W[5] / np.linalg.norm(W[5])).astype(np.float32)       # classes 5 and 40 lead, tied
    X[700:1400] += (6.0 * W[35] / np.linalg.norm(W[35])).astype(np.float32)  # 35 and 45
    X[-1] = 0.0
    z = orc.decision_function(X, W, b)
    for pattern in (ALL, dict(label=True, maxprob=True)):
        out = clf.score_rows(X, **pattern)
        parity.assert_parity({k: v for k, v in out.items() if v is not None}, z, orc.softmax(z), z.argmax(1),
                             f"two column halves, fixup={fixup}", z_atol=parity.Z_ATOL_TENSOR)
        lab = out["label"]
        assert not np.isin(lab, (40, 45)).any()                  # the higher twin never wins
        assert (lab[:700] == 5).sum() > 500 and (lab[700:1400] == 35).sum() > 500
    # all classes tied on every row: class 0
    clf0 = mlapi.B200LogisticRegression(np.zeros((c, d)), np.full(c, 0.25), np.arange(c))
    clf0.set_path(lib.PATH_GEMM)
    out = clf0.score_rows(X[:300], **ALL)
    assert (out["label"] == 0).all()
    np.testing.assert_allclose(out["proba"], 1.0 / c, rtol=1e-6)
    clf.close()
    clf0.close()


@pytest.mark.parametrize("d,c,n", [(96, 101, 129), (64, 50, 1000), (128, 1001, 300), (64, 150, 257), (40, 17, 130),
                                   (512, 100, 333)])
def test_gemm_outputs_stay_inside_their_rows(eng, d, c, n):
    """The epilogue's shared-memory store tiles write whole row segments: guard rows before and after P, the
    decision values, the labels and the max probabilities keep their canary, for class counts that take the tiles
    (not multiples of 8), ragged last tiles included.  Straight through the C ABI on caller-owned device buffers."""
    import ctypes as C
    import torch
    mlapi, lib = eng
    L = lib.load()
    W, b = synth.make_model(d, c, 3 * d + c)
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    X = torch.randn(n, d, device="cuda")
    canary = -7.25
    P = torch.full(((n + 2) * c,), canary, device="cuda")
    Z = torch.full(((n + 2) * c,), canary, device="cuda")
    lab = torch.full((n + 2,), -7, dtype=torch.int32, device="cuda")
    mp = torch.full((n + 2,), canary, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    lib.check(L.lr_score(clf._replicas[0].handle, C.c_void_p(X.data_ptr()), n, d, C.c_void_p(P.data_ptr() + 4 * c),
                         C.c_void_p(Z.data_ptr() + 4 * c), C.c_void_p(lab.data_ptr() + 4), C.c_void_p(mp.data_ptr() + 4),
                         C.c_void_p(st)))
    torch.cuda.synchronize()
    for t in (P, Z):
        t = t.view(n + 2, c)
        assert (t[0] == canary).all() and (t[-1] == canary).all()
        assert (t[1:-1] != canary).all()
    assert lab[0] == -7 and lab[-1] == -7 and mp[0] == canary and mp[-1] == canary
    Xh = X.cpu().numpy()
    z = orc.decision_function(Xh, W, b)
    parity.assert_parity({"proba": P.view(n + 2, c)[1:-1].cpu().numpy(), "decision": Z.view(n + 2, c)[1:-1].cpu().numpy(),
                          "label": lab[1:-1].cpu().numpy(), "maxprob": mp[1:-1].cpu().numpy()},
                         z, orc.softmax(z), z.argmax(1), f"guard rows d={d} C={c} n={n}", z_atol=parity.Z_ATOL_TENSOR)
    clf.close()


def test_labels_and_max_probability_without_a_p_matrix(eng, monkeypatch):
    """predict + predict_proba().max() in bulk (REF/main.py:21-22) on a many-class model: the batch goes through a
    scratch P in blocks of rows so that the re-scoring pass can pick the classes that matter; labels alone flag
    near-ties only.  Three blocks here (the scratch is shrunk to 1 MB), a ragged last one."""
    import torch
    mlapi, lib = eng
    monkeypatch.setenv("LRS_GEMM_SCRATCH_P_MB", "1")
    d, c, n = 128, 300, 10_000
    W, b = synth.make_model(d, c, 77)
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    Xh = synth.make_X(n, d, 78)
    Xh[:200] *= 3.0                                   # confident rows next to flat ones
    Xt = torch.from_numpy(Xh).cuda()
    z = orc.decision_function(Xh, W, b)
    P = orc.softmax(z)
    before = lib.load().lr_launch_count()
    out = clf.score_rows(Xt, label=True, maxprob=True)
    torch.cuda.synchronize()
    assert lib.load().lr_launch_count() - before >= 3 * 3          # three blocks, at least three kernels each
    parity.assert_parity({k: v.cpu().numpy() for k, v in out.items() if v is not None}, z, P, z.argmax(1), "label+maxprob, blocks")
    clf.take_rescored()
    out = clf.score_rows(Xt, label=True)
    torch.cuda.synchronize()
    parity.assert_parity({"label": out["label"].cpu().numpy()}, z, P, z.argmax(1), "labels alone")
    assert clf.take_rescored() < n // 100                          # near-ties only
    out = clf.score_rows(Xt, label=True, maxprob=True, decision=True)
    torch.cuda.synchronize()
    parity.assert_parity({k: v.cpu().numpy() for k, v in out.items() if v is not None}, z, P, z.argmax(1),
                         "label+maxprob+decision, blocks", z_atol=parity.Z_ATOL_TENSOR)
    clf.close()


def test_concurrent_calls_on_one_model(eng):
    """include/lrscore.h: a model is read-only between create and update, concurrent lr_score calls on distinct
    streams are safe.  Six threads, each with its own stream, batch and outputs, share one tensor-path model (per-call
    scratch is stream-ordered, the layout feedback and the counters are atomics) and one narrow-path model."""
    import threading
    import torch
    mlapi, lib = eng
    models = []
    for d, c in ((256, 300), (40, 3)):
        W, b = synth.make_model(d, c, 90 + c)
        models.append((mlapi.B200LogisticRegression(W, b, np.arange(c)), W, b, d))
    errors = []

    def worker(k):
        try:
            torch.cuda.set_device(0)
            stream = torch.cuda.Stream()
            rng = np.random.default_rng(1000 + k)
            with torch.cuda.stream(stream):
                for it in range(6):
                    clf, W, b, d = models[(k + it) % 2]
                    n = 700 + 97 * k + 13 * it
                    Xh = rng.standard_normal((n, d), dtype=np.float32)
                    Xt = torch.from_numpy(Xh).cuda(non_blocking=False)
                    out = clf.score_rows(Xt, **ALL)
                    stream.synchronize()
                    host = {key: v.cpu().numpy() for key, v in out.items()}
                    z = orc.decision_function(Xh, W, b)
                    parity.assert_parity(host, z, orc.softmax(z), z.argmax(1), f"thread {k} call {it}",
                                         z_atol=parity.Z_ATOL_TENSOR)
        except Exception as e:  # noqa: BLE001 - reported by the main thread
            errors.append(repr(e))

    ts = [threading.Thread(target=worker, args=(k,)) for k in range(6)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errors, errors
    for clf, *_ in models:
        clf.close()


# ---- dynamic range: what the 3xFP16 split must survive ---------------------------------------------
# fp16 halves hold 22 bits only while the `lo` half stays out of subnormals.  The engine balances W per
# feature at load time, puts each class row at the top of fp16's range, scales X per call from a robust
# row maximum and falls back to 3xTF32 when what is left of the batch's range would still cost precision.
# Every case below is held to the ordinary bar (1e-5 / 1e-5, labels outside the 1e-5 margin band), with a
# P matrix and with labels + max probability only (the /predict pattern).

def _check_both_patterns(clf, X, W, b, what, z_atol=parity.Z_ATOL_TENSOR, rows=None):
    z = orc.decision_function(X, W, b)
    P = orc.softmax(z)
    idx = z.argmax(1)
    sel = slice(None) if rows is None else rows
    out = clf.score_rows(X, **ALL)
    parity.assert_parity({k: v[sel] for k, v in out.items()}, z[sel], P[sel], idx[sel], what, z_atol=z_atol)
    out2 = clf.score_rows(X, label=True, maxprob=True)
    parity.assert_parity({k: v[sel] for k, v in out2.items() if v is not None}, z[sel], P[sel], idx[sel],
                         what + " (label+maxprob)")
    return out


@pytest.mark.parametrize("d,c", [(1024, 100), (256, 1000), (64, 32)])
def test_gemm_unstandardised_features(eng, d, c):
    """Feature scales spread over seven decades with the weights a fit would give them (inversely
    proportional): REF/Logistic Regression.ipynb:28-34 fits raw, unscaled columns."""
    mlapi, lib = eng
    rng = np.random.default_rng(1000 + d + c)
    n = 20_000
    W0, b = synth.make_model(d, c, 31)
    s = 10.0 ** rng.uniform(-3, 4, d)
    X = (rng.standard_normal((n, d)) * s).astype(np.float32)
    W = W0 / s
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    _check_both_patterns(clf, X, W, b, f"feature scales d={d} C={c}")
    assert clf.take_fallback() == 0                     # balanced at load time: stays on 3xFP16
    assert clf.take_rescored() < 2 * n * 0.02           # two calls
    clf.close()


def test_gemm_one_huge_feature(eng):
    mlapi, lib = eng
    d, c, n = 1024, 100, 20_000
    W, b = synth.make_model(d, c, 32)
    X = synth.make_X(n, d, 33)
    X[:, 7] *= 1e6
    W[:, 7] /= 1e6
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    _check_both_patterns(clf, X, W, b, "one feature x1e6")
    assert clf.take_fallback() == 0
    clf.close()


@pytest.mark.parametrize("sampled", [True, False])
def test_gemm_single_outlier_element(eng, sampled):
    """One element 1e5 times the rest, in a row the per-call sampler reads (rows i * (n // n_s)) or in one
    it does not: that row overflows fp16 and is re-scored in fp32; no other row may lose precision."""
    mlapi, lib = eng
    d, c, n = 1024, 100, 20_000
    W, b = synth.make_model(d, c, 34)
    X = synth.make_X(n, d, 35)
    step = n // min(n, max(64, min(2048, (1 << 20) // d)))
    row = 5 * step + (0 if sampled else 1)
    X[row, 11] = 1e5
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    # the outlier row itself has logits ~3e3: fp32 re-scoring leaves ~2e-4 on them
    _check_both_patterns(clf, X, W, b, f"outlier sampled={sampled}", z_atol=1e-3)
    assert clf.take_fallback() == 0
    assert clf.take_rescored() >= 2
    clf.close()


def test_gemm_class_row_spanning_six_decades(eng):
    mlapi, lib = eng
    d, c, n = 1024, 100, 20_000
    rng = np.random.default_rng(36)
    W, b = synth.make_model(d, c, 37)
    W[3] = W[3] * 10.0 ** rng.uniform(-6, 0, d) * 20.0
    W[40] = W[40] * 10.0 ** rng.uniform(-6, 0, d) * 20.0
    X = synth.make_X(n, d, 38)
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    _check_both_patterns(clf, X, W, b, "class rows spanning 6 decades")
    assert clf.take_fallback() == 0
    clf.close()


def test_gemm_fitted_on_raw_measurements(eng):
    """The iris situation at 20 classes: positive, unscaled measurements (means of a few units, spreads of a
    fraction of one, one column in the hundreds), a model fitted on them by LogisticRegression().fit."""
    from sklearn.linear_model import LogisticRegression
    mlapi, lib = eng
    rng = np.random.default_rng(39)
    d, c = 32, 20
    mu = np.tile([5.8, 3.0, 3.7, 1.2], d // 4)
    mu[5] = 250.0
    sd = np.tile([0.8, 0.4, 1.7, 0.7], d // 4)
    sd[5] = 40.0
    centres = mu + rng.standard_normal((c, d)) * sd
    ytr = rng.integers(0, c, 6000)
    Xtr = np.abs(centres[ytr] + rng.standard_normal((6000, d)) * sd * 0.7)
    est = LogisticRegression(max_iter=300).fit(Xtr, ytr)
    yte = rng.integers(0, c, 30_000)
    X = np.abs(centres[yte] + rng.standard_normal((30_000, d)) * sd * 0.7).astype(np.float32)
    clf = mlapi.B200LogisticRegression.from_sklearn(est)
    clf.set_path(lib.PATH_GEMM)
    _check_both_patterns(clf, X, est.coef_, est.intercept_, "fitted on raw measurements")
    assert clf.take_fallback() == 0
    clf.close()


@pytest.mark.parametrize("d,c,n", [(4096, 64, 6000), (4096, 1000, 3000), (16384, 64, 3000), (16384, 1000, 1500),
                                   (8192, 20, 3000), (2048, 48, 5000), (2500, 40, 3000), (5000, 17, 2000), (6000, 32, 2000),
                                   (12000, 30, 1500), (1536, 96, 3000), (4096, 128, 2000), (2048, 100, 2500), (1300, 65, 3000)])
def test_gemm_long_rows(eng, d, c, n):
    """The truncating accumulation error grows with the number of k-steps: the re-scoring band has to grow
    with the row length (it was measured at d = 1024); models of at most 64 classes sum long rows in up to
    three (seven: at most 32 classes) TMEM accumulators instead, and in seven (fifteen) without a second
    accumulator buffer when the rows are longer still (d > 2400 / 5600); 65..128 classes in up to three
    from d > 1280."""
    mlapi, lib = eng
    W, b = synth.make_model(d, c, 40 + c)
    X = synth.make_X(n, d, 41)
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    _check_both_patterns(clf, X, W, b, f"long rows d={d} C={c}", z_atol=parity.Z_ATOL_TENSOR * d / 1024)
    clf.close()


def test_gemm_range_guard_falls_back_to_tf32(eng):
    """2 % of the rows 1000 times larger than the rest -- more than the sampler's robust maximum sets aside:
    scaled for them, fp16 would push every ordinary row's low halves into subnormals (predicted logit error
    ~1e-4), so the call must run in 3xTF32 and say so; the ordinary rows stay inside the bar."""
    mlapi, lib = eng
    d, c, n = 1024, 100, 20_000
    W, b = synth.make_model(d, c, 42)
    X = synth.make_X(n, d, 43)
    big = np.zeros(n, bool)
    big[::50] = True
    X[big] *= 1000.0
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    clf.set_path(lib.PATH_GEMM)
    _check_both_patterns(clf, X, W, b, "range guard", rows=~big)
    assert clf.take_fallback() == 2 * n                 # both calls, every row
    # the large rows: logits ~1e3, fp32-grade arithmetic; labels hold wherever the margin is not tiny
    z = orc.decision_function(X[big], W, b)
    lab = clf.score_rows(X, label=True)["label"][big]
    safe = parity.margins(z) > 1.0
    assert (lab[safe] == z.argmax(1)[safe]).all()
    # an ordinary batch on the same model goes back to 3xFP16
    clf.take_fallback()
    clf.score_rows(synth.make_X(4096, d, 44), proba=True)
    assert clf.take_fallback() == 0
    clf.close()


# ---- short rows: 13..32 classes on dense rows of at most 64 floats -----------------------------------
@pytest.mark.parametrize("d,c", [(16, 26), (32, 20), (54, 20), (64, 32), (20, 13), (3, 17), (61, 31), (48, 16)])
def test_short_rows_tensor_form(eng, d, c):
    """The short-row kernel (W resident in shared memory, 1-D bulk copies of whole tiles, eight accumulators in
    flight): aligned and unaligned row lengths, ragged last tiles, unstandardised features, an outlier row."""
    import ctypes as C
    import torch
    mlapi, lib = eng
    L = lib.load()
    rng = np.random.default_rng(900 + d + c)
    W0, b = synth.make_model(d, c, 900 + c)
    s = 10.0 ** rng.uniform(-2, 3, d)
    W = W0 / s
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    for n in (4096, 4097, 50_003):
        X = (rng.standard_normal((n, d)) * s).astype(np.float32)
        X[n // 2, d // 2] = 3e4 * s[d // 2]                 # overflows the fp16 scale: that row is re-scored in fp32
        Xt = torch.from_numpy(X).cuda()
        path = C.c_int32(0)
        lib.check(L.lr_model_select_path(clf._replicas[0].handle, C.c_void_p(Xt.data_ptr()), n, d, C.byref(path)))
        assert path.value == lib.PATH_GEMM
        z = orc.decision_function(X, W, b)
        P = orc.softmax(z)
        idx = z.argmax(1)
        out = clf.score_rows(Xt, **ALL)
        torch.cuda.synchronize()
        host = {k: v.cpu().numpy() for k, v in out.items()}
        # the outlier row's logits are ~1e3: fp32 re-scoring leaves ~1e-4 on them
        parity.assert_parity(host, z, P, idx, f"short rows d={d} C={c} n={n}", z_atol=3e-4)
        out2 = clf.score_rows(Xt, label=True, maxprob=True)
        torch.cuda.synchronize()
        parity.assert_parity({k: v.cpu().numpy() for k, v in out2.items() if v is not None}, z, P, idx,
                             f"short rows d={d} C={c} n={n} (label+maxprob)")
    assert clf.take_fallback() == 0 and clf.take_rescored() >= 6 and clf.take_nonfinite() == 0
    Xb = torch.from_numpy(rng.standard_normal((5000, d)).astype(np.float32)).cuda()
    Xb[17, 0] = float("nan")
    clf.score_rows(Xb, proba=True)
    assert clf.take_nonfinite() == 1
    clf.close()


def test_short_rows_range_guard_falls_back_to_fp32(eng):
    import torch
    mlapi, lib = eng
    d, c, n = 32, 20, 20_000
    W, b = synth.make_model(d, c, 77)
    X = synth.make_X(n, d, 78)
    big = np.zeros(n, bool)
    big[::50] = True
    X[big] *= 3000.0
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    out = clf.score_rows(torch.from_numpy(X).cuda(), **ALL)
    torch.cuda.synchronize()
    host = {k: v.cpu().numpy()[~big] for k, v in out.items()}
    z = orc.decision_function(X[~big], W, b)
    parity.assert_parity(host, z, orc.softmax(z), z.argmax(1), "short rows, fp32 fallback")
    assert clf.take_fallback() == n
    clf.close()


# ---- more than 2048 classes: class groups on the tensor path -------------------------------------------
@pytest.mark.parametrize("d,c", [(64, 2049), (256, 3000), (32, 5000), (130, 4100)])
def test_more_than_2048_classes(eng, d, c):
    """Class groups of equal size, each scored by the tensor path into its columns of the caller's rows, then
    combined per row (log-sum-exp across groups, best class across groups, a tie between two groups' champions)."""
    import ctypes as C
    import torch
    mlapi, lib = eng
    L = lib.load()
    W, b = synth.make_model(d, c, 1200 + c)
    W[c - 1] = W[5]                                     # an exact tie between a class of the first and of the last group,
    b[c - 1] = b[5] = 12.0                              # far above every other class: the lower one must win every row
    clf = mlapi.B200LogisticRegression(W, b, np.arange(c))
    rng = np.random.default_rng(d + c)
    for n in (1, 129, 2500):
        X = rng.standard_normal((n, d), dtype=np.float32)
        Xt = torch.from_numpy(X).cuda()
        path = C.c_int32(0)
        lib.check(L.lr_model_select_path(clf._replicas[0].handle, C.c_void_p(Xt.data_ptr()), n, d, C.byref(path)))
        assert path.value == lib.PATH_GEMM
        z = orc.decision_function(X, W, b)
        P = orc.softmax(z)
        idx = z.argmax(1)
        out = clf.score_rows(X, **ALL)                  # host path (staged chunks)
        parity.assert_parity(out, z, P, idx, f"class groups d={d} C={c} n={n}", z_atol=parity.Z_ATOL_TENSOR)
        assert (out["label"] == 5).all()
        out2 = clf.score_rows(Xt, label=True, maxprob=True)     # device path, the /predict pattern
        assert (out2["label"].cpu().numpy() == 5).all()
        torch.cuda.synchronize()
        parity.assert_parity({k: v.cpu().numpy() for k, v in out2.items() if v is not None}, z, P, idx,
                             f"class groups d={d} C={c} n={n} (label+maxprob)")
    Xb = rng.standard_normal((300, d), dtype=np.float32)
    Xb[7, 1] = np.nan
    with pytest.raises(ValueError, match="NaN or infinity"):
        clf.predict_proba(Xb)
    clf.score_rows(torch.from_numpy(Xb).cuda(), proba=True)
    assert clf.take_nonfinite() == 1                    # one row, counted once whatever the number of groups
    clf.close()
