"""BASELINE.json's full sizes, through size-independent properties plus oracle checks on samples.
Inputs are generated on the device (torch is only the allocator / RNG here)."""
import numpy as np
import pytest

import parity
import synth
from oracle import lr_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    import __graft_entry__
    __graft_entry__.build()
    import mlapi_b200
    from mlapi_b200 import _lib
    return mlapi_b200, _lib


def _sample_check(clf, X, out, W, b, what, k=4096, seed=0, z_atol=parity.Z_ATOL):
    import torch
    n = X.shape[0]
    idx = np.unique(np.concatenate([np.random.default_rng(seed).integers(0, n, k), [0, n - 1]]))
    ti = torch.from_numpy(idx).cuda()
    Xs = X[ti].cpu().numpy()
    got = {key: (None if v is None else v[ti].cpu().numpy()) for key, v in out.items()}
    parity.assert_parity(got, orc.decision_function(Xs, W, b), orc.predict_proba(Xs, W, b),
                         orc.predict_index(Xs, W, b), what, z_atol=z_atol)


def test_cfg1_100m_rows_d32_binary(eng):
    """cfg 1 at full size: 100M x 32 fp32 (12.8 GB) resident, one launch."""
    import torch
    n, d = 100_000_000, 32
    W, b = synth.make_model(d, 1, 2)
    clf = eng[0].B200LogisticRegression(W, b, [0, 1])
    torch.manual_seed(1)
    X = torch.randn(n, d, device="cuda")
    out = clf.score_rows(X, proba=True, decision=True, label=True, maxprob=True)
    torch.cuda.synchronize()
    assert clf.take_nonfinite() == 0
    P, z, lab, mp = out["proba"], out["decision"], out["label"], out["maxprob"]
    # properties that hold for every one of the 100M rows
    assert float((P.sum(dim=1) - 1).abs().max()) <= 2e-7
    assert bool(((z > 0).int() == lab).all())                           # label rule
    assert bool((P.max(dim=1).values == mp).all())
    assert bool(((P[:, 1] > 0.5) <= (lab == 1)).all())
    # linearity checksum: sum of scores == (sum of rows) . w + n*b, in float64
    colsum = X.double().sum(dim=0).cpu().numpy()
    expect = float(colsum @ W[0] + n * b[0])
    got = float(z.double().sum())
    assert abs(got - expect) <= 1e-6 * n ** 0.5 * 10 + 1e-9 * abs(expect) + 5.0, (got, expect)
    # shard invariance: scoring a row block alone gives the same bits
    lo, hi = 37_000_001, 62_500_003
    part = clf.score_rows(X[lo:hi], proba=True, label=True)
    torch.cuda.synchronize()
    assert bool((part["proba"] == P[lo:hi]).all()) and bool((part["label"] == lab[lo:hi]).all())
    _sample_check(clf, X, out, W, b, "cfg1 full size")


def test_cfg2_d100k_binary_resident_chunk(eng):
    """cfg 2 rows are 400 KB each; 2M rows (800 GB) do not fit one GPU, so this scores a resident
    chunk (20 000 rows = 8 GB) the way bench.py does per round, and checks a sample."""
    import torch
    n, d = 20_000, 100_000
    W, b = synth.make_model(d, 1, 4)
    clf = eng[0].B200LogisticRegression(W, b, [0, 1])
    torch.manual_seed(3)
    X = torch.randn(n, d, device="cuda")
    out = clf.score_rows(X, proba=True, decision=True, label=True, maxprob=True)
    torch.cuda.synchronize()
    assert clf.take_nonfinite() == 0
    assert float((out["proba"].sum(dim=1) - 1).abs().max()) <= 2e-7
    assert bool(((out["decision"] > 0).int() == out["label"]).all())
    _sample_check(clf, X, out, W, b, "cfg2 chunk", k=64)
    part = clf.score_rows(X[777:1555], proba=True)
    torch.cuda.synchronize()
    # a different rows-per-CTA mapping may be chosen for the smaller batch: same values to 1e-6
    assert float((part["proba"] - out["proba"][777:1555]).abs().max()) <= 1e-6


def test_cfg3_multinomial_resident_chunk(eng):
    """cfg 3 (d=1024, C=1000): a 200 000-row resident chunk (0.8 GB X + 0.8 GB P)."""
    import torch
    n, d, c = 200_000, 1024, 1000
    W, b = synth.make_model(d, c, 6)
    clf = eng[0].B200LogisticRegression(W, b, np.arange(c))
    torch.manual_seed(5)
    X = torch.randn(n, d, device="cuda")
    out = clf.score_rows(X, proba=True, decision=True, label=True, maxprob=True)
    torch.cuda.synchronize()
    assert clf.take_nonfinite() == 0
    P = out["proba"]
    assert float((P.sum(dim=1) - 1).abs().max()) <= 1e-5
    assert bool((out["decision"].argmax(dim=1).int() == out["label"]).all())
    assert float((P.max(dim=1).values - out["maxprob"]).abs().max()) <= 1e-6
    _sample_check(clf, X, out, W, b, "cfg3 chunk", k=256, z_atol=parity.Z_ATOL_TENSOR)


@pytest.mark.parametrize("d,c,n", [(30, 1, 50_000_000), (54, 7, 20_000_000), (100_001, 1, 30_000)])
def test_unaligned_rows_full_size(eng, d, c, n):
    """The kernels for rows that are not 16-byte aligned at production sizes (flat-tile ring for d=30 binary
    and d=54 x 7 classes, 4-byte-load wide kernel for d=100001): row sums, label rule, shard invariance
    across the 4096-row batch threshold, oracle on sampled rows (first and last row included)."""
    import torch
    W, b = synth.make_model(d, c, 5 + d)
    clf = eng[0].B200LogisticRegression(W, b, list(range(max(c, 2))))
    torch.manual_seed(d)
    X = torch.empty((n, d), device="cuda")
    step = max(1, (1 << 27) // d)
    for lo in range(0, n, step):
        X[lo:lo + step].normal_()
    out = clf.score_rows(X, proba=True, decision=True, label=True, maxprob=True)
    torch.cuda.synchronize()
    assert clf.take_nonfinite() == 0
    P, z, lab, mp = out["proba"], out["decision"], out["label"], out["maxprob"]
    assert float((P.sum(dim=1) - 1).abs().max()) <= 4e-7
    assert bool((P.max(dim=1).values == mp).all())
    if c == 1:
        assert bool(((z > 0).int() == lab).all())
    else:
        assert bool((z.argmax(dim=1).int() == lab).all() or (z.gather(1, lab.long()[:, None]).squeeze(1) == z.max(dim=1).values).all())
    # a small block (below the ring's 4096-row threshold) and a large one score to the same bits
    for lo, hi in ((n // 3, n // 3 + 1000), (n // 2 + 1, n // 2 + 1 + min(n // 4, 2_000_000))):
        part = clf.score_rows(X[lo:hi], proba=True, label=True)
        torch.cuda.synchronize()
        assert bool((part["proba"] == P[lo:hi]).all()) and bool((part["label"] == lab[lo:hi]).all()), (lo, hi)
    _sample_check(clf, X, out, W, b, f"unaligned full size d={d} C={c}", k=1024)
    clf.close()
