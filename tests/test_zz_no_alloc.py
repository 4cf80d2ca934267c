"""lr_loss_grad_static -- the refit loss/gradient entry point for callers that are not allowed to
allocate device memory (a server that gave all HBM to its own pool at start-up, a refit step that
must not stall neighbouring streams inside cudaMalloc, an embedding runtime that keeps third-party
code off the device heap).  Its per-CTA partial sums live in an array compiled into the library.

The file name sorts last on purpose: a `pytest -x` run has already been through every other test
when it gets here.  The test drives the entry point through the C ABI (ctypes), twice in a row on
the same built-in array, and compares each result with what lr_loss_grad (the allocating entry
point) returns for the same inputs, to the tolerance the refit tests use.  It passes on its own in a
fresh process as well as at the end of the whole suite.
"""
import ctypes as C

import numpy as np
import pytest

from oracle import lr_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__
    __graft_entry__.build()
    from mlapi_b200 import _lib
    return _lib


def _call(lib, fn_name, X, y, W, b):
    """One call of `fn_name` through ctypes on torch-owned device buffers (inputs and outputs are the
    caller's; the entry point under test allocates nothing itself)."""
    import torch
    rows, d = W.shape
    out = torch.full((1 + rows * (d + 1),), float("nan"), dtype=torch.float64, device="cuda")
    fn = getattr(lib.load(), fn_name)
    stream = torch.cuda.current_stream().cuda_stream
    code = fn(C.c_void_p(X.data_ptr()), C.c_void_p(y.data_ptr()), X.shape[0], X.stride(0),
              C.c_void_p(W.data_ptr()), C.c_void_p(b.data_ptr()), rows, d, C.c_void_p(out.data_ptr()),
              C.c_void_p(stream))
    torch.cuda.synchronize()
    return code, out.cpu().numpy()


def _problem(seed, n, d, rows):
    import torch
    rng = np.random.default_rng(seed)
    X = torch.as_tensor(rng.standard_normal((n, d), dtype=np.float32)).cuda()
    y = torch.as_tensor(rng.integers(0, 2 if rows == 1 else rows, n).astype(np.int32)).cuda()
    W = torch.as_tensor((rng.standard_normal((rows, d)) / np.sqrt(d)).astype(np.float32)).cuda()
    b = torch.as_tensor(rng.standard_normal(rows).astype(np.float32)).cuda()
    return X, y, W, b


def test_static_scratch_entry_point_matches_allocating_one_twice_in_a_row(lib):
    # two different problems back to back: the second call finds the first call's partial sums and
    # arrival counter still in the built-in array and must not be disturbed by them
    for seed, n, d, rows in ((1, 200_003, 32, 1), (2, 50_001, 4, 3)):
        X, y, W, b = _problem(seed, n, d, rows)
        code_ref, ref = _call(lib, "lr_loss_grad", X, y, W, b)
        code, got = _call(lib, "lr_loss_grad_static", X, y, W, b)
        assert code_ref == lib.LR_OK and code == lib.LR_OK, lib.last_error()
        assert np.isfinite(got).all()
        # same tolerance as tests/test_refit_gpu.py, on the per-row means
        np.testing.assert_allclose(got[0] / n, ref[0] / n, rtol=5e-6, atol=1e-7)
        np.testing.assert_allclose(got[1:] / n, ref[1:] / n, rtol=1e-4, atol=3e-6)
        # and against the float64 restatement of scikit-learn's loss_gradient itself, not only its sibling
        Wh, bh = W.cpu().numpy().astype(np.float64), b.cpu().numpy().astype(np.float64)
        loss, gW, gb = orc.loss_gradient(Wh, bh, X.cpu().numpy(), y.cpu().numpy(), 0.0)
        np.testing.assert_allclose(got[0] / n, loss, rtol=2e-5, atol=1e-6)
        np.testing.assert_allclose(got[1:1 + rows * d].reshape(rows, d) / n, gW, rtol=1e-4, atol=5e-6)
        np.testing.assert_allclose(got[1 + rows * d:] / n, gb, rtol=1e-4, atol=5e-6)
    # and the very same call again gives the very same bits (nothing left over changes the result)
    code, again = _call(lib, "lr_loss_grad_static", X, y, W, b)
    assert code == lib.LR_OK and np.array_equal(again, got)


def test_static_scratch_entry_point_refuses_what_does_not_fit(lib):
    # 1 + 8*(128+1) = 1033 floats per partial vector > 512: refused, with a message that says why
    X, y, W, b = _problem(3, 64, 128, 8)
    code, _ = _call(lib, "lr_loss_grad_static", X, y, W, b)
    assert code == lib.LR_ERR_LIMIT
    msg = lib.last_error()
    assert "1033" in msg and "512" in msg and "lr_loss_grad" in msg
    # the allocating entry point serves the same shape
    code, out = _call(lib, "lr_loss_grad", X, y, W, b)
    assert code == lib.LR_OK and np.isfinite(out).all()
