"""The numpy oracle against the committed sklearn outputs (tests/golden/, made by make_golden.py)
and against the one number the reference itself records (notebook accuracy)."""
import numpy as np
import pytest

import synth
from oracle import lr_oracle as orc

TIGHT = dict(rtol=1e-12, atol=1e-14)


def _check(g, X, W, b, prefix=""):
    dec = orc.decision_function(X, W, b)
    np.testing.assert_allclose(dec, g[prefix + "decision"], rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(orc.predict_proba(X, W, b), g[prefix + "proba"], rtol=1e-10, atol=1e-14)
    # labels: bit-exact wherever the fp64 margin is not at rounding level
    idx = orc.predict_index(X, W, b)
    ref = g[prefix + "label_idx"]
    if dec.ndim == 1:
        safe = np.abs(dec) > 1e-9
    else:
        top2 = np.sort(dec, axis=1)[:, -2:]
        safe = (top2[:, 1] - top2[:, 0]) > 1e-9
    assert np.array_equal(idx[safe], ref[safe])
    return dec


def test_iris_model_matches_notebook_and_sklearn(golden):
    g = golden("iris_model.npz")
    assert g["notebook_score"] == 0.9666666666666667          # REF/Logistic Regression.ipynb:13
    assert list(g["classes"]) == synth.IRIS_CLASSES
    _check(g, g["X"], g["coef"], g["intercept"])
    labels = orc.predict(g["X"], g["coef"], g["intercept"], g["classes"])
    assert labels[0] == "Iris-setosa"
    # the two requests quoted in SURVEY.md section 8c
    p = orc.max_proba([[5.1, 3.5, 1.4, 0.2], [6.0, 2.9, 4.5, 1.5]], g["coef"], g["intercept"])
    np.testing.assert_allclose(p, [0.9793734699646953, 0.7493304124804981], rtol=1e-12)


def test_iris_fit_reproduces_model(golden):
    """The oracle's L-BFGS refit lands on the notebook model's coefficients."""
    g = golden("iris_model.npz")
    y = np.array([list(g["classes"]).index(v) for v in g["y_train"]])
    W, b, nit = orc.fit_lbfgs(g["X_train"], y, 3)
    np.testing.assert_allclose(W, g["coef"], atol=2e-3)
    np.testing.assert_allclose(b, g["intercept"], atol=2e-2)
    acc = (orc.predict_index(g["X"], W, b) == orc.predict_index(g["X"], g["coef"], g["intercept"])).mean()
    assert acc == 1.0


def test_iris_batch10k(golden):
    g = golden("iris_batch10k.npz")
    m = golden("iris_model.npz")
    rng = np.random.default_rng(0)
    X = (rng.random((10_000, 4), dtype=np.float32) * 8.0).astype(np.float32)
    assert synth.sha(X) == str(g["x_sha"])
    _check(g, X, m["coef"], m["intercept"])


@pytest.mark.parametrize("name", list(synth.CONFIGS))
def test_synthetic_configs(golden, name):
    g = golden(name + ".npz")
    X, W, b = synth.make_config(name, int(g["n"]))
    assert synth.sha(X) == str(g["x_sha"]) and synth.sha(W, b) == str(g["model_sha"])
    dec = _check(g, X, W, b)
    assert dec.ndim == (1 if W.shape[0] == 1 else 2)
    assert g["proba"].shape == (X.shape[0], max(2, W.shape[0]))


def test_adversarial_rows(golden):
    g = golden("adversarial.npz")
    for pre in ("bin_", "mul_"):
        X, W, b = g[pre + "X"], g[pre + "W"], g[pre + "b"]
        np.testing.assert_allclose(orc.predict_proba(X, W, b), g[pre + "proba"], **TIGHT)
        assert np.array_equal(orc.predict_index(X, W, b), g[pre + "label_idx"])   # ties included
    P = g["bin_proba"]
    assert list(g["bin_label_idx"][:2]) == [0, 0]            # score == 0 -> classes_[0]
    assert P[2, 1] == 1.0 and P[2, 0] == 0.0 and P[3, 0] == 1.0 and P[3, 1] == 0.0
    assert list(g["mul_label_idx"][:3]) == [0, 0, 2]         # first maximum wins


def test_expit_and_softmax_properties():
    z = np.array([-800.0, -40.0, -1.0, 0.0, 1.0, 40.0, 800.0])
    p = orc.expit(z)
    from scipy.special import expit as sp_expit
    np.testing.assert_allclose(p, sp_expit(z), rtol=1e-15, atol=0)
    s = orc.softmax(np.array([[1.0, 2.0, 3.0], [1000.0, 1000.0, -1000.0]]))
    np.testing.assert_allclose(s.sum(1), 1.0)
    np.testing.assert_allclose(s[1], [0.5, 0.5, 0.0])


def test_loss_gradient_matches_sklearn(golden):
    g = golden("loss_grad.npz")
    X, lam = g["X"], float(g["l2"])
    loss, gW, gb = orc.loss_gradient(g["bin_W"], g["bin_b"], X, g["bin_y"], lam)
    np.testing.assert_allclose(loss, g["bin_loss"], rtol=1e-12)
    np.testing.assert_allclose(np.concatenate([gW[0], gb]), g["bin_grad"], rtol=1e-10, atol=1e-14)
    loss, gW, gb = orc.loss_gradient(g["mul_W"], g["mul_b"], X, g["mul_y"], lam)
    np.testing.assert_allclose(loss, g["mul_loss"], rtol=1e-12)
    np.testing.assert_allclose(np.concatenate([gW, gb[:, None]], 1), g["mul_grad"], rtol=1e-10, atol=1e-14)
