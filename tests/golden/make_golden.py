"""Generate the golden vectors under tests/golden/ by RUNNING THE REFERENCE'S OWN IMPLEMENTATION
of the path in this container: scikit-learn's `LogisticRegression.decision_function /
predict_proba / predict` (the calls behind REF/main.py:21-22) and, for the refit row,
`LinearModelLoss.loss_gradient` (SK/linear_model/_linear_loss.py:291-378).

    python tests/golden/make_golden.py          # rewrites the .npz files next to this script

The reference ships no fixtures, and scikit-learn is a third-party
dependency rather than part of /root/reference, so these files are what pins the oracle.
Made with scikit-learn 1.9.0 / numpy 2.3.5 (the reference pins 0.24.1, REF/requirements.txt:11,
which cannot be installed here; the predict-side function is the same for a default
`LogisticRegression()`).  Inputs are regenerated from seeds by `tests/synth.py`; each file
stores a SHA-256 of them.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import synth  # noqa: E402


def outputs(clf, X):
    dec = clf.decision_function(X)
    proba = clf.predict_proba(X)
    pred = clf.predict(X)
    idx = np.searchsorted(clf.classes_, pred) if clf.classes_.dtype.kind != "O" else \
        np.array([list(clf.classes_).index(p) for p in pred])
    assert dec.dtype == np.float64 and proba.dtype == np.float64
    return dict(decision=dec, proba=proba, label_idx=idx.astype(np.int32))


def save(name, **arrays):
    path = os.path.join(HERE, name)
    np.savez_compressed(path, **arrays)
    print(f"{name}: {os.path.getsize(path) / 1024:.1f} KiB")


def main():
    import sklearn

    meta = np.array([sklearn.__version__, np.__version__])

    # --- cfg 0: the notebook's model ------------------------------------------------------
    clf, (Xtr, Xte, ytr, yte) = synth.fit_notebook_model()
    score = clf.score(Xte, yte)
    assert score == 0.9666666666666667, score       # REF/Logistic Regression.ipynb:13
    Xiris, yiris = synth.iris_uci()
    out = outputs(clf, Xiris)
    save("iris_model.npz", versions=meta, coef=np.ascontiguousarray(clf.coef_),
         intercept=clf.intercept_, classes=np.array(list(clf.classes_)),
         n_iter=clf.n_iter_, notebook_score=np.float64(score),
         X=Xiris.astype(np.float64), y=np.array(list(yiris)),
         X_train=Xtr.astype(np.float64), y_train=np.array(list(ytr)), **out)

    rng = np.random.default_rng(0)
    X10k = (rng.random((10_000, 4), dtype=np.float32) * 8.0).astype(np.float32)
    out = outputs(clf, X10k)
    save("iris_batch10k.npz", versions=meta, x_sha=np.array(synth.sha(X10k)), **out)

    # --- cfg 1..3 at oracle-sized n ---------------------------------------------------------
    for name, n in (("cfg1_d32_binary", 4096), ("cfg2_d100k_binary", 48), ("cfg3_d1024_c1000", 48)):
        X, W, b = synth.make_config(name, n)
        out = outputs(synth.sklearn_model(W, b), X)
        save(f"{name}.npz", versions=meta, n=np.int64(n),
             x_sha=np.array(synth.sha(X)), model_sha=np.array(synth.sha(W, b)), **out)

    # --- adversarial rows -------------------------------------------------------------------
    X, W, b = synth.adversarial_binary()
    ob = outputs(synth.sklearn_model(W, b), X)
    X2, W2, b2 = synth.adversarial_multi()
    om = outputs(synth.sklearn_model(W2, b2), X2)
    save("adversarial.npz", versions=meta,
         bin_X=X, bin_W=W, bin_b=b, **{f"bin_{k}": v for k, v in ob.items()},
         mul_X=X2, mul_W=W2, mul_b=b2, **{f"mul_{k}": v for k, v in om.items()})

    # --- refit row: loss and gradient from sklearn's own LinearModelLoss --------------------
    from sklearn._loss.loss import HalfBinomialLoss, HalfMultinomialLoss
    from sklearn.linear_model._linear_loss import LinearModelLoss

    rng = np.random.default_rng(11)
    n, d = 512, 32
    Xg = rng.standard_normal((n, d))
    Wb, bb = synth.make_model(d, 1, 12)
    yb = (rng.random(n) < 0.5).astype(np.float64)
    lam = 1.0 / n
    lml = LinearModelLoss(base_loss=HalfBinomialLoss(), fit_intercept=True)
    loss_b, grad_b = lml.loss_gradient(np.concatenate([Wb[0], bb]), Xg, yb,
                                       sample_weight=np.full(n, 1.0 / n), l2_reg_strength=lam)
    C = 5
    Wm, bm = synth.make_model(d, C, 13)
    ym = rng.integers(0, C, n).astype(np.float64)
    lmm = LinearModelLoss(base_loss=HalfMultinomialLoss(n_classes=C), fit_intercept=True)
    coef0 = np.asfortranarray(np.concatenate([Wm, bm[:, None]], axis=1))
    loss_m, grad_m = lmm.loss_gradient(coef0, Xg, ym, sample_weight=np.full(n, 1.0 / n),
                                       l2_reg_strength=lam)
    save("loss_grad.npz", versions=meta, X=Xg, l2=np.float64(lam),
         bin_W=Wb, bin_b=bb, bin_y=yb, bin_loss=np.float64(loss_b), bin_grad=np.asarray(grad_b),
         mul_W=Wm, mul_b=bm, mul_y=ym, mul_loss=np.float64(loss_m),
         mul_grad=np.ascontiguousarray(grad_m))


if __name__ == "__main__":
    main()
