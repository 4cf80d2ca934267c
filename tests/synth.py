"""Seeded synthetic models and inputs for the BASELINE.json configs.

Shared by `tests/golden/make_golden.py` (which commits sklearn's float64 answers) and by the
parity tests (which regenerate the same inputs from the same seeds).  numpy's PCG64
`Generator.standard_normal` stream is what the fixtures were made with; each fixture stores a
SHA-256 of the regenerated arrays so a drift in the generator is detected instead of silently
comparing different inputs.
"""
from __future__ import annotations

import hashlib

import numpy as np

IRIS_CLASSES = ["Iris-setosa", "Iris-versicolor", "Iris-virginica"]

# name -> (n_features, n_classes (1 = binary), seed_X, seed_model)
CONFIGS = {
    "cfg1_d32_binary": (32, 1, 1, 2),
    "cfg2_d100k_binary": (100_000, 1, 3, 4),
    "cfg3_d1024_c1000": (1024, 1000, 5, 6),
}


def sha(*arrays) -> str:
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def make_model(d: int, n_scores: int, seed: int):
    """W ~ N(0, 1/d) so logits are O(1); b ~ N(0, 1).  float64, as sklearn stores coef_."""
    rng = np.random.default_rng(seed)
    W = rng.standard_normal((n_scores, d)) / np.sqrt(d)
    b = rng.standard_normal(n_scores)
    return W, b


def make_X(n: int, d: int, seed: int) -> np.ndarray:
    """X ~ N(0,1), float32, C-contiguous."""
    rng = np.random.default_rng(seed)
    return rng.standard_normal((n, d), dtype=np.float32)


def make_config(name: str, n: int):
    d, c, sx, sm = CONFIGS[name]
    W, b = make_model(d, c, sm)
    return make_X(n, d, sx), W, b


def iris_uci():
    """UCI iris as the notebook downloads it (REF/Logistic Regression.ipynb:24-30): sklearn's
    bundled copy with the two rows sklearn corrected put back to their UCI values."""
    from sklearn.datasets import load_iris

    data = load_iris()
    X = data.data.copy()
    X[34] = [4.9, 3.1, 1.5, 0.1]
    X[37] = [4.9, 3.1, 1.5, 0.1]
    y = np.array(IRIS_CLASSES, dtype=object)[data.target]
    return X, y


def fit_notebook_model():
    """The notebook's single cell, offline (REF/Logistic Regression.ipynb:28-34)."""
    from sklearn.linear_model import LogisticRegression
    from sklearn.model_selection import train_test_split

    X, y = iris_uci()
    array = np.empty((150, 5), dtype=object)
    array[:, :4] = X
    array[:, 4] = y
    X_train, X_test, y_train, y_test = train_test_split(
        array[:, 0:4], array[:, 4], test_size=0.20, random_state=1, shuffle=True)
    clf = LogisticRegression()
    clf.fit(X_train, y_train)
    return clf, (X_train, X_test, y_train, y_test)


def sklearn_model(W, b, classes=None):
    """A fitted-looking sklearn LogisticRegression with injected coefficients."""
    from sklearn.linear_model import LogisticRegression

    W = np.atleast_2d(np.asarray(W, dtype=np.float64))
    clf = LogisticRegression()
    clf.coef_ = W
    clf.intercept_ = np.asarray(b, dtype=np.float64).reshape(-1)
    n_classes = 2 if W.shape[0] == 1 else W.shape[0]
    clf.classes_ = np.arange(n_classes) if classes is None else np.asarray(classes)
    clf.n_features_in_ = W.shape[1]
    clf.n_iter_ = np.array([1], dtype=np.int32)
    return clf


def adversarial_binary():
    """d=4 binary model with exact ties, saturating and sub-1e-5 margins (SURVEY.md 8c iii)."""
    W = np.array([[1.0, -2.0, 0.5, 0.25]])
    b = np.array([0.0])
    X = np.array([
        [0.0, 0.0, 0.0, 0.0],          # score exactly 0 -> classes_[0], p = 0.5
        [2.0, 1.0, 0.0, 0.0],          # exact tie again (2 - 2)
        [1000.0, 0.0, 0.0, 0.0],       # saturates to [0, 1]
        [-1000.0, 0.0, 0.0, 0.0],      # saturates to [1, 0]
        [0.0, 0.0, 0.0, 4e-6],         # |margin| = 1e-6 (label exempt)
        [0.0, 0.0, 0.0, -4e-6],
        [1e-30, 0.0, 0.0, 0.0],        # denormal-scale margin in fp32 terms
        [30.0, 0.0, 0.0, 0.0],
        [-30.0, 0.0, 0.0, 0.0],
        [0.5, 0.25, -1.0, 2.0],
        [88.0, 0.0, 0.0, 0.0],         # exp(-88) near the fp32 normal limit
        [-104.0, 0.0, 0.0, 0.0],       # exp underflows in fp32
    ], dtype=np.float32)
    return X, W, b


def adversarial_multi():
    """d=3, C=4 model with equal maxima (first index wins) and saturating rows."""
    W = np.array([[1.0, 0.0, 0.0],
                  [1.0, 0.0, 0.0],
                  [0.0, 1.0, 0.0],
                  [0.0, 0.0, 1.0]])
    b = np.zeros(4)
    X = np.array([
        [1.0, 0.0, 0.0],       # classes 0 and 1 tie -> 0
        [0.0, 0.0, 0.0],       # all tie -> 0
        [0.0, 2.0, 2.0],       # classes 2 and 3 tie -> 2
        [500.0, 0.0, 0.0],     # saturates: 0.5/0.5/0/0
        [0.0, 0.0, 1000.0],    # saturates: one-hot on 3
        [-1000.0, -1000.0, -1000.0],
        [0.3, 0.3 + 1e-6, 0.1],  # sub-1e-5 margin between 0/1 and 2
        [1.5, -0.5, 0.25],
    ], dtype=np.float32)
    return X, W, b
