"""mlapi_b200 -- B200-native scoring engine for the hot path behind mlAPI's POST /predict.

    from mlapi_b200 import B200LogisticRegression, load_model
    clf = load_model("LRClassifier.pkl")        # same pickle REF/main.py:19 opens
    clf.predict([[5.1, 3.5, 1.4, 0.2]]); clf.predict_proba([[5.1, 3.5, 1.4, 0.2]]).max()

The arithmetic runs in hand-written sm_100a CUDA kernels behind the C ABI of include/lrscore.h;
there is no CPU fallback.
"""
from .estimator import B200LogisticRegression, load_model  # noqa: F401

__all__ = ["B200LogisticRegression", "load_model"]
