"""Parity bar of BASELINE.json's north_star, in one place:
   - class labels bit-exact except on rows whose |margin| is below 1e-5
     (margin = |score| for a binary model, top-1 minus top-2 score otherwise);
   - probabilities within 1e-5 absolute + 1e-5 relative of sklearn's float64 output.
Decision values are not in the north_star; they are held to 2e-5 abs + 1e-5 rel, which is what
keeps the probability bound (|dp/dz| <= 1/4)."""
import numpy as np

P_ATOL = 1e-5
P_RTOL = 1e-5
MARGIN = 1e-5
Z_ATOL = 2e-5
Z_RTOL = 1e-5


def margins(decision64: np.ndarray) -> np.ndarray:
    if decision64.ndim == 1:
        return np.abs(decision64)
    top2 = np.partition(decision64, -2, axis=1)[:, -2:]
    return top2[:, 1] - top2[:, 0]


# The tcgen05 path accumulates in TMEM, whose fp32 adds truncate instead of rounding to nearest:
# 3 x 128 accumulation steps at d=1024 leave up to ~5e-5 on an O(1..6) logit (measured, see
# DESIGN.md "GEMM path: precision").  Probabilities and labels are held to the same bar as everywhere.
Z_ATOL_TENSOR = 1e-4


def assert_parity(out: dict, ref_decision, ref_proba, ref_label, what="", z_atol=Z_ATOL):
    ref_decision = np.asarray(ref_decision, dtype=np.float64)
    if out.get("proba") is not None:
        got = np.asarray(out["proba"], dtype=np.float64)
        assert got.shape == ref_proba.shape, (what, got.shape, ref_proba.shape)
        err = np.abs(got - ref_proba) - (P_ATOL + P_RTOL * np.abs(ref_proba))
        assert err.max() <= 0, f"{what}: proba off by {np.abs(got - ref_proba).max():.3e}"
    if out.get("decision") is not None:
        got = np.asarray(out["decision"], dtype=np.float64)
        assert got.shape == ref_decision.shape, (what, got.shape, ref_decision.shape)
        err = np.abs(got - ref_decision) - (z_atol + Z_RTOL * np.abs(ref_decision))
        assert err.max() <= 0, f"{what}: decision off by {np.abs(got - ref_decision).max():.3e}"
    if out.get("label") is not None:
        got = np.asarray(out["label"]).astype(np.int64)
        safe = margins(ref_decision) >= MARGIN
        bad = np.flatnonzero(safe & (got != np.asarray(ref_label).astype(np.int64)))
        assert bad.size == 0, f"{what}: {bad.size} label mismatches outside the margin band, first row {bad[:5]}"
    if out.get("maxprob") is not None:
        got = np.asarray(out["maxprob"], dtype=np.float64)
        ref = ref_proba.max(axis=1)
        err = np.abs(got - ref) - (P_ATOL + P_RTOL * np.abs(ref))
        assert err.max() <= 0, f"{what}: maxprob off by {np.abs(got - ref).max():.3e}"
