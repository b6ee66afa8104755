"""Comparison helpers shared by the CPU and GPU parity tests."""
import numpy as np


def rel_err(got, want, floor=1e-30):
    """max |got - want| / max(max |want|, floor)."""
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    return float(np.abs(got - want).max() / max(np.abs(want).max(), floor))


def grad_floor(case):
    """Scale below which a gradient is rounding noise: 1e-3 * max row weight / temperature.

    Gradients of NT-Xent are O(w / (tau * ||z||)); cases whose exact gradient is 0
    (identical or all-zero rows) would otherwise be compared against their own noise.
    """
    if case.reduction == "none":
        w = float(np.abs(case.grad_out).max())
    else:
        w = 1.0 if case.reduction == "sum" else 0.5 / case.n
    return 1e-3 * w / case.tau


def grad_err(got, want, case):
    return rel_err(got, want, grad_floor(case))
