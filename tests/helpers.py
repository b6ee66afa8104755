"""Comparison helpers shared by the CPU and GPU parity tests."""
import numpy as np


def rel_err(got, want, floor=1e-30):
    """max |got - want| / max(max |want|, floor)."""
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    return float(np.abs(got - want).max() / max(np.abs(want).max(), floor))


def grad_floor(case):
    """Size of the terms a gradient is built from: 0.5 * max row weight / (temperature * ||z||).

    d loss / d z is a difference of terms of size w / (tau * ||z||) (softmax part minus one-hot
    part, through the normalisation Jacobian 1/||z||).  When they cancel exactly (identical rows,
    N = 1) the result is rounding noise of those terms and has to be compared with their size,
    not with zero.  For ordinary cases this floor lies below max |gradient| and changes nothing.
    """
    if case.reduction == "none":
        w = float(np.abs(case.grad_out).max())
    else:
        w = 1.0 if case.reduction == "sum" else 0.5 / case.n
    norms = np.linalg.norm(np.concatenate([case.view0, case.view1]).astype(np.float64), axis=1)
    norm = float(np.median(norms)) or 1.0
    return 0.5 * w / (case.tau * norm)


def grad_err(got, want, case):
    return rel_err(got, want, grad_floor(case))
