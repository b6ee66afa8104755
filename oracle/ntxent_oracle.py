"""CPU oracle for the NT-Xent hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this module, and only as the checker or
as the timed CPU arm.  Nothing under ``simclr_b200/`` imports it: the product
path is the CUDA library and fails loudly without it.

Parity pin: the reference ships no tests or golden vectors for this path
(SURVEY.md section 8c), so the oracle is pinned against outputs of the
reference itself: ``tests/golden/make_golden.py`` imports
``/root/reference/loss.py`` in the build container, runs it on seeded inputs
and commits the input/output vectors under ``tests/golden/``;
``tests/test_oracle.py`` checks every function here against those vectors.

Restatements of the same algorithm that live here:

* :func:`ntxent_rows_fp64` -- numpy float64, row-chunked closed form (never
  materialises more than ``chunk x 2N`` logits), with the closed-form
  gradient.  It is the checker at sizes where the reference's O(N^2)
  temporaries do not fit.
* :func:`ntxent_sampled_fp64` -- the same closed form with every row's ``lse`` but the gradient
  of a chosen subset of rows only (large-size checks: full loss, sampled gradients).
* :func:`ntxent_reference_form` -- torch, dtype-following, autograd backward;
  it walks the reference's own operator sequence (three N x N similarity
  blocks, diagonal removal, concatenation, softmax cross-entropy,
  ``loss.py:33-65``) and is the timed CPU arm because it has the reference's
  cost profile.
* :func:`index_maps` -- integer formulas for the self column and the
  positive column of every query row, single-rank and rank-major sharded.
"""
from __future__ import annotations

import numpy as np

EPS = 1e-12  # F.normalize's eps, loss.py:33-34


# --------------------------------------------------------------------------
# index arithmetic (bit-exact part of the contract)
# --------------------------------------------------------------------------
def index_maps(n_local: int, world: int = 1, rank: int = 0):
    """Self column and positive column of each of this rank's 2*n_local rows.

    Rows are ordered [view0 rows ; view1 rows] (loss.py:51-52 puts the
    positive of row i of view0 at view1's row i and vice versa).  With more
    than one rank the key buffer is rank-major: chunk k holds
    [view0_k ; view1_k], so a row's self and positive columns both live in
    its own rank's chunk.
    """
    q_rows = 2 * n_local
    r = np.arange(q_rows, dtype=np.int64)
    row_off = rank * q_rows
    self_col = row_off + r
    pos_col = row_off + (r + n_local) % q_rows
    return self_col.astype(np.int32), pos_col.astype(np.int32)


# --------------------------------------------------------------------------
# float64 closed form
# --------------------------------------------------------------------------
def l2_normalise(z: np.ndarray):
    """x / max(||x||_2, eps) per row -- loss.py:33-34."""
    z = np.asarray(z, dtype=np.float64)
    norm = np.sqrt((z * z).sum(axis=1, keepdims=True))
    denom = np.maximum(norm, EPS)
    return z / denom, denom[:, 0]


def ntxent_rows_fp64(view0, view1, temperature: float, weights=None, chunk: int = 1024,
                     keys=None, n_local: int | None = None, rank: int = 0, world: int = 1):
    """Per-row NT-Xent loss (and closed-form gradient) in float64.

    Follows loss.py:33-65 mathematically: rows are L2-normalised, logits are
    cosine/temperature, the self column is excluded (the boolean mask at
    loss.py:38,46-47), the target is the other view of the same sample
    (``targets = arange(N)`` into the [sim01 | sim00'] concatenation,
    loss.py:37,51-52), the row loss is logsumexp - positive (loss.py:56-61).

    Returns dict with ``loss`` (2, N) as ``reduction="none"`` lays it out
    (loss.py:55-58), ``lse``, ``pos`` and, when ``weights`` (2, N) is given,
    ``dview0``/``dview1``: the gradient of ``sum(weights * loss)``.

    Sharded use: pass this rank's ``view0``/``view1`` plus ``keys`` = the
    rank-major concatenation of every rank's NORMALISED rows; the gradient is
    then only the query-side part and is not returned (use the single-process
    form on the concatenated batch for gradients).
    """
    v0 = np.asarray(view0, dtype=np.float64)
    v1 = np.asarray(view1, dtype=np.float64)
    if v0.ndim != 2 or v0.shape != v1.shape:
        raise ValueError("view0 and view1 must both be (N, D)")
    n = v0.shape[0]
    z = np.concatenate([v0, v1], axis=0)
    zh, denom = l2_normalise(z)
    m = 2 * n
    sharded = keys is not None
    kmat = np.asarray(keys, dtype=np.float64) if sharded else zh
    self_col, pos_col = index_maps(n, world, rank) if sharded else index_maps(n)
    inv_t = 1.0 / float(temperature)

    lse = np.empty(m)
    pos = np.empty(m)
    for lo in range(0, m, chunk):
        hi = min(m, lo + chunk)
        s = (zh[lo:hi] @ kmat.T) * inv_t
        rr = np.arange(lo, hi)
        pos[lo:hi] = s[rr - lo, pos_col[lo:hi]]
        s[rr - lo, self_col[lo:hi]] = -np.inf
        mx = s.max(axis=1, keepdims=True)
        lse[lo:hi] = mx[:, 0] + np.log(np.exp(s - mx).sum(axis=1))
    row_loss = lse - pos
    out = {"loss": row_loss.reshape(2, n), "lse": lse, "pos": pos, "zhat": zh, "denom": denom}
    if weights is None or sharded:
        return out

    w = np.asarray(weights, dtype=np.float64).reshape(m)
    # d/dzhat of sum_r w_r * loss_r :  (diag(w)(P - Y) + (P - Y)^T diag(w)) zhat / tau
    dzh = np.zeros_like(zh)
    for lo in range(0, m, chunk):
        hi = min(m, lo + chunk)
        s = (zh[lo:hi] @ zh.T) * inv_t
        rr = np.arange(lo, hi)
        s[rr - lo, self_col[lo:hi]] = -np.inf
        p = np.exp(s - lse[lo:hi, None])           # P_rc, zero on the self column
        p[rr - lo, pos_col[lo:hi]] -= 1.0          # P - Y
        g = p * w[lo:hi, None]                     # rows of diag(w)(P - Y)
        dzh[lo:hi] += g @ zh                       # query-side term
        dzh += g.T @ zh[lo:hi]                     # key-side term (transpose)
    dzh *= inv_t
    # normalisation Jacobian of x / max(||x||, eps)
    norm = np.sqrt((z * z).sum(axis=1))
    proj = (zh * dzh).sum(axis=1, keepdims=True)
    dz = np.where((norm > EPS)[:, None], (dzh - zh * proj) / denom[:, None], dzh / EPS)
    out["dview0"], out["dview1"] = dz[:n], dz[n:]
    return out


def ntxent_sampled_fp64(view0, view1, temperature: float, weights, rows, chunk: int = 2048):
    """All-row ``lse`` / ``pos`` / row loss plus the closed-form gradient of ONLY the rows in ``rows``.

    The checker at sizes where the full gradient pass of :func:`ntxent_rows_fp64` is too slow
    (BASELINE configs[3] inside ``bench.py``, configs[4] in the GPU tests; SURVEY.md section 8c:
    "full loss + gradients on a fixed random subset of rows and all-row lse").  Same mathematics,
    following loss.py:33-65: with ``P_rc = exp(s_rc - lse_r)`` (zero on the self column) and
    ``p(r)`` the positive of row ``r``,

        d/dzhat_r sum_q w_q loss_q = (1/tau) [ sum_c (w_r P_rc + w_c P_cr) zhat_c - (w_r + w_p(r)) zhat_p(r) ]

    and ``s`` is symmetric, so row ``r`` of ``S`` gives both ``P_rc`` and ``P_cr = exp(s_rc - lse_c)``
    once every row's ``lse`` is known.  ``rows`` index the concatenation [view0 ; view1].
    The matrix products run in torch float64 (threaded BLAS); everything else is numpy float64.
    ``weights`` is one (2, N) array or a list of them (several weightings share the all-row pass).
    Returns dict: ``loss`` (2, N), ``lse``, ``pos``, ``rows``, ``dz`` (len(rows), D) -- a list of such
    arrays when a list of weights was given.
    """
    import torch
    v0 = np.asarray(view0, dtype=np.float64)
    v1 = np.asarray(view1, dtype=np.float64)
    n = v0.shape[0]
    m = 2 * n
    z = np.concatenate([v0, v1], axis=0)
    zh, denom = l2_normalise(z)
    inv_t = 1.0 / float(temperature)
    zt = torch.from_numpy(zh)
    lse_t = torch.empty(m, dtype=torch.float64)
    pos_t = torch.empty(m, dtype=torch.float64)
    for lo in range(0, m, chunk):
        hi = min(m, lo + chunk)
        s = (zt[lo:hi] @ zt.T) * inv_t
        rr = torch.arange(lo, hi)
        pos_t[lo:hi] = s[rr - lo, (rr + n) % m]
        s[rr - lo, rr] = -float("inf")                          # self column (mask of loss.py:38,46-47)
        lse_t[lo:hi] = torch.logsumexp(s, dim=1)
    lse, pos = lse_t.numpy(), pos_t.numpy()
    out = {"loss": (lse - pos).reshape(2, n), "lse": lse, "pos": pos}
    rows = np.asarray(rows, dtype=np.int64)
    many = isinstance(weights, (list, tuple))
    ws = [np.asarray(w, dtype=np.float64).reshape(m) for w in (weights if many else [weights])]
    dzs = [np.empty((len(rows), zh.shape[1])) for _ in ws]
    for lo in range(0, len(rows), chunk):
        r = rows[lo:lo + chunk]
        s = (zt[r] @ zt.T).numpy() * inv_t                      # rows r of S == columns r of S
        i = np.arange(len(r))
        p_row, p_col = np.exp(s - lse[r, None]), np.exp(s - lse[None, :])     # P_rc and P_cr
        p = (r + n) % m
        zr = zh[r]
        norm = np.sqrt((z[r] * z[r]).sum(axis=1))
        for w, dz in zip(ws, dzs):
            coef = w[r, None] * p_row + w[None, :] * p_col
            coef[i, r] = 0.0                                    # self column (mask of loss.py:38,46-47)
            coef[i, p] -= w[r] + w[p]                           # the two one-hot terms of the positive pair
            dzh = (torch.from_numpy(coef) @ zt).numpy() * inv_t
            proj = (zr * dzh).sum(axis=1, keepdims=True)
            dz[lo:lo + len(r)] = np.where((norm > EPS)[:, None], (dzh - zr * proj) / denom[r, None], dzh / EPS)
    out["rows"], out["dz"] = rows, (dzs if many else dzs[0])
    return out


def reduce_rows(row_loss_2n: np.ndarray, reduction: str):
    """loss.py:54-65: 'none' -> (2, N); 'sum'; 'mean' = sum / N * 0.5."""
    if reduction == "none":
        return row_loss_2n
    total = row_loss_2n.sum()
    if reduction == "sum":
        return total
    if reduction == "mean":
        return total / row_loss_2n.shape[1] * 0.5
    raise AssertionError(reduction)


def weights_for(reduction: str, n: int, grad_output=None):
    """Row weights w such that backward of the reduced loss = grad of sum(w*loss)."""
    if reduction == "none":
        g = np.ones((2, n)) if grad_output is None else np.asarray(grad_output, dtype=np.float64)
        return g.reshape(2, n)
    g = 1.0 if grad_output is None else float(grad_output)
    scale = 1.0 if reduction == "sum" else 0.5 / n
    return np.full((2, n), g * scale)


# --------------------------------------------------------------------------
# torch restatement with the reference's operator sequence (timed CPU arm)
# --------------------------------------------------------------------------
def ntxent_reference_form(view0, view1, temperature: float, reduction: str = "mean"):
    """Torch port with the reference's cost profile; differentiable by autograd.

    Operator sequence of loss.py:33-65: normalise both views; three N x N
    similarity blocks divided by the temperature; the two same-view blocks
    lose their diagonal through a boolean mask; each view's logits are the
    cross-view block followed by its same-view block, so the target of row i
    is column i; softmax cross-entropy per view; reduction.
    """
    import torch
    import torch.nn.functional as F

    if not temperature > 0.0:
        raise AssertionError("temperature must be positive")      # loss.py:16
    if reduction not in ("none", "mean", "sum"):
        raise AssertionError("bad reduction")                     # loss.py:17
    a = F.normalize(view0, p=2, dim=1)
    b = F.normalize(view1, p=2, dim=1)
    n = a.shape[0]
    off_diag = ~torch.eye(n, dtype=torch.bool, device=a.device)
    cross = (a @ b.t()) / temperature
    logits = []
    for same, x in ((a, cross), (b, cross.t())):
        intra = ((same @ same.t()) / temperature)[off_diag].view(n, n - 1)
        logits.append(torch.cat([x, intra], dim=1))
    target = torch.arange(n, device=a.device)
    if reduction == "none":
        return torch.stack([F.cross_entropy(l, target, reduction="none") for l in logits])
    total = sum(F.cross_entropy(l, target, reduction="sum") for l in logits)
    return total if reduction == "sum" else total / n * 0.5
