"""The oracle against the reference's own outputs (tests/golden) and known answers."""
import math

import numpy as np
import pytest
import torch

from oracle import ntxent_oracle as orc


from helpers import grad_err, rel_err as _rel


def test_fp64_rows_match_reference_golden(golden_case):
    c = golden_case
    w = orc.weights_for(c.reduction, c.n, c.grad_out if c.reduction == "none" else None)
    out = orc.ntxent_rows_fp64(c.view0, c.view1, c.tau, weights=w)
    loss = orc.reduce_rows(out["loss"], c.reduction)
    # the golden numbers are fp32 outputs of the reference: tolerance is fp32 rounding
    assert _rel(loss, c.loss) < 2e-6
    assert grad_err(out["dview0"], c.dview0, c) < 2e-5
    assert grad_err(out["dview1"], c.dview1, c) < 2e-5


def test_reference_form_matches_reference_golden(golden_case):
    c = golden_case
    a = torch.tensor(c.view0, requires_grad=True)
    b = torch.tensor(c.view1, requires_grad=True)
    loss = orc.ntxent_reference_form(a, b, c.tau, c.reduction)
    if c.reduction == "none":
        loss.backward(torch.tensor(c.grad_out))
    else:
        loss.backward()
    assert _rel(loss.detach().numpy(), c.loss) < 1e-6
    assert grad_err(a.grad.numpy(), c.dview0, c) < 1e-5
    assert grad_err(b.grad.numpy(), c.dview1, c) < 1e-5


def test_known_answers():
    # identical unit rows / all-zero rows: every row loss is ln(2N-1)  
    for fill in (1.0, 0.0):
        v = np.full((4, 8), fill)
        rows = orc.ntxent_rows_fp64(v, v, 0.1)["loss"]
        assert rows.shape == (2, 4)
        assert np.allclose(rows, math.log(7.0), atol=1e-12)
    # N = 1: one logit, loss 0
    one = orc.ntxent_rows_fp64(np.random.randn(1, 8), np.random.randn(1, 8), 0.1)["loss"]
    assert np.allclose(one, 0.0, atol=1e-12)
    # reductions: mean == sum / (2N); none.sum() == sum
    rng = np.random.default_rng(0)
    a, b = rng.standard_normal((16, 12)), rng.standard_normal((16, 12))
    rows = orc.ntxent_rows_fp64(a, b, 0.5)["loss"]
    assert math.isclose(orc.reduce_rows(rows, "mean"), orc.reduce_rows(rows, "sum") / 32)
    assert math.isclose(orc.reduce_rows(rows, "none").sum(), orc.reduce_rows(rows, "sum"))


def test_ctor_argument_checks_of_port():
    a = torch.randn(4, 8)
    with pytest.raises(AssertionError):
        orc.ntxent_reference_form(a, a, 0.0)
    with pytest.raises(AssertionError):
        orc.ntxent_reference_form(a, a, 0.5, "avg")


def test_index_maps_single_rank_is_reference_layout():
    self_col, pos_col = orc.index_maps(5)
    assert self_col.tolist() == list(range(10))
    assert pos_col.tolist() == [5, 6, 7, 8, 9, 0, 1, 2, 3, 4]      # (r + N) mod 2N


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_rows_equal_reference_on_concatenated_batch(world):
    # rank-major sharding has the same negatives as the reference on the
    # concatenation of all ranks' view0 and all ranks' view1.
    name = f"shard_w{world}"
    from conftest import GOLDEN
    c = next(g for g in GOLDEN if g.name == name)
    full = orc.ntxent_rows_fp64(c.view0, c.view1, c.tau)["loss"]            # (2, N)
    nl = c.n // world
    chunks = []
    for k in range(world):
        zk = np.concatenate([c.view0[k * nl:(k + 1) * nl], c.view1[k * nl:(k + 1) * nl]])
        chunks.append(orc.l2_normalise(zk)[0])
    keys = np.concatenate(chunks)
    for k in range(world):
        sl = slice(k * nl, (k + 1) * nl)
        rows = orc.ntxent_rows_fp64(c.view0[sl], c.view1[sl], c.tau, keys=keys, rank=k, world=world)["loss"]
        assert np.allclose(rows, full[:, sl], atol=1e-12)
        self_col, pos_col = orc.index_maps(nl, world, k)
        assert self_col[0] == k * 2 * nl and pos_col[0] == k * 2 * nl + nl
    # mean over ranks of the per-rank mean equals the reference's global mean loss
    assert math.isclose(full.mean(), float(c.loss), rel_tol=2e-6)


def test_closed_form_gradient_matches_autograd_fp64():
    rng = np.random.default_rng(3)
    a, b = rng.standard_normal((20, 10)), rng.standard_normal((20, 10))
    w = rng.random((2, 20)) + 0.1
    out = orc.ntxent_rows_fp64(a, b, 0.3, weights=w)
    ta = torch.tensor(a, requires_grad=True)
    tb = torch.tensor(b, requires_grad=True)
    loss = orc.ntxent_reference_form(ta, tb, 0.3, "none")
    loss.backward(torch.tensor(w))
    assert _rel(out["loss"], loss.detach().numpy()) < 1e-12
    assert _rel(out["dview0"], ta.grad.numpy()) < 1e-11
    assert _rel(out["dview1"], tb.grad.numpy()) < 1e-11


def test_sampled_gradient_equals_full_closed_form():
    # the large-size checker (all-row lse, gradients of a subset of rows) against the full fp64 closed form,
    # with a clamped (all-zero) row and non-uniform row weights
    rng = np.random.default_rng(1)
    n, d = 300, 40
    a, b = rng.standard_normal((n, d)), rng.standard_normal((n, d))
    a[3] = 0.0
    w = rng.uniform(0.5, 1.5, (2, n))
    full = orc.ntxent_rows_fp64(a, b, 0.3, weights=w)
    rows = np.array([0, 3, 5, 299, 300, 303, 599, 17])
    part = orc.ntxent_sampled_fp64(a, b, 0.3, w, rows, chunk=128)
    want = np.concatenate([full["dview0"], full["dview1"]])[rows]
    assert np.abs(part["loss"] - full["loss"]).max() < 1e-12
    assert _rel(part["dz"][rows != 3], want[rows != 3]) < 1e-12
    assert _rel(part["dz"][rows == 3], want[rows == 3]) < 1e-12        # the eps-clamped row (huge gradient)
