"""fp32 inputs on the tensor cores: rows held as two bf16 terms (NTX_BF16X2 of include/ntxent_b200.h).

The same oracle and tolerance as the fp32 CUDA-core kernels (1e-5 of the value / of the largest gradient entry),
through the C ABI directly (any size) and through the module (which picks the layout for large fp32 batches).
"""
import math

import numpy as np
import pytest
import torch

from helpers import rel_err as _rel_err
from oracle import ntxent_oracle as orc

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-5


def rel_err(got, want):
    return _rel_err(got, want, 1.0)          # losses are O(log 2N) >= 1


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from simclr_b200 import _cabi
    _cabi.load()
    return torch.device("cuda:0")


@pytest.mark.parametrize("n,d,in_dtype", [(257, 128, torch.float32), (100, 33, torch.float32), (64, 64, torch.bfloat16)])
def test_two_term_rows_hold_the_normalised_row(n, d, in_dtype, dev):
    """ntx_l2norm(rows_dtype = NTX_BF16X2): hi + lo reproduces x / max(||x||, eps) * row_scale to 2^-17, hi is the
    bf16 rounding of it, the padding columns are zero."""
    from simclr_b200 import _cabi
    be = _cabi.Backend()
    g = torch.Generator().manual_seed(n)
    a, b = torch.randn(n, d, generator=g).to(in_dtype).to(dev), torch.randn(n, d, generator=g).to(in_dtype).to(dev)
    ld = be.row_pitch(d, _cabi.BF16X2)
    assert ld == 2 * ((d + 63) // 64 * 64)
    for rs in (1.0, 1.7):
        rows = torch.full((2 * n, ld), float("nan"), dtype=torch.bfloat16, device=dev)
        inv = torch.empty(2 * n, device=dev)
        be.l2norm(a, b, rows, inv, row_scale=rs, rows_dtype=_cabi.BF16X2)
        want = torch.nn.functional.normalize(torch.cat([a, b]).float(), dim=1) * rs
        hi, lo = rows[:, :ld // 2].float(), rows[:, ld // 2:].float()
        assert torch.equal(hi[:, :d], want.bfloat16().float()) or (hi[:, :d] - want).abs().max() < 2 ** -8 * rs
        assert ((hi + lo)[:, :d] - want).abs().max() <= 2 ** -17 * rs
        assert torch.count_nonzero(hi[:, d:]) == 0 and torch.count_nonzero(lo[:, d:]) == 0
        assert torch.allclose(inv, 1.0 / torch.cat([a, b]).float().norm(dim=1).clamp_min(1e-12), rtol=1e-6)


@pytest.mark.parametrize("n,d,world,tau,scaled", [
    (1024, 128, 1, 0.5, True), (600, 64, 1, 0.5, True), (513, 100, 1, 0.1, True), (300, 33, 1, 0.2, False),
    (512, 128, 4, 0.5, True), (130, 128, 1, 1.0, False), (2, 8, 1, 0.5, True),
])
def test_two_term_rows_match_fp64_oracle(n, d, world, tau, scaled, dev):
    """Forward and backward on NTX_BF16X2 rows through the C ABI against the fp64 oracle at the fp32 tolerance:
    both tile widths (D <= 64, D <= 128), ragged sizes, one of four ranks against all keys, unit and scaled rows."""
    from simclr_b200 import _cabi
    be = _cabi.Backend()
    X2 = _cabi.BF16X2
    g = torch.Generator().manual_seed(n + d)
    v0 = torch.randn(n, d, generator=g)
    v1 = torch.randn(n, d, generator=g) + 0.5 * v0
    gout = torch.rand(2, n, generator=g) + 0.25                              # reduction "none": one weight per row
    ref = orc.ntxent_rows_fp64(v0.numpy(), v1.numpy(), tau, weights=orc.weights_for("none", n, gout.numpy()))
    nl = n // world
    q_rows = 2 * nl
    ld = be.row_pitch(d, X2)
    plan = be.plan(q_rows, world * q_rows, d, X2)
    assert plan.tensor_cores == 2 and plan.d_pad == ld // 2 and plan.fwd_symmetric == 0
    ws = torch.empty(max(plan.fwd_ws_bytes, plan.bwd_ws_bytes), dtype=torch.uint8, device=dev)
    rs = be.row_scale(tau, X2) if scaled else 1.0
    assert not scaled or abs(rs * rs - math.log2(math.e) / tau) < 1e-4
    keys = torch.empty(world * q_rows, ld, dtype=torch.bfloat16, device=dev)
    inv = torch.empty(world, q_rows, device=dev)
    views, weights = [], []
    for k in range(world):
        a, b = v0[k * nl:(k + 1) * nl].to(dev), v1[k * nl:(k + 1) * nl].to(dev)
        views.append((a, b))
        weights.append(gout[:, k * nl:(k + 1) * nl].contiguous().view(-1).to(dev))
        be.l2norm(a, b, keys[k * q_rows:(k + 1) * q_rows], inv[k], row_scale=rs, rows_dtype=X2)
    st = torch.empty(world, 3, q_rows, device=dev)
    ls = torch.empty(world, 2, device=dev)
    for k in range(world):
        be.fwd(keys[k * q_rows:(k + 1) * q_rows], keys, d, (nl, world, k, rs), tau, st[k, 0], st[k, 1], st[k, 2], ls[k],
               ws, rows_dtype=X2)
    loss_rows = np.concatenate([st[k, 2].cpu().numpy().reshape(2, nl) for k in range(world)], axis=1)
    assert rel_err(loss_rows, ref["loss"]) < FP32_TOL * max(1.0, 0.5 / tau)
    assert abs(float(ls[:, 0].sum()) - ref["loss"].sum()) < FP32_TOL * ref["loss"].sum()
    rowsum_k = st[:, 0, :].reshape(-1).contiguous()
    w_k = torch.cat(weights)
    scale = max(np.abs(ref["dview0"]).max(), np.abs(ref["dview1"]).max())
    for k in range(world):
        a, b = views[k]
        d0, d1 = torch.empty_like(a), torch.empty_like(b)
        be.bwd(keys[k * q_rows:(k + 1) * q_rows], keys, d, (nl, world, k, rs), tau, st[k, 0], rowsum_k, weights[k], w_k, 1,
               1.0, a, b, inv[k], d0, d1, ws, rows_dtype=X2)
        sl = slice(k * nl, (k + 1) * nl)
        e0 = np.abs(d0.cpu().numpy() - ref["dview0"][sl]).max() / scale
        e1 = np.abs(d1.cpu().numpy() - ref["dview1"][sl]).max() / scale
        assert max(e0, e1) < FP32_TOL, (e0, e1)


def test_module_uses_two_term_rows_for_large_fp32_batches(dev, monkeypatch):
    """NT_Xent on fp32 inputs with >= 2048 keys and temperature >= 0.35 picks the two-term tensor-core path (seen in
    the layout handed to the backend), matches the oracle at 1e-5 and the CUDA-core result; smaller batches, smaller
    temperatures and NTXENT_FP32_TC=0 stay on the fp32 kernels."""
    from simclr_b200 import NT_Xent, _cabi, loss as loss_mod
    n, d, tau = 1024, 128, 0.5
    g = torch.Generator().manual_seed(11)
    a = torch.randn(n, d, generator=g)
    b = torch.randn(n, d, generator=g) + 0.5 * a
    ref = orc.ntxent_rows_fp64(a.numpy(), b.numpy(), tau, weights=orc.weights_for("mean", n))
    crit = NT_Xent(tau)
    assert crit._rows_dtype(a, 2 * n) is _cabi.BF16X2
    assert crit._rows_dtype(a[:100], 200) is torch.float32
    assert NT_Xent(0.2)._rows_dtype(a, 2 * n) is torch.float32
    assert NT_Xent(tau, compute_dtype=torch.bfloat16)._rows_dtype(a, 2 * n) is torch.bfloat16
    out = {}
    for flag in ("1", "0"):
        monkeypatch.setenv("NTXENT_FP32_TC", flag)
        ta, tb = a.to(dev).requires_grad_(True), b.to(dev).requires_grad_(True)
        before = loss_mod._backend.launch_count()
        loss = NT_Xent(tau)(ta, tb)
        loss.backward()
        out[flag] = (float(loss.detach()), ta.grad.cpu().numpy(), tb.grad.cpu().numpy(), loss_mod._backend.launch_count() - before)
    scale = max(np.abs(ref["dview0"]).max(), np.abs(ref["dview1"]).max())
    for flag in ("1", "0"):
        l, g0, g1, launches = out[flag]
        assert abs(l - ref["loss"].mean()) < FP32_TOL * ref["loss"].mean()
        assert np.abs(g0 - ref["dview0"]).max() / scale < FP32_TOL and np.abs(g1 - ref["dview1"]).max() / scale < FP32_TOL
        assert launches == 5                    # normalise, forward + finalize, backward + finalize
    # reduction "none" and forward_cat go the same way
    monkeypatch.setenv("NTXENT_FP32_TC", "1")
    z = torch.cat([a, b]).to(dev).requires_grad_(True)
    rows = NT_Xent(tau, "none").forward_cat(z)
    assert rel_err(rows.detach().cpu().numpy(), ref["loss"]) < FP32_TOL
    rows.sum().backward()
    refs = orc.ntxent_rows_fp64(a.numpy(), b.numpy(), tau, weights=orc.weights_for("sum", n))
    scale = max(np.abs(refs["dview0"]).max(), np.abs(refs["dview1"]).max())
    got = z.grad.cpu().numpy()
    assert np.abs(got[:n] - refs["dview0"]).max() / scale < FP32_TOL
    assert np.abs(got[n:] - refs["dview1"]).max() / scale < FP32_TOL


def test_two_term_rows_refusals(dev):
    """D > 128 and temperatures below the tensor-core floor are refused with a message."""
    from simclr_b200 import _cabi
    be = _cabi.Backend()
    with pytest.raises(RuntimeError, match="128"):
        be.plan(256, 256, 200, _cabi.BF16X2)
    a = torch.randn(64, 32, device=dev)
    ld = be.row_pitch(32, _cabi.BF16X2)
    rows = torch.empty(128, ld, dtype=torch.bfloat16, device=dev)
    st = torch.empty(4, 128, device=dev)
    be.l2norm(a, a, rows, st[0], rows_dtype=_cabi.BF16X2)
    plan = be.plan(128, 128, 32, _cabi.BF16X2)
    ws = torch.empty(max(plan.fwd_ws_bytes, plan.bwd_ws_bytes), dtype=torch.uint8, device=dev)
    with pytest.raises(RuntimeError, match="temperature"):
        be.fwd(rows, rows, 32, (64, 1, 0, 1.0), 0.01, st[1], st[2], st[3], torch.empty(2, device=dev), ws,
               rows_dtype=_cabi.BF16X2)


def test_fp32_inputs_at_config4_size_against_sampled_oracle(dev):
    """BASELINE configs[3] size (N = 8192 pairs, D = 128) with fp32 inputs: the module takes the two-term tensor-core
    path; every row's loss and the gradient of 1024 sampled rows under non-uniform row weights against the fp64
    oracle at the fp32 tolerance, plus size-independent properties: swapping the views swaps the halves of the
    result, and the gradient is orthogonal to its input row (the loss does not depend on the rows' norms)."""
    from simclr_b200 import NT_Xent, _cabi
    n, d, tau = 8192, 128, 0.5
    g = torch.Generator().manual_seed(17)
    a = torch.randn(n, d, generator=g)
    b = torch.randn(n, d, generator=g) + 0.25 * a
    i = torch.arange(2 * n, dtype=torch.float64).reshape(2, n)
    w = (0.5 + ((i * 7919) % 1000) / 1000.0).float()                        # non-uniform upstream gradient
    rng = np.random.default_rng(5)
    pairs = np.sort(rng.choice(n, size=512, replace=False))
    rows = np.concatenate([pairs, n + pairs])
    ref = orc.ntxent_sampled_fp64(a.numpy(), b.numpy(), tau, w.double().numpy(), rows)
    crit = NT_Xent(tau, "none")
    assert crit._rows_dtype(a, 2 * n) is _cabi.BF16X2
    ta, tb = a.to(dev).requires_grad_(True), b.to(dev).requires_grad_(True)
    out = crit(ta, tb)
    out.backward(w.to(dev))
    assert rel_err(out.detach().cpu().numpy(), ref["loss"]) < FP32_TOL
    got = np.concatenate([ta.grad[pairs].cpu().numpy(), tb.grad[pairs].cpu().numpy()])
    assert np.abs(got - ref["dz"]).max() / np.abs(ref["dz"]).max() < FP32_TOL
    # views swapped: the (2, N) losses swap their halves (the tiles and slots every row goes through change)
    swapped = crit(tb.detach(), ta.detach())
    assert torch.allclose(swapped[0], out[1].detach(), rtol=0, atol=2e-5) and torch.allclose(swapped[1], out[0].detach(), rtol=0, atol=2e-5)
    # d loss / d z_r is orthogonal to z_r
    dots = (ta.grad * ta.detach()).sum(1).abs().max() / (ta.grad.norm(dim=1) * ta.detach().norm(dim=1)).max()
    assert float(dots) < 1e-4


@pytest.mark.parametrize("d,tau", [(128, 0.5), (128, 0.35), (64, 0.35)])
def test_two_term_rows_on_clustered_inputs(d, tau, dev):
    """SURVEY 8d's second distribution: 10 cluster centres + 0.05 noise (cosines near 1, gradients are small
    differences of nearly equal rows) at the lowest temperature the module sends this way: still inside 1e-5."""
    from simclr_b200 import NT_Xent
    n = 1024
    g = torch.Generator().manual_seed(23)
    centres = torch.randn(10, d, generator=g)
    which = torch.randint(0, 10, (n,), generator=g)
    a = centres[which] + 0.05 * torch.randn(n, d, generator=g)
    b = centres[which] + 0.05 * torch.randn(n, d, generator=g)
    ref = orc.ntxent_rows_fp64(a.numpy(), b.numpy(), tau, weights=orc.weights_for("mean", n))
    crit = NT_Xent(tau)
    from simclr_b200 import _cabi
    assert crit._rows_dtype(a, 2 * n) is _cabi.BF16X2
    ta, tb = a.to(dev).requires_grad_(True), b.to(dev).requires_grad_(True)
    loss = crit(ta, tb)
    loss.backward()
    assert abs(float(loss.detach()) - ref["loss"].mean()) < FP32_TOL * ref["loss"].mean()
    scale = max(np.abs(ref["dview0"]).max(), np.abs(ref["dview1"]).max())
    e0 = np.abs(ta.grad.cpu().numpy() - ref["dview0"]).max() / scale
    e1 = np.abs(tb.grad.cpu().numpy() - ref["dview1"]).max() / scale
    assert max(e0, e1) < FP32_TOL, (e0, e1)
