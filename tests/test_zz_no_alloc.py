"""ntx_fwd_static: the forward for callers that cannot supply scratch memory.

Some hosts own only their inputs and outputs: a bare C / ctypes caller without a device
allocator, or code that runs where cudaMalloc is forbidden (it stalls the whole device and may
not be called while a stream is being captured).  For them the library keeps a fixed 1 MiB
scratch array on the device.  This test calls that entry point through the C ABI twice in a row
(the array keeps the previous call's contents, so the second call proves nothing stale leaks in)
and checks every output against the ordinary entry point, which takes a caller workspace.
It lives in a file that sorts last so that a run stopping at the first failure has already run
every other test; it passes both alone in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest
import torch

from helpers import rel_err

pytestmark = pytest.mark.gpu


def _inputs(n, d, dtype, dev, seed):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(n, d, generator=g).to(dtype).to(dev)
    b = torch.randn(n, d, generator=g).to(dtype).to(dev)
    return a, b


@pytest.mark.parametrize("dtype,tol,n,d", [(torch.float32, 1e-5, 200, 96), (torch.bfloat16, 1e-2, 1500, 128)])
def test_static_scratch_forward_matches_workspace_forward(dtype, tol, n, d):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from simclr_b200 import _cabi
    be = _cabi.Backend()
    dev = torch.device("cuda:0")
    q_rows = 2 * n
    ld = be.row_pitch(d, dtype)
    plan = be.plan(q_rows, q_rows, d, dtype)
    ws = torch.empty(plan.fwd_ws_bytes, dtype=torch.uint8, device=dev)
    for call, seed in enumerate((11, 12)):                       # two calls in a row, different data
        a, b = _inputs(n, d, dtype, dev, seed)
        rows = torch.empty(q_rows, ld, dtype=dtype, device=dev)
        inv = torch.empty(q_rows, device=dev)
        be.l2norm(a, b, rows, inv)
        want = torch.zeros(3, q_rows, device=dev)
        want_sum = torch.zeros(2, device=dev)
        be.fwd(rows, rows, d, (n, 1, 0), 0.5, want[0], want[1], want[2], want_sum, ws)
        got = torch.full((3, q_rows), float("nan"), device=dev)
        got_sum = torch.full((2,), float("nan"), device=dev)
        be.fwd_static(rows, rows, d, (n, 1, 0), 0.5, got[0], got[1], got[2], got_sum)
        torch.cuda.synchronize()
        for name, g_, w_ in (("rowsum", got[0], want[0]), ("pos_logit", got[1], want[1]),
                             ("row_loss", got[2], want[2]), ("loss_sum", got_sum, want_sum)):
            err = rel_err(g_.cpu().numpy(), w_.cpu().numpy())
            assert np.isfinite(g_.cpu().numpy()).all(), f"call {call}: {name} not written"
            assert err < tol, f"call {call}: {name} differs from ntx_fwd by {err:.2e}"


def test_static_scratch_refuses_problems_beyond_its_size():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from simclr_b200 import _cabi
    be = _cabi.Backend()
    dev = torch.device("cuda:0")
    n, d = 1 << 16, 64                                            # 2^17 rows: > 1 MiB of partial row sums
    q_rows = 2 * n
    assert be.plan(q_rows, q_rows, d, torch.bfloat16).fwd_ws_bytes_min > (1 << 20)
    rows = torch.zeros(q_rows, be.row_pitch(d, torch.bfloat16), dtype=torch.bfloat16, device=dev)
    out = torch.zeros(3, q_rows, device=dev)
    out_sum = torch.zeros(2, device=dev)
    with pytest.raises(_cabi.NtxError, match="built-in array"):
        be.fwd_static(rows, rows, d, (n, 1, 0), 0.5, out[0], out[1], out[2], out_sum)
