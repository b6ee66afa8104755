"""Host-side mirror of the reference module: constructor, state, argument errors."""
import pytest
import torch

from simclr_b200 import NT_Xent


def test_defaults_match_reference_constructor():
    m = NT_Xent()
    assert m.temperature == 0.1 and m.reduction == "mean" and m.device == torch.device("cpu")   # loss.py:9-10
    assert m.gather is False and m.compute_dtype is None
    assert list(m.parameters()) == [] and list(m.buffers()) == [] and m.state_dict() == {}


def test_constructor_asserts_like_reference():
    for bad in (0.0, -1.0):
        with pytest.raises(AssertionError):      # loss.py:16
            NT_Xent(temperature=bad)
    with pytest.raises(AssertionError):          # loss.py:17
        NT_Xent(reduction="avg")
    NT_Xent(temperature=0.5, device=0)           # main.py:83 passes an int device
    NT_Xent(0.5, "sum", "cuda:0")
    with pytest.raises(TypeError):
        NT_Xent(compute_dtype=torch.float16)


def test_new_arguments_are_keyword_only():
    with pytest.raises(TypeError):
        NT_Xent(0.5, "mean", torch.device("cpu"), True)


def test_env_switches(monkeypatch):
    monkeypatch.setenv("NTXENT_GATHER", "1")
    monkeypatch.setenv("NTXENT_COMPUTE_DTYPE", "bf16")
    m = NT_Xent()
    assert m.gather is True and m.compute_dtype == torch.bfloat16
    assert NT_Xent(gather=False).gather is False


def test_shape_and_dtype_errors():
    m = NT_Xent(0.5)
    with pytest.raises(IndexError):              # reference: mask shape mismatch -> IndexError
        m(torch.randn(4, 8), torch.randn(5, 8))
    with pytest.raises(RuntimeError):            # reference: matmul shape mismatch -> RuntimeError
        m(torch.randn(4, 8), torch.randn(4, 9))
    with pytest.raises(RuntimeError):
        m(torch.randn(2, 4, 8), torch.randn(2, 4, 8))
    with pytest.raises(TypeError):               # not a floating-point batch
        m(torch.ones(4, 8, dtype=torch.int64), torch.ones(4, 8, dtype=torch.int64))
    with pytest.raises(RuntimeError, match="no CPU path"):      # float64 / float16 are taken (through fp32): CUDA only
        m(torch.randn(4, 8).double(), torch.randn(4, 8).double())


def test_float64_and_float16_inputs_go_through_the_fp32_path(monkeypatch):
    """The reference returns the input dtype for float64 / float16 batches; so does the module,
    computing in fp32 with differentiable casts around the call (checked here with the CPU stand-in for the kernels)."""
    import numpy as np
    import simclr_b200.loss as loss_mod
    from oracle import ntxent_oracle as orc
    from standin_backend import StandInBackend
    monkeypatch.setattr(loss_mod, "_backend", StandInBackend())
    monkeypatch.setattr(loss_mod, "_require_cuda", lambda t: None)
    g = torch.Generator().manual_seed(3)
    v0, v1 = torch.randn(24, 16, generator=g), torch.randn(24, 16, generator=g)
    for dtype, tol in ((torch.float64, 1e-6), (torch.float16, 2e-3)):
        a, b = v0.to(dtype).requires_grad_(True), v1.to(dtype).requires_grad_(True)
        ref = orc.ntxent_rows_fp64(a.detach().double().numpy(), b.detach().double().numpy(), 0.5,
                                   weights=orc.weights_for("mean", 24))
        loss = NT_Xent(0.5)(a, b)
        assert loss.dtype == dtype and loss.dim() == 0
        loss.backward()
        assert a.grad.dtype == dtype and abs(float(loss.detach()) - ref["loss"].mean()) < tol * ref["loss"].mean()
        assert np.abs(a.grad.double().numpy() - ref["dview0"]).max() < tol * np.abs(ref["dview0"]).max() + 1e-4 * (dtype == torch.float16)
        rows = NT_Xent(0.5, "none").forward_cat(torch.cat([a.detach(), b.detach()]))
        assert rows.dtype == dtype and rows.shape == (2, 24)


def test_no_cpu_fallback():
    with pytest.raises(RuntimeError, match="no CPU path"):
        NT_Xent(0.5)(torch.randn(4, 8), torch.randn(4, 8))


def test_rows_layout_choice(monkeypatch):
    """Which rows the module asks the library for (pure host logic): the input dtype, except bf16 outside the
    tensor-core kernels' range -> fp32 rows, and large fp32 batches at temperature >= 0.35 -> two-term rows."""
    import torch
    from simclr_b200 import NT_Xent, _cabi
    from simclr_b200 import loss as loss_mod
    monkeypatch.delenv("NTXENT_FP32_TC", raising=False)
    monkeypatch.delenv("NTXENT_COMPUTE_DTYPE", raising=False)
    f32, bf16 = torch.empty(4, 128), torch.empty(4, 128, dtype=torch.bfloat16)
    big = loss_mod.FP32_TC_MIN_KEYS
    assert NT_Xent(0.5)._rows_dtype(f32, big) is _cabi.BF16X2
    assert NT_Xent(0.5)._rows_dtype(f32, big - 2) is torch.float32            # small batch: CUDA-core kernels
    assert NT_Xent(0.2)._rows_dtype(f32, big) is torch.float32                # rounding grows like 1 / temperature
    assert NT_Xent(0.35)._rows_dtype(f32, big) is _cabi.BF16X2
    assert NT_Xent(0.5)._rows_dtype(torch.empty(4, 129), big) is torch.float32    # one term must fit a 128-wide tile
    assert NT_Xent(0.5)._rows_dtype(bf16, big) is torch.bfloat16
    assert NT_Xent(0.5, compute_dtype=torch.bfloat16)._rows_dtype(f32, big) is torch.bfloat16
    assert NT_Xent(0.5, compute_dtype=torch.float32)._rows_dtype(bf16, big) is _cabi.BF16X2
    assert NT_Xent(0.5)._rows_dtype(torch.empty(4, 300, dtype=torch.bfloat16), big) is torch.float32
    assert NT_Xent(0.02)._rows_dtype(bf16, big) is torch.float32
    monkeypatch.setenv("NTXENT_FP32_TC", "0")
    assert NT_Xent(0.5)._rows_dtype(f32, big) is torch.float32
    assert _cabi.storage_dtype(_cabi.BF16X2) is torch.bfloat16 and _cabi.storage_dtype(torch.float32) is torch.float32
