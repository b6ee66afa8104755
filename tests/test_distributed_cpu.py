"""world_size-2 gloo tests of the multi-rank path: the differentiable row gather and the
NT_Xent(gather=True) autograd wiring (keys, row sums and row weights gathered; per-rank mean).
The CUDA kernels are replaced by tests/standin_backend.py here; the same scenario runs on NCCL
with the real kernels in tests/test_gpu_multi.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import GOLDEN

WORLD = 2


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _init(rank, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    from simclr_b200 import distributed_utils as du
    du.init_ddp({"distributed": {"local_rank": rank, "world_size": WORLD,
                                 "dist_url": f"tcp://127.0.0.1:{port}"}}, backend="gloo")
    return du


def _worker_loss(rank, port, reduction, out):
    _init(rank, port)
    import simclr_b200.loss as loss_mod
    from standin_backend import StandInBackend
    loss_mod._backend = StandInBackend()                            # test-only: kernels on CPU tensors
    loss_mod._require_cuda = lambda t: None                         # ... which the module otherwise refuses
    c = next(g for g in GOLDEN if g.name == "shard_w2")
    nl = c.n // WORLD
    sl = slice(rank * nl, (rank + 1) * nl)
    a = torch.tensor(c.view0[sl], requires_grad=True)
    b = torch.tensor(c.view1[sl], requires_grad=True)
    crit = loss_mod.NT_Xent(temperature=c.tau, reduction=reduction, gather=True)
    loss = crit(a, b)
    if reduction == "none":
        g = torch.full((2, nl), 0.5 + rank)                         # ranks hold different upstream gradients
        loss.backward(g)
    else:
        loss.backward()
    # DDP would average gradients over ranks; the reference gradient is for the GLOBAL mean loss
    res = {"loss": loss.detach().numpy().copy(), "g0": a.grad.numpy().copy(), "g1": b.grad.numpy().copy()}
    out[rank] = res
    dist.barrier()
    dist.destroy_process_group()


def _spawn(fn, *args):
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(fn, args=(_free_port(), *args, out), nprocs=WORLD, join=True)
    return dict(out)


def test_gathered_loss_matches_reference_on_concatenated_batch():
    out = _spawn(_worker_loss, "mean")
    c = next(g for g in GOLDEN if g.name == "shard_w2")
    nl = c.n // WORLD
    # mean over ranks of the per-rank mean loss == the reference's loss on the concatenated batch
    assert abs(np.mean([out[k]["loss"] for k in range(WORLD)]) - float(c.loss)) < 1e-6 * float(c.loss)
    for k in range(WORLD):
        sl = slice(k * nl, (k + 1) * nl)
        scale = np.abs(c.dview0).max()
        # each rank's loss is a mean over ITS rows -> gradients are W x the global-mean gradients (DDP's 1/W undoes it)
        assert np.abs(out[k]["g0"] / WORLD - c.dview0[sl]).max() < 1e-5 * scale
        assert np.abs(out[k]["g1"] / WORLD - c.dview1[sl]).max() < 1e-5 * scale


def test_gathered_loss_with_per_rank_row_weights():
    from oracle import ntxent_oracle as orc
    out = _spawn(_worker_loss, "none")
    c = next(g for g in GOLDEN if g.name == "shard_w2")
    nl = c.n // WORLD
    w = np.concatenate([np.full((2, nl), 0.5 + k) for k in range(WORLD)], axis=1)       # (2, N) global weights
    ref = orc.ntxent_rows_fp64(c.view0, c.view1, c.tau, weights=w)
    for k in range(WORLD):
        sl = slice(k * nl, (k + 1) * nl)
        assert np.abs(out[k]["loss"] - ref["loss"][:, sl]).max() < 1e-5
        scale = np.abs(ref["dview0"]).max()
        assert np.abs(out[k]["g0"] - ref["dview0"][sl]).max() < 1e-5 * scale
        assert np.abs(out[k]["g1"] - ref["dview1"][sl]).max() < 1e-5 * scale


def _worker_two_term(rank, port, n, d, tau, out):
    _init(rank, port)
    import simclr_b200.loss as loss_mod
    from simclr_b200 import _cabi
    from standin_backend import StandInBackend
    loss_mod._backend = StandInBackend()
    loss_mod._require_cuda = lambda t: None
    seen = []
    orig = StandInBackend.l2norm

    def spy(self, view0, view1, rows, inv_norm, row_scale=1.0, rows_dtype=None):
        seen.append((rows_dtype, rows.dtype, tuple(rows.shape)))
        return orig(self, view0, view1, rows, inv_norm, row_scale, rows_dtype)

    StandInBackend.l2norm = spy
    g = torch.Generator().manual_seed(9)
    v0 = torch.randn(n, d, generator=g)
    v1 = torch.randn(n, d, generator=g) + 0.5 * v0
    nl = n // WORLD
    a = v0[rank * nl:(rank + 1) * nl].clone().requires_grad_(True)
    b = v1[rank * nl:(rank + 1) * nl].clone().requires_grad_(True)
    loss = loss_mod.NT_Xent(temperature=tau, gather=True)(a, b)
    loss.backward()
    out[rank] = {"loss": float(loss.detach()), "g0": a.grad.numpy().copy(), "g1": b.grad.numpy().copy(),
                 "layout": seen[0][0] is _cabi.BF16X2, "store": seen[0][1] == torch.bfloat16, "shape": seen[0][2]}
    dist.barrier()
    dist.destroy_process_group()


def test_gathered_fp32_batch_goes_through_two_term_rows():
    """fp32 inputs with 2048 gathered keys: the module asks for two-term rows (bf16 storage of twice the bf16 pitch),
    names the layout to every library call, gathers them like any row buffer, and the result keeps fp32 accuracy."""
    from oracle import ntxent_oracle as orc
    n, d, tau = 1024, 24, 0.5
    out = _spawn(_worker_two_term, n, d, tau)
    g = torch.Generator().manual_seed(9)
    v0 = torch.randn(n, d, generator=g)
    v1 = torch.randn(n, d, generator=g) + 0.5 * v0
    ref = orc.ntxent_rows_fp64(v0.numpy(), v1.numpy(), tau, weights=orc.weights_for("mean", n))
    nl = n // WORLD
    scale = max(np.abs(ref["dview0"]).max(), np.abs(ref["dview1"]).max())
    for k in range(WORLD):
        assert out[k]["layout"] and out[k]["store"] and out[k]["shape"] == (2 * nl, 128)
        sl = slice(k * nl, (k + 1) * nl)
        assert abs(out[k]["loss"] - float(ref["loss"][:, sl].mean())) < 1e-5 * float(ref["loss"].mean())
        assert np.abs(out[k]["g0"] / WORLD - ref["dview0"][sl]).max() < 1e-5 * scale
        assert np.abs(out[k]["g1"] / WORLD - ref["dview1"][sl]).max() < 1e-5 * scale
