"""NCCL run of the row-sharded loss with the real kernels (needs >= 2 GPUs; skipped otherwise).
Oracle: the reference on the concatenation of every rank's view0 and view1."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, name, dtype_name, out, backend="nccl", steps=1):
    import torch.distributed as dist
    from simclr_b200 import NT_Xent, distributed_utils as du
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    du.init_ddp({"distributed": {"local_rank": rank, "world_size": world, "dist_url": f"tcp://127.0.0.1:{port}"}})
    dev = torch.device("cuda", rank)
    c = next(g for g in GOLDEN if g.name == name)
    nl = c.n // world
    sl = slice(rank * nl, (rank + 1) * nl)
    dtype = getattr(torch, dtype_name)
    a = torch.tensor(c.view0[sl], device=dev, requires_grad=True)
    b = torch.tensor(c.view1[sl], device=dev, requires_grad=True)
    crit = NT_Xent(temperature=c.tau, reduction="mean", device=rank, gather=True, compute_dtype=dtype,
                   gather_backend=backend)
    for _ in range(steps):                       # several steps: the peer backend alternates two key buffers
        a.grad = b.grad = None
        loss = crit(a, b)
        loss.backward()
    torch.cuda.synchronize()
    out[rank] = {"loss": float(loss.detach()), "g0": a.grad.cpu().numpy(), "g1": b.grad.cpu().numpy()}
    dist.barrier()
    du.cleanup()


@pytest.mark.parametrize("backend,steps", [("nccl", 1), ("peer", 3)])
@pytest.mark.parametrize("dtype_name,tol", [("float32", 1e-5), ("bfloat16", 1e-2)])
def test_sharded_loss_over_nccl(dtype_name, tol, backend, steps):
    """backend "peer": the same loss with the keys and the per-row vectors exchanged by peer-memory stores
    from the kernels (ntx_l2norm_peers / ntx_scatter_peers) instead of NCCL all-gathers."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2
    name = "shard_w2"
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), name, dtype_name, out, backend, steps), nprocs=world, join=True)
    c = next(g for g in GOLDEN if g.name == name)
    nl = c.n // world
    assert abs(np.mean([out[k]["loss"] for k in range(world)]) - float(c.loss)) < tol * float(c.loss)
    scale = np.abs(c.dview0).max()
    for k in range(world):
        sl = slice(k * nl, (k + 1) * nl)
        assert np.abs(out[k]["g0"] / world - c.dview0[sl]).max() < tol * scale
        assert np.abs(out[k]["g1"] / world - c.dview1[sl]).max() < tol * scale


def _worker_fp32_tc(rank, world, port, n, d, tau, out, backend="nccl", steps=1):
    import torch.distributed as dist
    from simclr_b200 import NT_Xent, _cabi, distributed_utils as du
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    du.init_ddp({"distributed": {"local_rank": rank, "world_size": world, "dist_url": f"tcp://127.0.0.1:{port}"}})
    dev = torch.device("cuda", rank)
    g = torch.Generator().manual_seed(21)
    v0 = torch.randn(n, d, generator=g)
    v1 = torch.randn(n, d, generator=g) + 0.5 * v0
    nl = n // world
    a = v0[rank * nl:(rank + 1) * nl].to(dev).requires_grad_(True)
    b = v1[rank * nl:(rank + 1) * nl].to(dev).requires_grad_(True)
    crit = NT_Xent(temperature=tau, reduction="mean", device=rank, gather=True, gather_backend=backend)
    picked = crit._rows_dtype(a, 2 * nl * world) is _cabi.BF16X2
    for _ in range(steps):
        a.grad = b.grad = None
        loss = crit(a, b)
        loss.backward()
    torch.cuda.synchronize()
    out[rank] = {"loss": float(loss.detach()), "g0": a.grad.cpu().numpy(), "g1": b.grad.cpu().numpy(), "picked": picked}
    dist.barrier()
    du.cleanup()


@pytest.mark.parametrize("backend,steps", [("nccl", 1), ("peer", 3)])
def test_sharded_fp32_inputs_on_tensor_cores(backend, steps):
    """fp32 inputs, 2048 gathered keys on two ranks: the module holds the rows as two bf16 terms (NTX_BF16X2), the
    all-gather or the kernels' peer-memory stores move them like any row buffer, and loss and gradients match the
    fp64 oracle at the fp32 tolerance."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    from oracle import ntxent_oracle as orc
    world, n, d, tau = 2, 1024, 96, 0.5
    g = torch.Generator().manual_seed(21)
    v0 = torch.randn(n, d, generator=g)
    v1 = torch.randn(n, d, generator=g) + 0.5 * v0
    ref = orc.ntxent_rows_fp64(v0.numpy(), v1.numpy(), tau, weights=orc.weights_for("mean", n))
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_fp32_tc, args=(world, _free_port(), n, d, tau, out, backend, steps), nprocs=world, join=True)
    nl = n // world
    want = float(ref["loss"].mean())
    scale = max(np.abs(ref["dview0"]).max(), np.abs(ref["dview1"]).max())
    for k in range(world):
        assert out[k]["picked"]
        sl = slice(k * nl, (k + 1) * nl)
        # each rank's loss is the mean over ITS rows; its gradient that of the sum over ranks of those means
        assert abs(out[k]["loss"] - float(ref["loss"][:, sl].mean())) < 1e-5 * want
        assert np.abs(out[k]["g0"] / world - ref["dview0"][sl]).max() < 1e-5 * scale
        assert np.abs(out[k]["g1"] / world - ref["dview1"][sl]).max() < 1e-5 * scale
