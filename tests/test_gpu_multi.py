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
