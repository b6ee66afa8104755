"""Process-group helpers with the reference's surface plus the cross-rank row gather.

``init_ddp(conf)`` and ``cleanup()`` keep the names and argument meaning of the
reference's distributed_utils.py:8-24 (``conf["distributed"]["local_rank" |
"world_size" | "dist_url"]``, NCCL backend, ``torch.cuda.set_device(rank)``).

``gather_rows`` / ``GatherRows`` are new: the reference has no cross-rank collective
on the loss path at all.  They build the rank-major key
buffer the sharded NT-Xent kernels expect: chunk k = rows of rank k.
"""
from __future__ import annotations

from typing import Any, Mapping

import torch
import torch.distributed as dist


def init_ddp(conf: Mapping[str, Any], backend: str = "nccl") -> None:
    rank = conf["distributed"]["local_rank"]
    world_size = conf["distributed"]["world_size"]
    dist_url = conf["distributed"]["dist_url"]

    dist.init_process_group(backend=backend, init_method=dist_url, world_size=world_size, rank=rank)
    if backend == "nccl":
        torch.cuda.set_device(rank)


def cleanup() -> None:
    dist.destroy_process_group()


def gather_rows(rows: torch.Tensor, group=None) -> torch.Tensor:
    """All-gather equally sized row blocks into one rank-major buffer (no autograd)."""
    world = dist.get_world_size(group)
    out = torch.empty((world * rows.shape[0],) + tuple(rows.shape[1:]), dtype=rows.dtype, device=rows.device)
    dist.all_gather_into_tensor(out, rows.contiguous(), group=group)
    return out


class GatherRows(torch.autograd.Function):
    """Differentiable row gather: all-gather forward, SUM reduce-scatter backward.

    Every rank consumes the full gathered buffer, so the gradient of a rank's rows is
    the sum over ranks of the gradients each rank holds for that chunk.  With the
    usual DDP 1/W gradient averaging downstream this reproduces the single-process
    gradient on the concatenated batch.
    """

    @staticmethod
    def forward(ctx, rows, group=None):
        ctx.group = group
        ctx.local_rows = rows.shape[0]
        return gather_rows(rows, group)

    @staticmethod
    def backward(ctx, grad):
        grad = grad.contiguous()
        out = torch.empty((ctx.local_rows,) + tuple(grad.shape[1:]), dtype=grad.dtype, device=grad.device)
        dist.reduce_scatter_tensor(out, grad, op=dist.ReduceOp.SUM, group=ctx.group)
        return out, None


def gather_rows_with_grad(rows: torch.Tensor, group=None) -> torch.Tensor:
    return GatherRows.apply(rows, group)
