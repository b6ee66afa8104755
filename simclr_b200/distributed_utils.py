"""Process-group helpers with the reference's surface.

``init_ddp(conf)`` and ``cleanup()`` keep the names and argument meaning of the
reference's distributed_utils.py:8-24 (``conf["distributed"]["local_rank" |
"world_size" | "dist_url"]``, NCCL backend, ``torch.cuda.set_device(rank)``).

The reference has no cross-rank collective on the loss path at all.  The gathered loss of
this package does its exchange inside ``NT_Xent`` (``gather=True``): NCCL all-gathers of the rows and of the row
sums, or the kernels' own peer-memory stores (``simclr_b200/peer_exchange.py``).  A differentiable row gather
with a reduce-scatter backward (the north star's other backward variant) is deliberately absent: DESIGN.md
section 5 gives the measured cost that ruled it out.
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
