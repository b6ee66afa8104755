"""Cross-rank exchange over NVLink peer memory for the gathered NT-Xent (one node, <= 8 ranks).

The reference has no cross-rank gather (distributed_utils.py:8-24 is only init_ddp / cleanup); the one
BASELINE.json's north_star asks for is an all-gather of the normalised rows.  Here every rank owns one
*exchange block* in symmetric memory (``torch.distributed._symmetric_memory``: the same allocation on every
rank, mapped into every process): control words, the row sums and row weights of every key, and two key
buffers.  The library's kernels move and order the data themselves (include/ntxent_b200.h, "Multi-rank
exchange"): the normalise kernel stores its rows into every block, the forward's finalize its row sums, the
backward's first kernel its row weights; each producer then writes the step number into a flag word of every
block, and each consumer polls the flag words of its own block before it first reads a rank's data.  There is
no barrier kernel, no collective-library launch and no staging copy, so a step is six kernels that can sit in
one CUDA graph.

Hazards and how they are closed:
 * keys of step t are read by the backward of step t on every rank while a fast rank may already write
   step t+1: two key buffers.  The step count lives on the device (a word of the block, advanced by the
   normalise kernel), so a replayed CUDA graph alternates buffers like eager calls do.  A rank cannot reach
   step t+2 before every rank has sent its rows of step t+1, i.e. has finished the backward of step t.
 * row sums and weights of step t+1 are written after the sender's forward of step t+1, which waited for
   every rank's rows of that step, i.e. for the end of every rank's backward of step t: one buffer is enough.
 * at most two forward calls may be outstanding without their backward (checked on the host).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

MAX_PEERS = 8


def available() -> bool:
    try:
        import torch.distributed._symmetric_memory  # noqa: F401
    except Exception:
        return False
    return torch.cuda.is_available()


class PeerExchange:
    """The exchange block of this rank, the peer addresses of every rank's block, for one problem shape."""

    def __init__(self, be, group, q_rows: int, ld: int, rows_dtype: torch.dtype, device: torch.device):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        if self.world > MAX_PEERS:
            raise RuntimeError(f"peer-memory gather spans one node: world {self.world} > {MAX_PEERS}")
        self.q_rows, self.ld, self.rows_dtype = q_rows, ld, rows_dtype
        nbytes = be.peer_block_bytes(q_rows, ld, rows_dtype, self.world)
        self.block = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
        self.block.zero_()                                   # flags, step count and key buffers start at zero
        self.handle = symm_mem.rendezvous(self.block, group=self.group.group_name)
        # NVSwitch multicast mapping of the blocks when the platform has one (0 otherwise): one multimem.st per datum
        # instead of one store per rank.  NTXENT_PEER_MULTICAST=0 turns it off.
        mc = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        if os.environ.get("NTXENT_PEER_MULTICAST", "1").lower() in ("0", "false", "no", "off"):
            mc = 0
        self.multicast = mc
        self.blocks = be.peer_blocks([int(p) for p in self.handle.buffer_ptrs], self.rank, mc)
        torch.cuda.synchronize(device)
        dist.barrier(group=self.group)                       # every block is zeroed before any rank stores into one
        self.outstanding = 0                                 # forwards whose backward has not run yet

    def begin_forward(self, needs_backward: bool) -> None:
        if self.outstanding >= 2:
            raise RuntimeError(
                "NT_Xent(gather_backend='peer') keeps two key buffers: at most two forward calls may be outstanding "
                "without their backward. Use gather_backend='nccl' for other call patterns.")
        if needs_backward:
            self.outstanding += 1

    def end_backward(self) -> None:
        self.outstanding = max(0, self.outstanding - 1)


_exchanges: dict = {}


def exchange_for(be, group, q_rows: int, ld: int, rows_dtype: torch.dtype, device: torch.device) -> PeerExchange:
    """One PeerExchange per (group, shape); creating it is a collective call (all ranks, same order)."""
    key = (id(group), q_rows, ld, rows_dtype, device)
    ex = _exchanges.get(key)
    if ex is None:
        ex = _exchanges[key] = PeerExchange(be, group, q_rows, ld, rows_dtype, device)
    return ex


def close_all() -> None:
    """Drop every PeerExchange (symmetric blocks and their rendezvous handles).  Call it on all ranks, after the
    CUDA graphs that captured the blocks have been destroyed and before ``destroy_process_group``."""
    _exchanges.clear()
