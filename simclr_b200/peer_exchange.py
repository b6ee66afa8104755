"""Cross-rank exchange over NVLink peer memory for the gathered NT-Xent (one node, <= 8 ranks).

The reference has no cross-rank gather (distributed_utils.py:8-24 is only init_ddp / cleanup); the one
BASELINE.json's north_star asks for is an all-gather of the normalised rows.  Here every rank's key
buffer lives in symmetric memory (the same allocation made
on every rank and mapped into every process, ``torch.distributed._symmetric_memory``); the normalise
kernel stores its rows directly into all of them (``ntx_l2norm_peers``), the per-row vectors of the
backward go out the same way (``ntx_scatter_peers``), and one cross-rank barrier kernel on the stream
orders the stores against the readers.  No collective-library launch, no staging copy.

Hazards and how they are closed (all on the caller's stream, so the whole step can sit in a CUDA graph):
 * keys of step t are read by the backward of step t on every rank, while a fast rank may already
   write step t+1: two key buffers, used alternately.  A rank cannot reach step t+2 before every rank
   has passed the barrier of step t+1, i.e. has finished the backward of step t.
 * the per-row vectors are written after the forward of step t+1, i.e. after that step's first
   barrier, which every rank passes only once its backward of step t is done: one buffer is enough.
 * inside a CUDA graph capture the buffer choice is frozen into the graph, so every replay writes the
   same key buffer; there an extra barrier before the stores makes every rank finish its previous
   backward first (three barrier kernels per step instead of two).
"""
from __future__ import annotations

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
    """Symmetric buffers and their peer addresses for one problem shape and process group."""

    def __init__(self, group, q_rows: int, ld: int, rows_dtype: torch.dtype, device: torch.device):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        if self.world > MAX_PEERS:
            raise RuntimeError(f"peer-memory gather spans one node: world {self.world} > {MAX_PEERS}")
        self.q_rows, self.ld, self.rows_dtype = q_rows, ld, rows_dtype
        name = self.group.group_name
        self.keys, self.key_ptrs, self.key_hdl = [], [], []
        for _ in range(2):
            buf = symm_mem.empty((self.world * q_rows, ld), dtype=rows_dtype, device=device)
            hdl = symm_mem.rendezvous(buf, group=name)
            self.keys.append(buf)
            self.key_hdl.append(hdl)
            self.key_ptrs.append([int(p) for p in hdl.buffer_ptrs])
        self.vec = symm_mem.empty((self.world, 2, q_rows), dtype=torch.float32, device=device)
        self.vec_hdl = symm_mem.rendezvous(self.vec, group=name)
        self.vec_ptrs = [int(p) for p in self.vec_hdl.buffer_ptrs]
        self.step = 0
        self.held = [False, False]          # key buffer saved for a backward that has not run yet

    def gather_rows(self, be, view0, view1, inv_norm, hold: bool):
        """Normalise this rank's rows into every rank's key buffer.  Returns (the local, complete key buffer,
        its index); ``hold`` marks it as saved for a backward call, which must ``release`` it."""
        which = self.step & 1
        if self.held[which]:
            raise RuntimeError(
                "NT_Xent(gather_backend='peer') keeps two key buffers and the one due now is still saved for a "
                "backward call that has not run: at most two forward calls may be outstanding. Use "
                "gather_backend='nccl' for other call patterns.")
        self.step += 1
        self.held[which] = hold
        if torch.cuda.is_current_stream_capturing():
            # a replayed graph writes the SAME buffer every step, so the alternation cannot protect a slow
            # rank that still reads it in its previous backward: wait until every rank has got here
            self.key_hdl[which].barrier(channel=2)
        be.l2norm_peers(view0, view1, self.key_ptrs[which], self.rank, self.ld, self.rows_dtype, inv_norm)
        self.key_hdl[which].barrier(channel=0)
        return self.keys[which], which

    def release(self, which: int) -> None:
        self.held[which] = False

    def gather_vectors(self, be, mine: torch.Tensor) -> torch.Tensor:
        """``mine`` = (2, q_rows) [row sums ; row weights] of this rank -> (world, 2, q_rows) of every rank."""
        be.scatter_peers(mine, self.vec_ptrs, self.rank * 2 * self.q_rows * 4)
        self.vec_hdl.barrier(channel=1)
        return self.vec


_exchanges: dict = {}


def exchange_for(group, q_rows: int, ld: int, rows_dtype: torch.dtype, device: torch.device) -> PeerExchange:
    """One PeerExchange per (group, shape); creating it is a collective call (all ranks, same order)."""
    key = (id(group), q_rows, ld, rows_dtype, device)
    ex = _exchanges.get(key)
    if ex is None:
        ex = _exchanges[key] = PeerExchange(group, q_rows, ld, rows_dtype, device)
    return ex


def close_all() -> None:
    """Drop every PeerExchange (symmetric buffers and their rendezvous handles).  Call it on all ranks, after the
    CUDA graphs that captured the buffers have been destroyed and before ``destroy_process_group``."""
    _exchanges.clear()
