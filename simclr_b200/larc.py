"""Drop-in for ``apex.parallel.LARC`` around ``torch.optim.SGD`` on B200.

The reference builds its optimizer as (main.py:85-94)

    optimizer = torch.optim.SGD(params=exclude_from_wt_decay(...), lr=..., momentum=..., nesterov=False, weight_decay=0.)
    optimizer = LARC(optimizer=optimizer, trust_coefficient=0.001, clip=False)

and uses ``optimizer.optim`` (for the LR scheduler), ``optimizer.param_groups``, ``zero_grad()`` and ``step()``
(main.py:96-116).  ``apex`` cannot be installed on the B200 box; this class has the same constructor and the same
surface, so only the import line of ``main.py`` changes.  ``step()`` is ONE cooperative kernel launch of
``libntxent_b200.so`` (``ntx_larc_sgd_step``) over every parameter tensor: layer-wise trust ratio, weight decay folded
into the gradient, SGD momentum update -- no host synchronisation (apex's ``if param_norm != 0 and grad_norm != 0``
syncs twice per tensor).  Momentum buffers live in the wrapped optimizer's ``state[p]["momentum_buffer"]`` as SGD
keeps them, so ``state_dict()`` / ``load_state_dict()`` round-trip with plain ``torch.optim.SGD``.

fp32 CUDA parameters only (what the reference trains); anything else raises -- there is no CPU path.
"""
from __future__ import annotations

import ctypes

import torch

from . import _cabi

CHUNK_ELEMS = 2048          # elements per work item of the kernel (8 KB of each tensor): ~5 items per CTA on a ResNet-18


class LARC:
    """
    :param optimizer: a ``torch.optim.SGD`` (``nesterov=False``, ``dampening=0``); its per-group ``lr``, ``momentum`` and
        ``weight_decay`` are read at every step, like apex does.
    :param trust_coefficient: trust coefficient for the layer-wise adaptive rate (apex default 0.02; the reference passes 0.001).
    :param clip: apex's clipping mode, ``min(adaptive_lr / lr, 1)`` (apex default True; the reference passes False).
    :param eps: added to the denominator of the adaptive rate (apex default 1e-8).
    """

    def __init__(self, optimizer: torch.optim.Optimizer, trust_coefficient: float = 0.02, clip: bool = True,
                 eps: float = 1e-8) -> None:
        if not isinstance(optimizer, torch.optim.SGD):
            raise TypeError("simclr_b200.larc.LARC fuses LARC with torch.optim.SGD; got " + type(optimizer).__name__)
        self.optim = optimizer
        self.trust_coefficient = trust_coefficient
        self.eps = eps
        self.clip = clip
        self._cache = {}            # group index -> (pointer signature, table, chunks, workspace, tensors kept alive)

    # ---- the surface apex.parallel.LARC forwards to the wrapped optimizer
    def __getstate__(self):
        return self.optim.__getstate__()

    def __setstate__(self, state):
        self.optim.__setstate__(state)

    @property
    def state(self):
        return self.optim.state

    def __repr__(self):
        return self.optim.__repr__()

    @property
    def param_groups(self):
        return self.optim.param_groups

    @param_groups.setter
    def param_groups(self, value):
        self.optim.param_groups = value

    def state_dict(self):
        return self.optim.state_dict()

    def load_state_dict(self, state_dict):
        self.optim.load_state_dict(state_dict)
        self._cache.clear()

    def zero_grad(self, *args, **kwargs):
        self.optim.zero_grad(*args, **kwargs)

    def add_param_group(self, param_group):
        self.optim.add_param_group(param_group)
        self._cache.clear()

    # ---- the fused step
    def _plan(self, gi: int, group, params):
        """Device tables for one parameter group, rebuilt only when a pointer changed (gradients are re-allocated
        by zero_grad(set_to_none=True))."""
        # cheap per-step check: the gradients' addresses (parameters and momentum buffers only move through
        # load_state_dict / add_param_group, which drop the cache)
        sig = (len(params), group["weight_decay"]) + tuple(p.grad.data_ptr() for p in params)
        hit = self._cache.get(gi)
        if hit is not None and hit[0] == sig:
            return hit
        state = self.optim.state
        bufs = []
        for p in params:
            st = state[p]
            buf = st.get("momentum_buffer")
            if buf is None:                                   # zero buffer == SGD's first step (buf = grad)
                buf = st["momentum_buffer"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            bufs.append(buf)
        dev = params[0].device
        table = (_cabi.NtxLarcParam * len(params))()
        chunk_list = []
        for t, (p, b) in enumerate(zip(params, bufs)):
            n = p.numel()
            nch = -(-n // CHUNK_ELEMS)
            e = table[t]
            e.param, e.grad, e.momentum = p.data_ptr(), p.grad.data_ptr(), b.data_ptr()
            e.numel, e.weight_decay = n, float(group["weight_decay"])
            e.first_chunk, e.n_chunks = len(chunk_list) // 2, nch
            for c in range(nch):
                chunk_list += (t, c)
        raw = torch.frombuffer(bytearray(bytes(table)), dtype=torch.uint8).clone()
        table_dev = raw.to(dev)
        chunks_dev = torch.tensor(chunk_list, dtype=torch.int32).to(dev)
        n_chunks = len(chunk_list) // 2
        ws = torch.zeros(_cabi.load().ntx_larc_workspace_bytes(len(params), n_chunks), dtype=torch.uint8, device=dev)
        hit = (sig, table_dev, chunks_dev, ws, n_chunks)
        self._cache[gi] = hit
        return hit

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.optim.param_groups):
            if group.get("nesterov", False) or group.get("dampening", 0) != 0 or group.get("maximize", False):
                raise ValueError("the fused LARC step implements SGD with nesterov=False, dampening=0, maximize=False "
                                 "(main.py:85-91)")
            params = [p for p in group["params"] if p.grad is not None]
            if not params:
                continue
            for p in params:
                if p.device.type != "cuda" or p.dtype != torch.float32 or p.grad.dtype != torch.float32:
                    raise TypeError("the fused LARC step needs fp32 CUDA parameters and gradients (no CPU path)")
                if not (p.is_contiguous() and p.grad.is_contiguous()) and not (
                        p.is_contiguous(memory_format=torch.channels_last) and
                        p.grad.is_contiguous(memory_format=torch.channels_last)):
                    raise ValueError("parameters and gradients must be dense and share a memory layout")
            with torch.cuda.device(params[0].device):
                _, table_dev, chunks_dev, ws, n_chunks = self._plan(gi, group, params)
                _cabi.Backend().larc_sgd_step(table_dev, len(params), chunks_dev, n_chunks, CHUNK_ELEMS,
                                              group["lr"], group["momentum"], self.trust_coefficient, self.eps,
                                              self.clip, ws)
        return loss
