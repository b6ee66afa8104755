"""CPU oracle for the fused LARC + SGD-momentum step.  TEST INFRASTRUCTURE ONLY (same rules as ntxent_oracle).

What it restates: the optimizer of main.py:85-94 -- ``torch.optim.SGD(momentum, nesterov=False, weight_decay=0.)``
over the two parameter groups of ``exclude_from_wt_decay`` (main.py:18-36), wrapped in
``apex.parallel.LARC(trust_coefficient=0.001, clip=False)`` -- stepped at main.py:116.

``apex`` is a third-party dependency that is NOT vendored in /root/reference and cannot be installed here (the
reference pins no version: ``from apex.parallel.LARC import LARC``, main.py:7).  Its published algorithm
(NVIDIA/apex, apex/parallel/LARC.py, ``LARC.step``) is restated below; PARITY UNPINNED for that file itself.
The SGD half IS pinned: ``tests/test_larc_cpu.py`` runs this oracle against the real ``torch.optim.SGD`` fed with
gradients edited the way LARC.step edits them.

    for each group (lr, momentum, weight_decay), for each p with a gradient:
        param_norm, grad_norm = ||p||_2, ||g||_2
        if param_norm != 0 and grad_norm != 0:
            adaptive_lr = trust_coefficient * param_norm / (grad_norm + param_norm * weight_decay + eps)
            if clip: adaptive_lr = min(adaptive_lr / lr, 1)
            g = (g + weight_decay * p) * adaptive_lr
        # torch.optim.SGD with weight_decay = 0 (LARC zeroes it for the step), dampening = 0, nesterov = False:
        buf = g                      (first step)    |    buf = momentum * buf + g
        p = p - lr * buf
"""
from __future__ import annotations

import numpy as np


def larc_sgd_step(params, grads, bufs, lr: float, momentum: float, weight_decay: float,
                  trust_coefficient: float = 0.001, clip: bool = False, eps: float = 1e-8):
    """One step over the tensors of one parameter group, float64.  ``bufs[i]`` is None before the first step.
    Returns (new params, new momentum buffers)."""
    out_p, out_b = [], []
    for p, g, b in zip(params, grads, bufs):
        p = np.asarray(p, dtype=np.float64)
        g = np.asarray(g, dtype=np.float64)
        pn, gn = np.sqrt((p * p).sum()), np.sqrt((g * g).sum())
        if pn != 0 and gn != 0:
            a = trust_coefficient * pn / (gn + pn * weight_decay + eps)
            if clip:
                a = min(a / lr, 1.0)
            g = (g + weight_decay * p) * a
        b = g.copy() if b is None else momentum * np.asarray(b, dtype=np.float64) + g
        out_p.append(p - lr * b)
        out_b.append(b)
    return out_p, out_b
