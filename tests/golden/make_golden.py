"""Generate the golden vectors under tests/golden/ from the REFERENCE itself.

Run in the build container only (``/root/reference`` does not exist on the
GPU box):   python tests/golden/make_golden.py

It imports ``/root/reference/loss.py`` unmodified, runs ``NT_Xent`` forward and
autograd backward on seeded CPU inputs, and stores inputs and outputs as
``ntxent_golden.npz``.  The reference has no tests or fixtures of its own for
this path, so these vectors are the parity pin for
``oracle/`` and, through it, for the CUDA path.
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def _inputs(kind, n, d, seed):
    g = torch.Generator().manual_seed(seed)
    if kind == "randn":
        return torch.randn(n, d, generator=g), torch.randn(n, d, generator=g)
    if kind == "clustered":           # cosine similarities close to 1
        centres = torch.randn(10, d, generator=g)
        pick = torch.randint(0, 10, (n,), generator=g)
        a = centres[pick] + 0.05 * torch.randn(n, d, generator=g)
        b = centres[pick] + 0.05 * torch.randn(n, d, generator=g)
        return a, b
    if kind == "ones":
        return torch.ones(n, d), torch.ones(n, d)
    if kind == "zeros":
        return torch.zeros(n, d), torch.zeros(n, d)
    if kind == "onehot":              # hand-placed positives expose index errors
        a = torch.zeros(n, d)
        b = torch.zeros(n, d)
        for i in range(n):
            a[i, i % d] = 1.0
            b[i, (i * 3 + 1) % d] = 2.0
        return a, b
    raise ValueError(kind)


CASES = [
    # name, kind, N, D, tau, reduction, seed, bf16-rounded inputs?
    ("c1_mean", "randn", 256, 128, 0.5, "mean", 7, False),        # BASELINE configs[0]
    ("small_mean", "randn", 8, 16, 0.1, "mean", 1, False),
    ("small_sum", "randn", 8, 16, 0.1, "sum", 2, False),
    ("small_none", "randn", 8, 16, 0.1, "none", 3, False),
    ("ragged_sum", "randn", 33, 20, 0.5, "sum", 4, False),
    ("ragged_none", "randn", 77, 50, 0.2, "none", 5, False),
    ("clustered_t05", "clustered", 64, 32, 0.5, "mean", 6, False),
    ("clustered_t01", "clustered", 64, 32, 0.1, "mean", 8, False),
    ("tile_edge", "randn", 192, 64, 0.5, "mean", 9, False),       # 2N = 384 = 3 tiles of 128
    ("tile_edge_p1", "randn", 129, 96, 0.07, "mean", 10, False),
    ("bf16_inputs", "randn", 160, 128, 0.5, "mean", 11, True),
    ("bf16_none", "clustered", 96, 64, 0.2, "none", 12, True),
    ("kat_ones", "ones", 4, 8, 0.1, "none", 0, False),            # every row ln(2N-1)
    ("kat_zeros", "zeros", 4, 8, 0.1, "none", 0, False),          # eps clamp -> ln(2N-1)
    ("kat_n1", "randn", 1, 8, 0.1, "mean", 13, False),            # single logit -> 0
    ("kat_onehot", "onehot", 12, 6, 0.5, "none", 0, False),
    ("shard_w2", "randn", 32, 24, 0.5, "mean", 14, False),        # also used by the 2-rank tests
    ("shard_w4", "randn", 64, 40, 0.3, "mean", 15, False),
]


def main():
    sys.path.insert(0, REF)
    from loss import NT_Xent  # the unmodified reference, loss.py:4

    out = {}
    names = []
    for name, kind, n, d, tau, red, seed, as_bf16 in CASES:
        a, b = _inputs(kind, n, d, seed)
        if as_bf16:  # reference run in fp32 on the up-cast of bf16 inputs
            a, b = a.bfloat16().float(), b.bfloat16().float()
        a.requires_grad_(True)
        b.requires_grad_(True)
        loss = NT_Xent(temperature=tau, reduction=red)(a, b)
        if red == "none":
            g = torch.Generator().manual_seed(1000 + seed)
            gout = torch.rand(2, n, generator=g) + 0.25       # non-uniform upstream gradient
            loss.backward(gout)
        else:
            gout = torch.tensor(1.0)
            loss.backward()
        names.append(name)
        out[f"{name}/view0"] = a.detach().numpy()
        out[f"{name}/view1"] = b.detach().numpy()
        out[f"{name}/loss"] = loss.detach().numpy()
        out[f"{name}/grad_out"] = gout.numpy()
        out[f"{name}/dview0"] = a.grad.numpy()
        out[f"{name}/dview1"] = b.grad.numpy()
        out[f"{name}/meta"] = np.array([n, d, tau, {"none": 0, "mean": 1, "sum": 2}[red], int(as_bf16)],
                                       dtype=np.float64)
    out["names"] = np.array(names)
    out["torch_version"] = np.array(torch.__version__)
    path = os.path.join(HERE, "ntxent_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(names), "cases")


if __name__ == "__main__":
    main()
