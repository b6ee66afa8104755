"""Context measurements for BASELINE.md (not the bench contract): this repo's module vs the
reference-form port run on the SAME B200 through stock PyTorch, at several BASELINE configs.
Test infrastructure: imports oracle/."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ntxent_oracle as orc  # noqa: E402
from simclr_b200 import NT_Xent  # noqa: E402

dev = torch.device("cuda:0")


def timeit(fn, iters):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def case(n, d, dtype, tau=0.5, iters=50, ref=True):
    g = torch.Generator().manual_seed(7)
    v0 = torch.randn(n, d, generator=g).to(dtype).to(dev).requires_grad_(True)
    v1 = torch.randn(n, d, generator=g).to(dtype).to(dev).requires_grad_(True)
    crit = NT_Xent(tau)

    def ours():
        v0.grad = v1.grad = None
        crit(v0, v1).backward()

    t_ours = timeit(ours, iters)
    t_ref = None
    if ref:
        r0 = v0.detach().float().requires_grad_(True)
        r1 = v1.detach().float().requires_grad_(True)

        def theirs():
            r0.grad = r1.grad = None
            orc.ntxent_reference_form(r0, r1, tau, "mean").backward()
        try:
            t_ref = timeit(theirs, max(3, iters // 5))
        except torch.OutOfMemoryError:
            t_ref = None
    flops = 18.0 * n * n * d
    print(f"N={n:6d} D={d:4d} {str(dtype).split('.')[-1]:8s} ours {t_ours * 1e3:9.1f} us ({n / t_ours / 1e3:8.2f} M pairs/s, "
          f"{flops / t_ours / 1e9:7.1f} TFLOP/s by 18N^2D)"
          + (f" | reference form on this GPU (PyTorch fp32, eager) {t_ref * 1e3:9.1f} us -> x{t_ref / t_ours:.1f}" if t_ref else ""),
          flush=True)


if __name__ == "__main__":
    case(256, 128, torch.float32)            # config 1 shape
    case(512, 128, torch.float32)            # config 2/3 loss size
    case(512, 128, torch.bfloat16)
    case(2048, 128, torch.bfloat16)
    case(8192, 128, torch.bfloat16, iters=30)    # config 4 on one GPU (eager launches)
    case(8192, 128, torch.float32, iters=5, ref=False)
    case(8192, 256, torch.bfloat16, iters=20, ref=False)
    case(16384, 256, torch.bfloat16, iters=10, ref=False)   # half of config 5 on one GPU
    case(32768, 256, torch.bfloat16, iters=5, ref=False)    # config 5 on ONE GPU
