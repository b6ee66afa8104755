#!/usr/bin/env python
"""Step time (forward + backward of the module, CUDA-graph replays, L2 flushed between steps) and error against the
fp64 oracle's sampled rows for the three ways a batch can go through the library:

    fp32 inputs, two-term rows on the tensor cores (NTX_BF16X2; the module's choice for large fp32 batches)
    fp32 inputs, fp32 rows on the CUDA cores      (NTXENT_FP32_TC=0)
    bf16 inputs, bf16 rows on the tensor cores

    python tools/fp32_tc_times.py [n] [d] [tau] [steps]
"""
import os
import statistics
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ntxent_oracle as orc  # noqa: E402
from simclr_b200 import NT_Xent  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
d = int(sys.argv[2]) if len(sys.argv) > 2 else 128
tau = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 20
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(3)
v0 = torch.randn(n, d, generator=g)
v1 = torch.randn(n, d, generator=g) + 0.5 * v0
sample = np.arange(0, 2 * n, max(1, 2 * n // 256))                       # rows the oracle checks
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def run(label, dtype, env):
    os.environ["NTXENT_FP32_TC"] = env
    a = v0.to(dtype).to(dev).requires_grad_(True)
    b = v1.to(dtype).to(dev).requires_grad_(True)
    crit = NT_Xent(tau)

    def step():
        a.grad = b.grad = None
        loss = crit(a, b)
        loss.backward()
        return loss

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    a.grad = b.grad = None
    with torch.cuda.graph(graph, capture_error_mode="thread_local"):
        loss = step()
    times = []
    for _ in range(steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        graph.replay()
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) * 1e3)
    w = orc.weights_for("mean", n)
    ref = orc.ntxent_sampled_fp64(a.detach().float().cpu().numpy(), b.detach().float().cpu().numpy(), tau, w, sample)
    got = torch.cat([a.grad, b.grad]).float().cpu().numpy()[sample]
    gerr = np.abs(got - ref["dz"]).max() / np.abs(ref["dz"]).max()
    print(f"{label:46s} {statistics.median(times):9.1f} us/step   gradient error (sampled rows) {gerr:.2e}   "
          f"loss {float(loss.detach()):.6f}")


if __name__ == "__main__":
    print(f"N = {n} pairs, D = {d}, temperature {tau}, {steps} graph replays, L2 flushed between steps")
    run("fp32 inputs, two-term rows, tensor cores", torch.float32, "1")
    run("fp32 inputs, fp32 rows, CUDA cores", torch.float32, "0")
    run("bf16 inputs, bf16 rows, tensor cores", torch.bfloat16, "1")
