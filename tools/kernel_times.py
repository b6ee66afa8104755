"""Main-kernel times (CUDA events through the library's measurement hook) and accuracy for a library build.

    NTXENT_B200_LIB=... python tools/kernel_times.py [n] [d] [iters] [world]

world > 1: the shape one rank of `world` sees (n / world pairs of its own against all 2n keys), run on one GPU.
"""
import os
import statistics
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simclr_b200 import _cabi  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
d = int(sys.argv[2]) if len(sys.argv) > 2 else 128
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 30
world = int(sys.argv[4]) if len(sys.argv) > 4 else 1
dev = torch.device("cuda:0")
be = _cabi.Backend()
g = torch.Generator().manual_seed(7)
v0 = torch.randn(n, d, generator=g).bfloat16().to(dev)
v1 = torch.randn(n, d, generator=g).bfloat16().to(dev)
q_rows = 2 * n
ld = be.row_pitch(d, torch.bfloat16)
rows = torch.empty(q_rows, ld, dtype=torch.bfloat16, device=dev)
stats = torch.empty(4, q_rows, device=dev)
loss_sum = torch.empty(2, device=dev)
nl = n // world                      # pairs of "this rank" (rank 0)
ql = 2 * nl
plan = be.plan(ql, q_rows, d, torch.bfloat16)
ws = torch.empty(max(plan.fwd_ws_bytes, plan.bwd_ws_bytes), dtype=torch.uint8, device=dev)
g0, g1 = torch.empty_like(v0), torch.empty_like(v1)
one = torch.ones(1, device=dev)
rs = be.row_scale(0.5, torch.bfloat16) if os.environ.get("NTX_UNIT_ROWS") is None else 1.0
rsk = torch.full((q_rows,), 5000.0, device=dev)          # stand-in row sums of every key (world > 1 timing only)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
tf, tb, tt = [], [], []
for it in range(iters + 5):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
    for e in ev:
        e.record()
    flush.zero_()
    ev[4].record()
    be.l2norm(v0, v1, rows, stats[0], row_scale=rs)
    be.profile_next(0, ev[0], ev[1])
    if world == 1:
        be.fwd(rows, rows, d, (n, 1, 0, rs), 0.5, stats[1], stats[2], stats[3], loss_sum, ws)
        be.profile_next(1, ev[2], ev[3])
        be.bwd(rows, rows, d, (n, 1, 0, rs), 0.5, stats[1], stats[1], one, one, 0, 1.0 / q_rows, v0, v1, stats[0], g0, g1, ws)
    else:
        # rows laid out rank-major would interleave the views; for timing the first ql rows stand in for rank 0's
        be.fwd(rows[:ql], rows, d, (nl, world, 0, rs), 0.5, stats[1][:ql], stats[2][:ql], stats[3][:ql], loss_sum, ws)
        be.profile_next(1, ev[2], ev[3])
        be.bwd(rows[:ql], rows, d, (nl, world, 0, rs), 0.5, stats[1][:ql], rsk, one, one, 0, 1.0 / q_rows, v0[:nl], v1[:nl],
               stats[0][:ql], g0[:nl], g1[:nl], ws)
    ev[5].record()
    torch.cuda.synchronize()
    if it >= 5:
        tf.append(ev[0].elapsed_time(ev[1])); tb.append(ev[2].elapsed_time(ev[3])); tt.append(ev[4].elapsed_time(ev[5]))
print(f"{os.environ.get('NTXENT_B200_LIB', 'default'):40s} fwd {statistics.mean(tf)*1e3:7.1f} us  bwd {statistics.mean(tb)*1e3:7.1f} us  "
      f"step(5 kernels) {statistics.mean(tt)*1e3:7.1f} us  loss {float(loss_sum[1]):.6f} |g0| {float(g0.float().abs().sum()):.6f}")
