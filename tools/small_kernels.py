"""Back-to-back timing of each library call (warm L2, CUDA events around 200 launches)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simclr_b200 import _cabi  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
d = int(sys.argv[2]) if len(sys.argv) > 2 else 128
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
plan = be.plan(q_rows, q_rows, d, torch.bfloat16)
ws = torch.empty(max(plan.fwd_ws_bytes, plan.bwd_ws_bytes), dtype=torch.uint8, device=dev)
g0, g1 = torch.empty_like(v0), torch.empty_like(v1)
one = torch.ones(1, device=dev)
lay = (n, 1, 0)


def timed(name, fn, reps=200):
    for _ in range(5):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    print(f"{name:28s} {a.elapsed_time(b) / reps * 1e3:8.2f} us per call")
    return a.elapsed_time(b) / reps * 1e3


t_l2 = timed("l2norm", lambda: be.l2norm(v0, v1, rows, stats[0]))
t_f = timed("fwd (main + finalize)", lambda: be.fwd(rows, rows, d, lay, 0.5, stats[1], stats[2], stats[3], loss_sum, ws))
t_b = timed("bwd (main + finalize)", lambda: be.bwd(rows, rows, d, lay, 0.5, stats[1], stats[1], one, one, 0, 1.0 / q_rows,
                                                    v0, v1, stats[0], g0, g1, ws))
print(f"sum {t_l2 + t_f + t_b:.1f} us")
