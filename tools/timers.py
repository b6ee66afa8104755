"""Read the per-CTA cycle counters of the -DNTX_TIMING build after one fwd+bwd step.

    NTXENT_B200_LIB=simclr_b200/libntxent_b200_timing.so python tools/timers.py [n] [d] [world]

world > 1: the shapes of one rank of `world` (its 2n/world rows against all 2n keys), run on this one GPU.

NTX_TIMELINE=1 also prints per-CTA stamps (cycles from kernel entry) of a build with both defines:
    python -c "import os, simclr_b200.build as b; b.build(force=True, defines=('NTX_TIMING', 'NTX_TIMELINE'), \
               out=os.path.join(b.PKG, 'libntxent_b200_timeline.so'))"
"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simclr_b200 import _cabi  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
d = int(sys.argv[2]) if len(sys.argv) > 2 else 128
world = int(sys.argv[3]) if len(sys.argv) > 3 else 1
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
nl = n // world
ql = 2 * nl
plan = be.plan(ql, q_rows, d, torch.bfloat16)
ws = torch.empty(max(plan.fwd_ws_bytes, plan.bwd_ws_bytes), dtype=torch.uint8, device=dev)
g0, g1 = torch.empty_like(v0), torch.empty_like(v1)
one = torch.ones(1, device=dev)
rs = be.row_scale(0.5, torch.bfloat16)
rsk = torch.full((q_rows,), 5000.0, device=dev)          # stand-in row sums of every key (world > 1)
for _ in range(3):
    be.l2norm(v0, v1, rows, stats[0], row_scale=rs)
    if world == 1:
        be.fwd(rows, rows, d, (n, 1, 0, rs), 0.5, stats[1], stats[2], stats[3], loss_sum, ws)
        be.bwd(rows, rows, d, (n, 1, 0, rs), 0.5, stats[1], stats[1], one, one, 0, 1.0 / q_rows, v0, v1, stats[0], g0, g1, ws)
    else:
        be.fwd(rows[:ql], rows, d, (nl, world, 0, rs), 0.5, stats[1][:ql], stats[2][:ql], stats[3][:ql], loss_sum, ws)
        be.bwd(rows[:ql], rows, d, (nl, world, 0, rs), 0.5, stats[1][:ql], rsk, one, one, 0, 1.0 / q_rows, v0[:nl], v1[:nl],
               stats[0][:ql], g0[:nl], g1[:nl], ws)
torch.cuda.synchronize()
buf = np.zeros((148, 24), dtype=np.int64)
rc = _cabi.load().ntx_debug_timers(buf.ctypes.data_as(ctypes.c_void_p))
assert rc == 0
m = buf.mean(axis=0)
tiles = plan.row_blocks * plan.key_tiles / plan.grid
names = ["fwd mma total", "fwd mma wait k_full", "fwd mma wait s_empty", "fwd mma issue", "fwd epi total",
         "fwd epi wait s_full", "", "bwd epi drain (wait + store)", "bwd mma total", "bwd mma wait k_full", "bwd mma wait buf_free",
         "bwd gemm2 wait c_full", "bwd gemm1 issue", "bwd epi total", "bwd epi wait s_full", "bwd gemm2 issue"]
print(f"tiles per CTA {tiles:.1f}")
for name, v in zip(names, m):
    if name:
        print(f"{name:24s} {v:10.0f} cycles  {v / tiles:8.1f} per tile")

if os.environ.get("NTX_TIMELINE"):
    def line(label, cols):
        used = buf[:plan.grid] if plan.grid < 148 else buf
        print(label + ": " + ", ".join(f"{name} {used[:, c].mean():.0f} / {used[:, c].max()}" for name, c in cols))
    print("cycles from kernel entry, mean / max over CTAs (full-tile forward only; -DNTX_TIMING -DNTX_TIMELINE build)")
    line("fwd", [("previous kernel done", 20), ("first S tile ready", 21), ("tiles and flush done", 22)])
    line("bwd", [("previous kernel done", 16), ("first S tile ready", 17), ("last tile of group 0 done", 18),
                 ("after the final drain", 19)])
