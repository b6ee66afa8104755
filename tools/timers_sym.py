"""Cycle counters of the symmetric forward (-DNTX_TIMING build); forward only."""
import ctypes, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simclr_b200 import _cabi  # noqa: E402
n, d = 8192, 128
dev = torch.device("cuda:0"); be = _cabi.Backend()
g = torch.Generator().manual_seed(7)
v0 = torch.randn(n, d, generator=g).bfloat16().to(dev); v1 = torch.randn(n, d, generator=g).bfloat16().to(dev)
q_rows = 2 * n; ld = be.row_pitch(d, torch.bfloat16)
rows = torch.empty(q_rows, ld, dtype=torch.bfloat16, device=dev); stats = torch.empty(4, q_rows, device=dev)
loss_sum = torch.empty(2, device=dev); plan = be.plan(q_rows, q_rows, d, torch.bfloat16)
ws = torch.empty(max(plan.fwd_ws_bytes, plan.bwd_ws_bytes), dtype=torch.uint8, device=dev)
be.l2norm(v0, v1, rows, stats[0])
for _ in range(3):
    be.fwd(rows, rows, d, (n, 1, 0), 0.5, stats[1], stats[2], stats[3], loss_sum, ws)
torch.cuda.synchronize()
buf = np.zeros((148, 16), dtype=np.int64)
assert _cabi.load().ntx_debug_timers(buf.ctypes.data_as(ctypes.c_void_p)) == 0
m = buf.mean(axis=0); tiles = plan.row_blocks * (plan.row_blocks + 1) / 2 / plan.fwd_grid
names = ["gemm1 total", "gemm1 wait k_full", "gemm1 wait s_free", "gemm1 issue", "epi total", "epi wait s_full",
         "epi wait e_free", "epi compute loop (own)", "epi colsum store (own)"]
print(f"tiles per CTA {tiles:.1f}, symmetric={plan.fwd_symmetric}")
for name, v in zip(names, m):
    print(f"{name:26s} {v:10.0f} cycles  {v / tiles:8.1f} per tile")
