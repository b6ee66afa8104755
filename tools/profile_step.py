"""A few fwd+bwd steps of the benchmark workload straight through the C ABI (for ncu).

    python tools/profile_step.py [n_pairs] [d] [steps] [world]
    NTX_ROWS=x2 python tools/profile_step.py ...      # fp32 inputs held as two bf16 terms (NTX_BF16X2 rows)

world > 1: the kernels of ONE rank of a `world`-rank split of the same global batch (its 2n/world rows against all
2n keys, the shapes a rank of `bench.py --gpus world` launches), emulated on this one GPU without any exchange:
every rank's forward runs first (they supply the row sums), then `steps` forward + backward passes of rank 0.
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simclr_b200 import _cabi  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
d = int(sys.argv[2]) if len(sys.argv) > 2 else 128
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
world = int(sys.argv[4]) if len(sys.argv) > 4 else 1
dev = torch.device("cuda:0")
be = _cabi.Backend()
g = torch.Generator().manual_seed(7)
split = os.environ.get("NTX_ROWS", "") == "x2"
kind = _cabi.BF16X2 if split else torch.bfloat16
kw = {"rows_dtype": kind} if split else {}
v0 = torch.randn(n, d, generator=g).to(torch.float32 if split else torch.bfloat16).to(dev)
v1 = torch.randn(n, d, generator=g).to(torch.float32 if split else torch.bfloat16).to(dev)
nl = n // world
q_rows = 2 * nl
ld = be.row_pitch(d, kind)
keys = torch.empty(world * q_rows, ld, dtype=torch.bfloat16, device=dev)
stats = torch.empty(world, 4, q_rows, device=dev)
loss_sum = torch.empty(2, device=dev)
plan = be.plan(q_rows, world * q_rows, d, kind)
ws = torch.empty(max(plan.fwd_ws_bytes, plan.bwd_ws_bytes), dtype=torch.uint8, device=dev)
one = torch.ones(1, device=dev)
rs = be.row_scale(0.5, kind)                        # the module's bf16 rows carry sqrt(log2 e / tau)
views = [(v0[k * nl:(k + 1) * nl], v1[k * nl:(k + 1) * nl]) for k in range(world)]
g0, g1 = torch.empty_like(views[0][0]), torch.empty_like(views[0][1])


def forward(k):
    q = keys[k * q_rows:(k + 1) * q_rows]
    be.l2norm(views[k][0], views[k][1], q, stats[k, 0], row_scale=rs, **kw)
    be.fwd(q, keys, d, (nl, world, k, rs), 0.5, stats[k, 1], stats[k, 2], stats[k, 3], loss_sum, ws, **kw)


for k in range(world - 1, 0, -1):                 # every rank's rows first, then their forward passes
    be.l2norm(views[k][0], views[k][1], keys[k * q_rows:(k + 1) * q_rows], stats[k, 0], row_scale=rs, **kw)
be.l2norm(views[0][0], views[0][1], keys[:q_rows], stats[0, 0], row_scale=rs, **kw)
for k in range(world - 1, 0, -1):
    forward(k)
for _ in range(steps):
    forward(0)
    rowsum_k = stats[:, 1, :].reshape(-1).contiguous() if world > 1 else stats[0, 1]
    be.bwd(keys[:q_rows], keys, d, (nl, world, 0, rs), 0.5, stats[0, 1], rowsum_k, one, one, 0, 1.0 / q_rows,
           views[0][0], views[0][1], stats[0, 0], g0, g1, ws, **kw)
torch.cuda.synchronize()
print("loss", float(loss_sum[1]))
