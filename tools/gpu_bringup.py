"""Bring-up diagnostics on a real B200: stage-by-stage errors against the fp64 oracle.

    python tools/gpu_bringup.py [fp32|bf16|all]

Each mode runs in its own subprocess (a device-side trap poisons the CUDA context).
Test infrastructure: imports oracle/.
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def stage(mode, n, d, tau, world=1):
    import numpy as np
    import torch
    from oracle import ntxent_oracle as orc
    from simclr_b200 import _cabi

    be = _cabi.Backend()
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(n * 7 + d)
    v0 = torch.randn(n, d, generator=g)
    v1 = torch.randn(n, d, generator=g)
    in_dtype = torch.float32 if mode == "fp32" else torch.bfloat16
    rows_dtype = in_dtype
    v0 = v0.to(in_dtype)
    v1 = v1.to(in_dtype)
    ref = orc.ntxent_rows_fp64(v0.float().numpy(), v1.float().numpy(), tau,
                               weights=orc.weights_for("mean", n))
    a, b = v0.to(dev), v1.to(dev)
    q_rows = 2 * n
    ld = be.row_pitch(d, rows_dtype)
    rows = torch.empty(q_rows, ld, dtype=rows_dtype, device=dev)
    stats = torch.zeros(4, q_rows, device=dev)
    loss_sum = torch.zeros(2, device=dev)
    be.l2norm(a, b, rows, stats[0])
    torch.cuda.synchronize()
    e_rows = np.abs(rows[:, :d].float().cpu().numpy() - ref["zhat"]).max()
    e_pad = float(rows[:, d:].float().abs().max()) if ld > d else 0.0
    plan = be.plan(q_rows, q_rows, d, rows_dtype)
    ws = torch.empty(max(plan.fwd_ws_bytes, plan.bwd_ws_bytes, 16), dtype=torch.uint8, device=dev)
    be.fwd(rows, rows, d, (n, 1, 0), tau, stats[1], stats[2], stats[3], loss_sum, ws)
    torch.cuda.synchronize()
    lse = (torch.log(stats[1]) + 1.0 / tau).cpu().numpy()
    e_lse = np.abs(lse - ref["lse"]).max()
    e_pos = np.abs(stats[2].cpu().numpy() - ref["pos"]).max()
    e_loss = abs(float(loss_sum[1]) - ref["loss"].mean()) / abs(ref["loss"].mean())
    gout = torch.ones(1, device=dev)
    d0 = torch.empty_like(a)
    d1 = torch.empty_like(b)
    be.bwd(rows, rows, d, (n, 1, 0), tau, stats[1], stats[1], gout, gout, 0, 1.0 / q_rows, a, b, stats[0],
           d0, d1, ws)
    torch.cuda.synchronize()
    scale = max(np.abs(ref["dview0"]).max(), np.abs(ref["dview1"]).max())
    e_g0 = np.abs(d0.float().cpu().numpy() - ref["dview0"]).max() / scale
    e_g1 = np.abs(d1.float().cpu().numpy() - ref["dview1"]).max() / scale
    print(f"[{mode}] N={n} D={d} tau={tau} grid={plan.grid} slots={plan.max_slots} tc={plan.tensor_cores} | "
          f"rows {e_rows:.2e} pad {e_pad:.1e} lse {e_lse:.2e} pos {e_pos:.2e} loss_rel {e_loss:.2e} "
          f"g0 {e_g0:.2e} g1 {e_g1:.2e}", flush=True)


def child(mode):
    import torch
    print(torch.cuda.get_device_name(0), flush=True)
    sizes = [(8, 16, 0.5), (64, 128, 0.5), (33, 20, 0.5), (256, 128, 0.5), (300, 96, 0.1), (1024, 128, 0.5),
             (777, 256, 0.2), (2048, 64, 0.5)]
    for n, d, tau in sizes:
        try:
            stage(mode, n, d, tau)
        except Exception as e:  # noqa: BLE001
            print(f"[{mode}] N={n} D={d}: FAILED {type(e).__name__}: {e}", flush=True)
            if "CUDA" in str(e) or "cuda" in str(e):
                return 1
    return 0


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        sys.exit(child(sys.argv[2]))
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    rc = 0
    for mode in (("fp32", "bf16") if which == "all" else (which,)):
        p = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", mode], timeout=600)
        print(f"== {mode} exit {p.returncode}", flush=True)
        rc |= p.returncode
    sys.exit(rc)
