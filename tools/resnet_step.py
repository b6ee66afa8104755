#!/usr/bin/env python
"""BASELINE configs[1] (SimCLR ResNet-18 CIFAR-10 step, batch 512, d = 128, tau = 0.5, one B200) with the reference's
OWN model: ``ContrastiveModel("resnet18", d=128)`` imported from the unmodified copy of model.py under the git-ignored
baseline/_ref/ (staged by ``__graft_entry__.build()``).  The call pattern of main.py:104-122 is rebuilt by hand
(hydra / apex are not installable): zero_grad, two model calls, loss, backward, optimizer step, on synthetic
``torch.rand(512, 3, 32, 32)`` views.

Two identically initialised models are stepped side by side, one with this repository's ``NT_Xent`` and one with the
reference's loss (baseline/_ref/loss.py on the GPU when present, else the oracle's reference-form port):
  * the loss values must agree step for step (fp32 inputs -> CUDA-core kernels, 1e-5);
  * the repository's loss call and its backward run under ``torch.cuda.set_sync_debug_mode("error")``: any host
    synchronisation raises (the reference's loss synchronises twice per call: mask upload, ``nonzero``);
  * loss-only time (forward + backward of the loss on detached embeddings, CUDA events) is reported for both.

    python tools/resnet_step.py [--steps 3] [--batch 512] [--fused-larc]      # prints one JSON line
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF_DIR = os.path.join(ROOT, "baseline", "_ref")


def load_ref(name):
    path = os.path.join(REF_DIR, name + ".py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("simclr_reference_" + name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def run(steps=3, batch=512, d=128, tau=0.5, lr=0.05, fused_larc=False, device="cuda:0"):
    import torch
    from simclr_b200 import NT_Xent
    dev = torch.device(device)
    model_mod = load_ref("model")
    if model_mod is None:
        raise FileNotFoundError("baseline/_ref/model.py is absent: run __graft_entry__.build() where /root/reference exists")
    loss_mod = load_ref("loss")
    torch.manual_seed(7)                                              # conf/config.yaml:9
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False
    ours_m = model_mod.ContrastiveModel("resnet18", d=d).to(dev).train()
    ref_m = model_mod.ContrastiveModel("resnet18", d=d).to(dev).train()
    ref_m.load_state_dict(ours_m.state_dict())
    ours_l = NT_Xent(temperature=tau, device=dev.index or 0)          # main.py:83
    if loss_mod is not None:
        ref_l, ref_kind = loss_mod.NT_Xent(temperature=tau, device=dev), "reference loss.py (unmodified copy)"
    else:
        from oracle import ntxent_oracle as orc
        ref_l, ref_kind = (lambda a, b: orc.ntxent_reference_form(a, b, tau, "mean")), "oracle reference-form port"

    def make_opt(model):
        sgd = torch.optim.SGD(model.parameters(), lr=lr, momentum=0.9, nesterov=False, weight_decay=0.0)
        if fused_larc:
            from simclr_b200.larc import LARC
            return LARC(optimizer=sgd, trust_coefficient=0.001, clip=False)
        return sgd

    ours_o, ref_o = make_opt(ours_m), torch.optim.SGD(ref_m.parameters(), lr=lr, momentum=0.9)
    g = torch.Generator().manual_seed(7)
    losses, t_ours, t_ref = [], [], []

    def timed_loss(crit, z0, z1):
        a, b = z0.detach().clone().requires_grad_(True), z1.detach().clone().requires_grad_(True)
        crit(a, b).backward()                                          # warm
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.grad = b.grad = None
        e0.record()
        crit(a, b).backward()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3

    for _ in range(steps):
        v0 = torch.rand(batch, 3, 32, 32, generator=g).to(dev)
        v1 = torch.rand(batch, 3, 32, 32, generator=g).to(dev)
        # ---- this repository's loss inside main.py's step; no host synchronisation allowed around the loss
        ours_o.zero_grad()
        z0, z1 = ours_m(v0), ours_m(v1)
        torch.cuda.synchronize()
        torch.cuda.set_sync_debug_mode("error")
        try:
            loss = ours_l(z0, z1)
            loss.backward()
        finally:
            torch.cuda.set_sync_debug_mode("default")
        ours_o.step()
        # ---- the reference's loss, same pattern
        ref_o.zero_grad()
        r0, r1 = ref_m(v0), ref_m(v1)
        rloss = ref_l(r0, r1)
        rloss.backward()
        ref_o.step()
        losses.append((float(loss), float(rloss)))
        t_ours.append(timed_loss(ours_l, z0, z1))
        t_ref.append(timed_loss(ref_l, r0, r1))
    return {"config": f"ContrastiveModel(resnet18, d={d}) on rand({batch},3,32,32) x2, tau={tau}, {steps} steps",
            "losses_ours_vs_reference": losses,
            "max_rel_loss_diff": max(abs(a - b) / abs(b) for a, b in losses),
            "loss_fwd_bwd_us_ours": statistics.median(t_ours), "loss_fwd_bwd_us_reference": statistics.median(t_ref),
            "reference_loss": ref_kind, "optimizer": "fused LARC + SGD (simclr_b200.larc)" if fused_larc else "torch SGD",
            "host_sync_in_our_loss": False}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--fused-larc", action="store_true")
    a = ap.parse_args()
    print(json.dumps(run(a.steps, a.batch, fused_larc=a.fused_larc)))
