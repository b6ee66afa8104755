#!/usr/bin/env python
"""Time of one optimizer step on a ResNet-18 / ResNet-50 parameter set (the reference's base_cnn choices):
the fused LARC + SGD kernel (simclr_b200.larc.LARC) against the reference's recipe written with torch ops
(apex LARC.step's per-tensor norms and gradient edits + torch.optim.SGD; apex itself is not installable here).

    python tools/larc_times.py [--arch resnet18] [--iters 50]

Prints one JSON line: microseconds per step (CUDA events, graph-free, warm), the HBM rate of the fused kernel by the
algorithmic count of 3 reads + 2 writes x 4 bytes per element, and the measured copy peak beside it."""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="resnet18")
    ap.add_argument("--iters", type=int, default=50)
    args = ap.parse_args()
    import torch
    import torchvision
    from simclr_b200.larc import LARC
    dev = torch.device("cuda:0")
    torch.manual_seed(0)

    def groups(net):
        named = list(net.named_parameters())
        skip = ("bias", "bn")
        return [{"params": [p for n, p in named if not any(s in n for s in skip)], "weight_decay": 1e-4},
                {"params": [p for n, p in named if any(s in n for s in skip)], "weight_decay": 0.0}]

    def timed(fn):
        for _ in range(5):
            fn()
        times = []
        for _ in range(args.iters):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            times.append(a.elapsed_time(b) * 1e3)
        return statistics.median(times)

    net_a = getattr(torchvision.models, args.arch)(num_classes=128).to(dev)
    net_b = getattr(torchvision.models, args.arch)(num_classes=128).to(dev)
    for net in (net_a, net_b):
        for p in net.parameters():
            p.grad = torch.randn_like(p)
    numel = sum(p.numel() for p in net_a.parameters())
    fused = LARC(torch.optim.SGD(groups(net_a), lr=0.3, momentum=0.9), trust_coefficient=0.001, clip=False)
    ref = torch.optim.SGD(groups(net_b), lr=0.3, momentum=0.9)

    def ref_step():
        with torch.no_grad():
            decays = []
            for group in ref.param_groups:
                wd = group["weight_decay"]
                decays.append(wd)
                group["weight_decay"] = 0
                for p in group["params"]:
                    pn, gn = torch.norm(p.data), torch.norm(p.grad.data)
                    if pn != 0 and gn != 0:                       # two host syncs per tensor, as in apex
                        a = 0.001 * pn / (gn + pn * wd + 1e-8)
                        p.grad.data += wd * p.data
                        p.grad.data *= a
            ref.step()
            for group, wd in zip(ref.param_groups, decays):
                group["weight_decay"] = wd

    t_fused = timed(fused.step)
    t_ref = timed(ref_step)
    # the kernel alone: the step captured in a CUDA graph, so that no host time sits between the events
    t_graph = None
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fused.step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            fused.step()
        t_graph = timed(graph.replay)
    except Exception as e:  # noqa: BLE001
        print(f"[larc_times] graph capture of the cooperative launch failed: {e!r}", file=sys.stderr)
    peak = None
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        peak = json.load(open(path)).get("hbm_gbs")
    gbs = 20.0 * numel / ((t_graph or t_fused) * 1e-6) / 1e9
    print(json.dumps({"arch": args.arch, "parameters": numel, "tensors": len(list(net_a.parameters())),
                      "fused_us": t_fused, "fused_graph_replay_us": t_graph, "torch_recipe_us": t_ref,
                      "speedup": t_ref / t_fused,
                      "fused_gbs_algorithmic_20B_per_element": gbs, "hbm_peak_gbs": peak,
                      "frac": gbs / peak if peak else None}))


if __name__ == "__main__":
    main()
