#!/usr/bin/env python
"""Per-rank kernel timeline of one NT-Xent step (forward + backward of the module, CUDA-graph replays).

    python tools/step_timeline.py [--out profiles/x.csv]                                   # one GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/step_timeline.py --out ...

Every rank replays the captured step `--steps` times under the CUPTI activity tracer (torch.profiler); rank 0
writes one CSV line per kernel of the step: its start relative to the step's first kernel and its duration,
both as medians over the replays, plus the gap to the previous kernel's end.  The L2 is NOT flushed between
steps here and the tracer adds a little per-kernel overhead: the numbers place the microseconds, bench.py is
the measurement.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def short_name(full: str) -> str:
    """'void (anonymous namespace)::tc_bwd_kernel<2>(CUtensorMap, ...)' -> 'tc_bwd_kernel<2>'"""
    name = full.replace("(anonymous namespace)::", "")
    if name.startswith("void "):
        name = name[5:]
    depth, out = 0, []
    for ch in name:                      # cut at the first '(' outside template brackets
        if ch == "<":
            depth += 1
        elif ch == ">":
            depth -= 1
        elif ch == "(" and depth == 0:
            break
        out.append(ch)
    return "".join(out)[:90].replace(",", ";")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=8192, help="global batch (pairs)")
    ap.add_argument("--d", type=int, default=128)
    ap.add_argument("--tau", type=float, default=0.5)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--gather-backend", default="peer")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    from simclr_b200 import NT_Xent

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    quiet = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
    if quiet:
        quiet(False)
    nl = args.n // world
    g = torch.Generator().manual_seed(7)
    v0 = torch.randn(args.n, args.d, generator=g).bfloat16()[rank * nl:(rank + 1) * nl]
    v1 = torch.randn(args.n, args.d, generator=g).bfloat16()[rank * nl:(rank + 1) * nl]
    a = v0.to(dev).requires_grad_(True)
    b = v1.to(dev).requires_grad_(True)
    crit = NT_Xent(temperature=args.tau, device=local_rank, gather=world > 1, gather_backend=args.gather_backend)

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
        step()
    for _ in range(5):
        graph.replay()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()

    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(args.steps):
            graph.replay()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()              # every rank starts a step together, as in a training loop
        torch.cuda.synchronize()
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "trace.json")
        prof.export_chrome_trace(path)
        trace = json.load(open(path))
    kernels = sorted((e for e in trace["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")),
                     key=lambda e: e["ts"])
    # the NCCL barrier's kernels separate the steps: split on them
    steps, cur = [], []
    for e in kernels:
        if "nccl" in e["name"].lower() and args.gather_backend != "nccl":
            if cur:
                steps.append(cur)
                cur = []
            continue
        cur.append(e)
    if cur:
        steps.append(cur)
    if world == 1:
        per = len(kernels) // args.steps
        steps = [kernels[i * per:(i + 1) * per] for i in range(args.steps)]
    steps = [s for s in steps if len(s) == len(steps[len(steps) // 2])]
    if rank == 0:
        lines = ["idx,kernel,start_us,duration_us,gap_before_us"]
        nk = len(steps[0])
        total = []
        for i in range(nk):
            name = short_name(steps[0][i]["name"])
            start = statistics.median(s[i]["ts"] - s[0]["ts"] for s in steps)
            dur = statistics.median(s[i]["dur"] for s in steps)
            gap = statistics.median((s[i]["ts"] - (s[i - 1]["ts"] + s[i - 1]["dur"])) if i else 0.0 for s in steps)
            lines.append(f"{i},{name},{start:.2f},{dur:.2f},{gap:.2f}")
        total = statistics.median(s[-1]["ts"] + s[-1]["dur"] - s[0]["ts"] for s in steps)
        lines.append(f"# rank 0 of {world}; global batch {args.n}, d={args.d}; {len(steps)} replays; "
                     f"first kernel start -> last kernel end: {total:.2f} us (median)")
        text = "\n".join(lines)
        print(text)
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            open(args.out, "w").write(text + "\n")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
