#!/usr/bin/env python
"""What the two backward variants of SURVEY.md section 8e would pay for their exchange step, measured with NCCL on
this box (BASELINE configs[3]: N = 8192 pairs, D = 128):

  north-star variant : reduce-scatter (SUM) of the key gradients, fp32 (2N x D per rank in, 2N/W x D out)
  symmetric variant  : all-gather of the 2N fp32 row sums (what this repository builds; its backward needs no dK GEMM)
  forward (both)     : all-gather of the normalised keys, bf16 (2N x D in total)

    python -m torch.distributed.run --nproc-per-node W --master-addr 127.0.0.1 tools/exchange_costs.py

Rank 0 prints one JSON line with the time per collective (50 back to back between two CUDA events, max over ranks)."""
import json
import os
import sys

import torch
import torch.distributed as dist

N, D = 8192, 128


def main():
    world = int(os.environ["WORLD_SIZE"])
    rank = int(os.environ["RANK"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    m = 2 * N
    rows = torch.randn(m // world, D, device=dev).bfloat16()
    keys = torch.empty(m, D, dtype=torch.bfloat16, device=dev)
    dk = torch.randn(m, D, device=dev)
    dk_mine = torch.empty(m // world, D, device=dev)
    rs = torch.randn(m // world, device=dev)
    rs_all = torch.empty(m, device=dev)
    ops = {
        "all_gather_keys_bf16": lambda: dist.all_gather_into_tensor(keys, rows),
        "reduce_scatter_dK_fp32": lambda: dist.reduce_scatter_tensor(dk_mine, dk, op=dist.ReduceOp.SUM),
        "all_gather_rowsum_fp32": lambda: dist.all_gather_into_tensor(rs_all, rs),
    }
    out = {"world": world, "bytes": {"all_gather_keys_bf16": keys.numel() * 2, "reduce_scatter_dK_fp32": dk.numel() * 4,
                                     "all_gather_rowsum_fp32": rs_all.numel() * 4}}
    for name, fn in ops.items():
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        # 50 back-to-back collectives between two events: the GPU-side cost per collective once the launch queue
        # is full (host launch time hidden behind the previous collective)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(50):
            fn()
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) * 1e3 / 50], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[name + "_us"] = round(float(t), 2)
    if rank == 0:
        print(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    sys.exit(main())
