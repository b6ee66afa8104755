"""Probe: whole-step CUDA graph (module forward + backward with NCCL gathers inside) under torchrun."""
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simclr_b200 import NT_Xent  # noqa: E402

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
n = 8192 // world
g = torch.Generator().manual_seed(7 + rank)
a = torch.randn(n, 128, generator=g).bfloat16().to(dev).requires_grad_(True)
b = torch.randn(n, 128, generator=g).bfloat16().to(dev).requires_grad_(True)
crit = NT_Xent(0.5, gather=True)


def step():
    loss = crit(a, b)
    loss.backward()
    return loss


for _ in range(3):
    a.grad = b.grad = None
    eager = step()
torch.cuda.synchronize()
print(rank, "eager ok", float(eager.detach()), flush=True)
# fresh leaves: AccumulateGrad nodes remember the stream they were first used on, and the eager
# run above used the legacy default stream, which may not be touched during capture
a = a.detach().clone().requires_grad_(True)
b = b.detach().clone().requires_grad_(True)
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    for _ in range(3):
        a.grad = b.grad = None
        step()
torch.cuda.current_stream().wait_stream(side)
torch.cuda.synchronize()
print(rank, "side-stream warmup ok", flush=True)
graph = torch.cuda.CUDAGraph()
a.grad = b.grad = None
with torch.cuda.graph(graph, capture_error_mode="thread_local"):
    static_loss = step()
torch.cuda.synchronize()
print(rank, "captured", flush=True)
for _ in range(5):
    graph.replay()
torch.cuda.synchronize()
dist.barrier()
t0 = time.perf_counter()
for _ in range(100):
    graph.replay()
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 100
print(rank, f"replay ok: loss {float(static_loss):.4f}, {dt * 1e6:.1f} us/step", flush=True)
dist.barrier()
dist.destroy_process_group()
