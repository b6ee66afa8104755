"""Probe: torch symmetric memory between the ranks of one node (peer pointers, barrier, graph capture).

    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/symm_probe.py
"""
import os
import sys
import time

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dev = torch.device("cuda", torch.cuda.current_device())
dist.init_process_group("nccl", device_id=dev)
group = dist.group.WORLD.group_name
buf = symm_mem.empty(world * 1024, dtype=torch.float32, device=dev)
hdl = symm_mem.rendezvous(buf, group=group)
print(f"rank {rank}: buffer_ptrs {[hex(p) for p in hdl.buffer_ptrs]} multicast_ptr {hex(hdl.multicast_ptr)} "
      f"signal_pad_size {hdl.signal_pad_size}", flush=True)
buf.fill_(-1.0)
hdl.barrier()
# every rank writes its chunk into every peer's buffer through the peer tensors
mine = torch.full((1024,), float(rank), device=dev)
for p in range(world):
    hdl.get_buffer(p, (world * 1024,), torch.float32)[rank * 1024:(rank + 1) * 1024].copy_(mine)
hdl.barrier()
want = torch.arange(world, device=dev, dtype=torch.float32).repeat_interleave(1024)
assert torch.equal(buf, want), buf[::1024]
# the same under CUDA graph capture
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
        for p in range(world):
            hdl.get_buffer(p, (world * 1024,), torch.float32)[rank * 1024:(rank + 1) * 1024].copy_(mine + 10)
        hdl.barrier()
    for _ in range(3):
        g.replay()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(20):
        g.replay()
    ev1.record()
torch.cuda.synchronize()
assert torch.equal(buf, want + 10)
print(f"rank {rank}: ok; graph of {world} peer copies + barrier: {ev0.elapsed_time(ev1) / 20 * 1e3:.1f} us", flush=True)
# barrier alone
with torch.cuda.stream(side):
    ev0.record()
    for _ in range(50):
        hdl.barrier()
    ev1.record()
torch.cuda.synchronize()
print(f"rank {rank}: barrier alone {ev0.elapsed_time(ev1) / 50 * 1e3:.1f} us (eager launches)", flush=True)
# NCCL all-gather of the same size for comparison
out = torch.empty(world * 1024, device=dev)
for _ in range(5):
    dist.all_gather_into_tensor(out, mine)
torch.cuda.synchronize()
ev0.record()
for _ in range(50):
    dist.all_gather_into_tensor(out, mine)
ev1.record()
torch.cuda.synchronize()
print(f"rank {rank}: NCCL all_gather 4 KB/rank {ev0.elapsed_time(ev1) / 50 * 1e3:.1f} us (eager launches)", flush=True)
dist.barrier()
os._exit(0)
