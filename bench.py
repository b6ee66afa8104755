#!/usr/bin/env python
"""NT-Xent forward+backward throughput on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N ...            # the reference algorithm on host CPU cores

Workload (BASELINE.json configs[3], the one the metric and target are quoted on):
global batch N = 8192 pairs, D = 128, bf16 embeddings, tau = 0.5, synthetic N(0,1)
inputs (seed 7).  With N GPUs the global batch is sharded by rows (strong scaling):
each rank owns N/W pairs and all-gathers the normalised keys.

A "step" = one forward + backward of the loss module (loss value + dview0, dview1).
value   : pairs/s with inputs resident in HBM (CUDA events per step, L2 flushed between steps).
e2e     : pairs/s through the public module from PINNED HOST buffers, H2D copies and the
          D2H read of the loss inside the timed region; every step copies its own 2*N*D inputs, the
          copy of step k+1 running on a copy stream while step k computes (serial figure also given).
roofline: dominant kernel (tcgen05 backward), algorithmic flops 12*N^2*D / W per launch.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_GLOBAL, DIM, TAU, SEED = 8192, 128, 0.5, 7
METRIC, UNIT = "ntxent_fwd_bwd_pairs_per_sec", "pairs/s"
FALLBACK_TFLOPS = 1590.0          # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent
KERNELS_PER_STEP = 5              # l2norm, fwd, fwd_finalize, bwd, bwd_finalize (+ the scatter kernel with peer gather)
# dram__bytes_read.sum + dram__bytes_write.sum of ONE tc_bwd_kernel<2> launch at config 4 on one GPU, from the
# committed `ncu --set full` capture (profiles/r01b_all_kernels_ncu_full_summary.csv): the keys once, nothing else
BWD_TRAFFIC_BYTES_C4_1GPU = 4328192


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="c4", choices=["c4", "c5"],
                    help="c4 = BASELINE configs[3] (the metric's config, default); c5 = configs[4]: "
                         "global batch 32768, d=256 (documentation runs only)")
    ap.add_argument("--gather-backend", default="peer", choices=["peer", "nccl"],
                    help="multi-GPU exchange: kernels store into every rank's buffers over NVLink peer memory "
                         "(falls back to NCCL all-gathers when symmetric memory cannot be set up), or NCCL")
    ap.add_argument("--no-graphs", action="store_true",
                    help="launch every step eagerly instead of replaying CUDA graphs of the module")
    return ap.parse_args()


GATHER_BACKEND = "nccl"


def workload_config(world, eager=False):
    return {"workload": f"NT-Xent fwd+bwd, global batch {N_GLOBAL} pairs ({2 * N_GLOBAL} views), d={DIM}, "
                        f"tau={TAU}, bf16 (BASELINE configs[{3 if N_GLOBAL == 8192 else 4}])",
            "global_batch": N_GLOBAL, "d": DIM, "temperature": TAU,
            "parallelism": (f"rows sharded over {world} rank(s), keys gathered by "
                            + ("peer-memory stores from the normalise kernel + barrier kernel (NVLink)"
                               if GATHER_BACKEND == "peer" else "NCCL all-gather")) if world > 1 else "single GPU",
            "l2": "flushed between timed steps (256 MiB write)",
            "launch": "eager" if eager else "one CUDA graph of the module's forward + backward "
                                            "(kernels and cross-rank gathers), replayed per step"}


def synthetic(dtype):
    import torch
    g = torch.Generator().manual_seed(SEED)
    v0 = torch.randn(N_GLOBAL, DIM, generator=g, dtype=torch.float32)
    v1 = torch.randn(N_GLOBAL, DIM, generator=g, dtype=torch.float32)
    return v0.to(dtype), v1.to(dtype)


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock and throttle reasons polled through NVML every ~2 ms while the timed regions run
    (the regions last tens of milliseconds: `nvidia-smi -lms` is too coarse for them)."""

    def __init__(self, index):
        self.index, self.samples, self.stop, self.thread, self.max_mhz = index, [], False, None, None

    def __enter__(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            uuid = None
            try:
                import torch
                uuid = str(torch.cuda.get_device_properties(self.index).uuid)
            except Exception:  # noqa: BLE001
                pass
            handle = None
            if uuid:
                try:
                    handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
                except Exception:  # noqa: BLE001
                    handle = None
            if handle is None:
                handle = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(handle, pynvml.NVML_CLOCK_SM))

            def poll():
                while not self.stop:
                    try:
                        mhz = pynvml.nvmlDeviceGetClockInfo(handle, pynvml.NVML_CLOCK_SM)
                        reasons = pynvml.nvmlDeviceGetCurrentClocksEventReasons(handle)
                        util = pynvml.nvmlDeviceGetUtilizationRates(handle).gpu
                        self.samples.append((float(mhz), int(reasons), int(util)))
                    except Exception:  # noqa: BLE001
                        pass
                    time.sleep(0.002)
            self.thread = threading.Thread(target=poll, daemon=True)
            self.thread.start()
        except Exception:  # noqa: BLE001
            self.thread = None
        return self

    def __exit__(self, *exc):
        self.stop = True
        if self.thread:
            self.thread.join(timeout=2)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        bits = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                "sw_power_cap": 0x4, "hw_power_brake_slowdown": 0x80}
        seen = set()
        for _, r, _ in self.samples:
            for name, bit in bits.items():
                if r & bit:
                    seen.add(name)
        return {"sm_mhz": statistics.median(m for m, _, _ in self.samples), "sm_max_mhz": self.max_mhz,
                "sm_min_mhz": min(m for m, _, _ in self.samples), "reasons": sorted(seen),
                "samples": len(self.samples), "how": "NVML polled every 2 ms across the timed regions"}


# ----------------------------------------------------------------------------- CPU arm
def cpu_steps(n, steps, warmup, threads=None):
    """fwd+bwd of the reference-form port (oracle) on host cores; returns per-step seconds."""
    import torch
    from oracle import ntxent_oracle as orc
    if threads:
        torch.set_num_threads(threads)
    v0, v1 = synthetic(torch.bfloat16)                       # same bf16-rounded values the GPU sees,
    v0 = v0[:n].float().requires_grad_(True)                 # computed in fp32 like the reference would
    v1 = v1[:n].float().requires_grad_(True)
    times = []
    for i in range(warmup + steps):
        v0.grad = v1.grad = None
        t0 = time.perf_counter()
        orc.ntxent_reference_form(v0, v1, TAU, "mean").backward()
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    return times


def cpu_baseline_block():
    import torch
    cores = torch.get_num_threads()
    t = cpu_steps(N_GLOBAL, steps=2, warmup=1)
    sec = statistics.median(t)
    return {"value": N_GLOBAL / sec, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"2 timed + 1 warm-up fwd+bwd steps of oracle.ntxent_reference_form at the full "
                      f"N={N_GLOBAL}, D={DIM} (fp32 arithmetic on the bf16-rounded inputs), "
                      f"{sec * 1e3:.0f} ms/step"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import torch
    cores = torch.get_num_threads()
    probe = cpu_steps(N_GLOBAL, steps=1, warmup=0)[0]       # also serves as a first warm-up
    budget = 240.0
    total = args.steps + args.warmup
    n = N_GLOBAL
    while n > 512 and probe * (n / N_GLOBAL) ** 2 * total > budget:
        n //= 2
    times = cpu_steps(n, args.steps, args.warmup)
    sec = statistics.mean(times)
    scale = (N_GLOBAL / n) ** 2                              # work is quadratic in the batch
    sec_full = sec * scale
    value = N_GLOBAL / sec_full
    sample = (f"each step = fwd+bwd of the reference-form port on {n} of {N_GLOBAL} pairs"
              + ("" if n == N_GLOBAL else f", time scaled by (N/n)^2 = {scale:.0f} to the full batch"))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec_full * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "fp32",
            "data": "synthetic", "config": workload_config(1),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from simclr_b200 import NT_Xent, _cabi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the CUDA library is the only implementation")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if N_GLOBAL % world:
        raise SystemExit(f"global batch {N_GLOBAL} is not divisible by {world} ranks")
    _cabi.load()
    n_local = N_GLOBAL // world
    K, W = args.steps, max(args.warmup, 3)

    v0_all, v1_all = synthetic(torch.bfloat16)
    sl = slice(rank * n_local, (rank + 1) * n_local)
    host0, host1 = v0_all[sl].contiguous().pin_memory(), v1_all[sl].contiguous().pin_memory()
    d0 = host0.to(dev).requires_grad_(True)
    d1 = host1.to(dev).requires_grad_(True)
    global GATHER_BACKEND
    GATHER_BACKEND = args.gather_backend if world > 1 else "nccl"
    if GATHER_BACKEND == "peer":
        # every rank must take the same branch: agree on whether symmetric memory can be set up
        from simclr_b200 import peer_exchange
        ok = torch.tensor([1 if peer_exchange.available() and world <= peer_exchange.MAX_PEERS else 0], device=dev)
        if ok.item():
            try:
                ld = _cabi.Backend().row_pitch(DIM, torch.bfloat16)
                peer_exchange.exchange_for(None, 2 * n_local, ld, torch.bfloat16, dev)
            except Exception as e:  # noqa: BLE001
                print(f"[bench] rank {rank}: peer-memory gather unavailable ({e!r}); using NCCL", file=sys.stderr)
                ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if not ok.item():
            GATHER_BACKEND = "nccl"
    crit = NT_Xent(temperature=TAU, reduction="mean", device=local_rank, gather=world > 1,
                   gather_backend=GATHER_BACKEND)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ctl = dist.new_group(backend="gloo") if world > 1 else None     # control plane: barriers, max over ranks

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(group=ctl)

    clocks = ClockSampler(local_rank)
    clocks.__enter__()                                               # sampled over every timed region below

    # ---- dominant kernel alone (tcgen05 backward) through the C ABI, CUDA events on its stream
    be = _cabi.Backend()
    q_rows = 2 * n_local
    ld = be.row_pitch(DIM, torch.bfloat16)
    with torch.no_grad():
        rows = torch.empty(q_rows, ld, dtype=torch.bfloat16, device=dev)
        stats = torch.empty(4, q_rows, device=dev)
        be.l2norm(d0, d1, rows, stats[0])
        keys = rows
        if world > 1:
            keys = torch.empty(world * q_rows, ld, dtype=torch.bfloat16, device=dev)
            dist.all_gather_into_tensor(keys, rows)
        q = keys[rank * q_rows:(rank + 1) * q_rows]
        plan = be.plan(q_rows, keys.shape[0], DIM, torch.bfloat16)
        ws = torch.empty(max(plan.fwd_ws_bytes, plan.bwd_ws_bytes), dtype=torch.uint8, device=dev)
        loss_sum = torch.empty(2, device=dev)
        lay = (n_local, world, rank)
        g0, g1 = torch.empty_like(d0), torch.empty_like(d1)
        one = torch.ones(1, device=dev)
        kt = {0: [], 1: []}
        for i in range(W + K):
            evs = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            for e in evs:
                e.record()
            flush.zero_()
            be.profile_next(0, evs[0], evs[1])
            be.fwd(q, keys, DIM, lay, TAU, stats[1], stats[2], stats[3], loss_sum, ws)
            rowsum_k = stats[1]
            if world > 1:
                rowsum_k = torch.empty(world * q_rows, device=dev)
                dist.all_gather_into_tensor(rowsum_k, stats[1].contiguous())
            be.profile_next(1, evs[2], evs[3])
            be.bwd(q, keys, DIM, lay, TAU, stats[1], rowsum_k, one, one, 0, 1.0 / q_rows, d0, d1, stats[0],
                   g0, g1, ws)
            torch.cuda.synchronize()
            if i >= W:
                kt[0].append(evs[0].elapsed_time(evs[1]))
                kt[1].append(evs[2].elapsed_time(evs[3]))
    fwd_ms, bwd_ms = statistics.mean(kt[0]), statistics.mean(kt[1])


    # ---- the step: forward + backward of the public module on static inputs.  By default the whole
    # step (five kernels + the NCCL gathers) is captured once into a CUDA graph and replayed.
    # Fresh leaves: they must first be used on the capture side stream, never on the legacy stream.
    sa = d0.detach().clone().requires_grad_(True)
    sb = d1.detach().clone().requires_grad_(True)

    def eager_step():
        sa.grad = sb.grad = None
        loss = crit(sa, sb)
        loss.backward()
        return loss

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(W):
            static_loss = eager_step()
    torch.cuda.current_stream().wait_stream(side)
    barrier()
    if args.no_graphs:
        def step():
            nonlocal static_loss
            static_loss = eager_step()
    else:
        graph = torch.cuda.CUDAGraph()
        sa.grad = sb.grad = None
        with torch.cuda.graph(graph, capture_error_mode="thread_local"):
            static_loss = eager_step()
        step = graph.replay
    for _ in range(W):
        step()
    barrier()

    # ---- device-resident timing: per-step CUDA events, L2 flushed between steps
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    stops = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    for i in range(K):
        flush.zero_()
        starts[i].record()
        step()
        stops[i].record()
    barrier()
    dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, stops))

    # ---- end to end: pinned host -> device inputs, module fwd+bwd, loss value back on the host
    def e2e_step():
        with torch.no_grad():
            sa.copy_(host0, non_blocking=True)
            sb.copy_(host1, non_blocking=True)
        step()
        return float(static_loss.detach())
    for _ in range(W):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        last_loss = e2e_step()
    barrier()
    e2e_serial_ms = (time.perf_counter() - t0) * 1e3

    # The same with the input copy of step k+1 overlapping step k, as a training loop's prefetcher does:
    # a copy stream fills one of two staging sets from pinned memory while the other feeds the step;
    # every step still brings its 2*N*D inputs over PCIe and its loss value is read on the host (one
    # step later, from pinned memory), all inside the timed region.
    copy_stream = torch.cuda.Stream()
    stage = [(torch.empty_like(sa.detach()), torch.empty_like(sb.detach())) for _ in range(2)]
    copied = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    finished = [torch.cuda.Event() for _ in range(2)]
    loss_host = [torch.zeros((), dtype=static_loss.dtype).pin_memory() for _ in range(2)]
    main = torch.cuda.current_stream()

    def prefetch(k):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[k & 1])
            stage[k & 1][0].copy_(host0, non_blocking=True)
            stage[k & 1][1].copy_(host1, non_blocking=True)
            copied[k & 1].record(copy_stream)

    def pipelined_steps(count):
        losses = []
        for ev in consumed:
            ev.record(main)
        prefetch(0)
        for k in range(count):
            if k + 1 < count:
                prefetch(k + 1)
            main.wait_event(copied[k & 1])
            with torch.no_grad():
                sa.copy_(stage[k & 1][0], non_blocking=True)
                sb.copy_(stage[k & 1][1], non_blocking=True)
            consumed[k & 1].record(main)
            step()
            loss_host[k & 1].copy_(static_loss.detach(), non_blocking=True)
            finished[k & 1].record(main)
            if k >= 1:
                finished[(k - 1) & 1].synchronize()
                losses.append(float(loss_host[(k - 1) & 1]))
        finished[(count - 1) & 1].synchronize()
        losses.append(float(loss_host[(count - 1) & 1]))
        return losses

    pipelined_steps(W)
    barrier()
    t0 = time.perf_counter()
    e2e_losses = pipelined_steps(K)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    assert len(e2e_losses) == K and abs(e2e_losses[-1] - last_loss) <= 1e-2 * abs(last_loss), (e2e_losses[-1], last_loss)
    clocks.__exit__(None, None, None)

    if world > 1:
        t = torch.tensor([dev_ms, e2e_ms, fwd_ms, bwd_ms, e2e_serial_ms], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=ctl)            # max over ranks
        dev_ms, e2e_ms, fwd_ms, bwd_ms, e2e_serial_ms = t.tolist()

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        peak_sustained = None
        if os.path.exists(peaks_path):
            peaks = json.load(open(peaks_path))
            peak, peak_src = peaks["bf16_tflops"], "measured (MEASURED_PEAKS.json bf16_tflops, burst)"
            peak_sustained = peaks.get("bf16_tflops_sustained")
        else:
            peak, peak_src = FALLBACK_TFLOPS, "fallback (B200_PROFILING.md)"
        # what the kernel actually issues: S is recomputed, and both GEMMs run over all (2N)^2 logits
        bwd_executed = 16.0 * N_GLOBAL * N_GLOBAL * DIM / world
        bwd_flops = 12.0 * N_GLOBAL * N_GLOBAL * DIM / world      # algorithmic, SURVEY.md 8a row a8
        fwd_flops = 6.0 * N_GLOBAL * N_GLOBAL * DIM / world
        achieved = bwd_flops / (bwd_ms * 1e-3) / 1e12
        ms_per_step = dev_ms / K
        line = {
            "metric": METRIC, "value": N_GLOBAL / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(world, args.no_graphs),
            "e2e": {"value": N_GLOBAL / (e2e_ms / K * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": 2 * N_GLOBAL * DIM * 2, "d2h_bytes_per_step": 2 * world,
                    "ms_per_step": e2e_ms / K, "loss": last_loss,
                    "how": "every step: inputs pinned host -> device (double-buffered on a copy stream, so the "
                           "copy of step k+1 overlaps step k), module forward + backward, loss read on the host",
                    "serial_ms_per_step": e2e_serial_ms / K,
                    "serial_value": N_GLOBAL / (e2e_serial_ms / K * 1e-3)},
            "gpu_launches": (KERNELS_PER_STEP + (1 if GATHER_BACKEND == "peer" else 0)) * K,
            "roofline": {"bound": "tensor", "kernel": f"tc_bwd_kernel<{DIM // 64}>", "achieved": achieved, "peak": peak,
                         "unit": "TFLOP/s", "frac": achieved / peak,
                         "traffic": BWD_TRAFFIC_BYTES_C4_1GPU if (world == 1 and N_GLOBAL == 8192 and DIM == 128) else None,
                         "traffic_source": "ncu --set full, profiles/r01b_all_kernels_ncu_full_summary.csv (bytes per launch)",
                         "peak_source": peak_src, "algorithmic_flops_per_launch": bwd_flops,
                         "kernel_ms": bwd_ms,
                         "executed_flops_per_launch": bwd_executed,
                         "executed_tflops": bwd_executed / (bwd_ms * 1e-3) / 1e12,
                         "peak_sustained": peak_sustained,
                         "executed_frac_of_sustained": (bwd_executed / (bwd_ms * 1e-3) / 1e12 / peak_sustained
                                                        if peak_sustained else None),
                         "fwd_kernel": {"kernel": ("tc_fwd_sym_kernel" if world == 1 and DIM <= 128 else "tc_fwd_kernel") + f"<{DIM // 64}>", "kernel_ms": fwd_ms,
                                        "achieved": fwd_flops / (fwd_ms * 1e-3) / 1e12},
                         "whole_step_frac": 18.0 * N_GLOBAL * N_GLOBAL * DIM / world / (ms_per_step * 1e-3) / 1e12 / peak},
            "clocks": clocks.summary(),
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_block()
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier(group=ctl)
        sys.stdout.flush()
        os._exit(0)          # skip NCCL teardown: nothing left to do and it can stall after graph replays
    return 0


if __name__ == "__main__":
    a = parse()
    if a.workload == "c5":
        N_GLOBAL, DIM = 32768, 256
    sys.exit(run_reference(a) if a.impl == "reference" else run_ours(a))
