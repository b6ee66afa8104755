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
parity  : before any timing every rank checks what it computes against the fp64 oracle on the same seed-7
          global batch: its row sums (C ABI), and -- through the module, the chosen exchange back end and the
          CAPTURED graph, over three replays with two different input sets -- its per-row losses, its mean loss
          and 512 sampled gradient rows (tolerance 1e-2 of the largest reference value, the bf16 mode's
          contract).  A failure exits non-zero.
"""
from __future__ import annotations

import argparse
import importlib.util
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
PARITY_TOL = 1e-2                 # bf16 mode: 1e-2 relative (BASELINE.json north_star)
PARITY_GRAD_PAIRS = 256           # sampled pairs per rank: both views' rows -> 512 gradient rows
# dram__bytes_read.sum + dram__bytes_write.sum of ONE tc_bwd_kernel<2> launch at config 4 on one GPU, from the
# committed `ncu --set full` capture (profiles/r02_all_kernels_ncu_full_summary.csv, end of round 2: 4.347 MB read + 2.6 KB written): the keys once, nothing else
BWD_TRAFFIC_BYTES_C4_1GPU = 4349952


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle check in front of the timing")
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
PEER_MULTICAST = False


def workload_config(world, eager=False):
    return {"workload": f"NT-Xent fwd+bwd, global batch {N_GLOBAL} pairs ({2 * N_GLOBAL} views), d={DIM}, "
                        f"tau={TAU}, bf16 (BASELINE configs[{3 if N_GLOBAL == 8192 else 4}])",
            "global_batch": N_GLOBAL, "d": DIM, "temperature": TAU,
            "parallelism": (f"rows sharded over {world} rank(s), keys gathered by "
                            + ("the kernels' own peer-memory stores and flag words (NVLink"
                               + (", multimem.st through the NVSwitch multicast mapping" if PEER_MULTICAST else "")
                               + "), no barrier kernel"
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
def host_cores():
    """Cores this process may run on (cgroup / affinity aware)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def reference_module():
    """The reference's own ``NT_Xent`` from an UNMODIFIED copy of its loss.py under the git-ignored
    baseline/_ref/ (put there by ``__graft_entry__.build()`` in the build container; it travels to the GPU box
    like the built .so), or None when that copy is absent -- the oracle's reference-form port is timed then."""
    path = os.path.join(ROOT, "baseline", "_ref", "loss.py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("simclr_reference_loss", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.NT_Xent


def cpu_arm():
    """(kind, step function): one fwd+bwd of the reference loss on host cores in fp32."""
    ref_cls = reference_module()
    if ref_cls is not None:
        crit = ref_cls(temperature=TAU, reduction="mean")          # device = cpu, as the reference defaults
        return "reference", lambda a, b: crit(a, b).backward()
    from oracle import ntxent_oracle as orc
    return "port", lambda a, b: orc.ntxent_reference_form(a, b, TAU, "mean").backward()


def cpu_steps(steps, warmup, threads):
    """fwd+bwd of the reference loss at the FULL batch on `threads` host cores; returns (kind, per-step seconds)."""
    import torch
    torch.set_num_threads(threads)        # torch.distributed.run exports OMP_NUM_THREADS=1: undo that
    kind, step = cpu_arm()
    v0, v1 = synthetic(torch.bfloat16)                       # same bf16-rounded values the GPU sees,
    v0 = v0.float().requires_grad_(True)                     # computed in fp32 like the reference would
    v1 = v1.float().requires_grad_(True)
    times = []
    for i in range(warmup + steps):
        v0.grad = v1.grad = None
        t0 = time.perf_counter()
        step(v0, v1)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    return kind, times


def cpu_what(kind):
    return ("the reference's own loss.py (unmodified copy under baseline/_ref/)" if kind == "reference"
            else "oracle.ntxent_reference_form (port of loss.py's operator sequence; baseline/_ref/loss.py absent)")


def cpu_baseline_block():
    cores = host_cores()
    kind, t = cpu_steps(steps=2, warmup=1, threads=cores)
    sec = statistics.median(t)
    return {"value": N_GLOBAL / sec, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"2 timed + 1 warm-up fwd+bwd steps of {cpu_what(kind)} at the full "
                      f"N={N_GLOBAL}, D={DIM} (fp32 arithmetic on the bf16-rounded inputs), "
                      f"{sec * 1e3:.0f} ms/step on {cores} threads"}


def run_reference(args):
    """The reference loss on the host CPU cores at the full configuration: never a sub-sampled batch.  All the
    cores this process may use, whatever the launcher exported.  When the requested steps would not fit a
    ~4 minute budget the number of timed steps is cut (and the line says so), never the batch."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = host_cores()
    kind, probe = cpu_steps(steps=1, warmup=0, threads=cores)      # also the first warm-up
    budget = 240.0
    warmup = max(0, args.warmup - 1)
    steps = args.steps
    if probe[0] * (steps + warmup) > budget:
        warmup = min(warmup, 1)
        steps = max(2, min(steps, int(budget / probe[0]) - warmup))
    kind, times = cpu_steps(steps, warmup, cores)
    sec = statistics.mean(times)
    value = N_GLOBAL / sec
    sample = (f"each step = fwd+bwd of {cpu_what(kind)} on all {N_GLOBAL} pairs, {cores} threads"
              + ("" if steps == args.steps else f"; {steps} timed steps instead of {args.steps} to stay inside {budget:.0f} s"))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "fp32",
            "data": "synthetic", "config": workload_config(1),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------- parity (checker = the oracle)
def rel_max(got, want):
    import numpy as np
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    return float(np.abs(got - want).max() / max(np.abs(want).max(), 1e-30))


class ParityOracle:
    """fp64 oracle numbers for one rank: every row's lse on the GLOBAL batch (loss.py:25-65 on the
    concatenation of all ranks' view0 / view1) and the gradient of this rank's sampled rows under
    per-row weights `w_local` (2, n_local) that every rank derives from the same formula."""

    def __init__(self, world, rank, n_local, threads):
        import numpy as np
        import torch
        from oracle import ntxent_oracle as orc
        torch.set_num_threads(max(1, threads))
        self.world, self.rank, self.nl = world, rank, n_local
        v0, v1 = synthetic(torch.bfloat16)
        v0, v1 = v0.float().numpy(), v1.float().numpy()
        rng = np.random.default_rng(SEED + 1000 * rank)
        pairs = np.sort(rng.choice(n_local, size=min(PARITY_GRAD_PAIRS, n_local), replace=False))
        self.pairs = pairs
        g = rank * n_local + pairs
        self.rows = np.concatenate([g, N_GLOBAL + g])                  # rows of [view0_all ; view1_all]
        self.w_global = self.weights_global()
        # second weighting: 1 / (2 n_local) on every row = reduction "mean" on each rank
        self.ref = orc.ntxent_sampled_fp64(v0, v1, TAU, [self.w_global, np.full((2, N_GLOBAL), 0.5 / n_local)],
                                           self.rows)
        sl = slice(rank * n_local, (rank + 1) * n_local)
        self.loss_rows = self.ref["loss"][:, sl]                       # (2, n_local) of this rank
        lse = self.ref["lse"].reshape(2, N_GLOBAL)[:, sl]
        self.rowsum = np.exp(lse - 1.0 / TAU).reshape(-1)              # what ntx_fwd calls rowsum

    @staticmethod
    def weights_global():
        """Non-uniform, deterministic row weights (the upstream gradient of reduction='none'), (2, N_GLOBAL),
        rounded to bf16 because the module's output -- and so its upstream gradient -- is bf16."""
        import torch
        i = torch.arange(2 * N_GLOBAL, dtype=torch.float64).reshape(2, N_GLOBAL)
        w = (1.0 + 0.5 * torch.sin(0.37 * i)) / (2.0 * N_GLOBAL)
        return w.to(torch.bfloat16).to(torch.float64).numpy()

    def weights_local(self):
        return self.w_global[:, self.rank * self.nl:(self.rank + 1) * self.nl]

    def grads(self, which):
        dz = self.ref["dz"][0 if which == "none" else 1]
        k = len(self.pairs)
        return dz[:k], dz[k:]                                          # view0 rows, view1 rows


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
    quiet = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
    if quiet:
        quiet(False)        # the leaves are created on the default stream and stepped on the capture stream, on purpose
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
                ex = peer_exchange.exchange_for(_cabi.Backend(), None, 2 * n_local, ld, torch.bfloat16, dev)
                global PEER_MULTICAST
                PEER_MULTICAST = bool(ex.multicast)
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
        rs = be.row_scale(TAU, torch.bfloat16)             # what the module stores into bf16 rows
        be.l2norm(d0, d1, rows, stats[0], row_scale=rs)
        keys = rows
        if world > 1:
            keys = torch.empty(world * q_rows, ld, dtype=torch.bfloat16, device=dev)
            dist.all_gather_into_tensor(keys, rows)
        q = keys[rank * q_rows:(rank + 1) * q_rows]
        plan = be.plan(q_rows, keys.shape[0], DIM, torch.bfloat16)
        ws = torch.empty(max(plan.fwd_ws_bytes, plan.bwd_ws_bytes), dtype=torch.uint8, device=dev)
        loss_sum = torch.empty(2, device=dev)
        lay = (n_local, world, rank, rs)
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
    # step (the kernels + the cross-rank exchange) is captured once into a CUDA graph and replayed.
    # Fresh leaves: they must first be used on the capture side stream, never on the legacy stream.
    sa = d0.detach().clone().requires_grad_(True)
    sb = d1.detach().clone().requires_grad_(True)
    crit_none = NT_Xent(temperature=TAU, reduction="none", device=local_rank, gather=world > 1,
                        gather_backend=GATHER_BACKEND)
    gw = torch.empty(2, n_local, dtype=torch.bfloat16, device=dev)      # upstream gradient of the (2, n) row losses

    def mean_step():
        sa.grad = sb.grad = None
        loss = crit(sa, sb)
        loss.backward()
        return loss

    def none_step():
        sa.grad = sb.grad = None
        rows_out = crit_none(sa, sb)
        rows_out.backward(gw)
        return rows_out

    def build(step_fn):
        """(callable that runs one step, its static output, its static (dview0, dview1), library kernels per step)"""
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(W):
                out = step_fn()
        torch.cuda.current_stream().wait_stream(side)
        barrier()
        if args.no_graphs:
            box = {"out": out}

            def run():
                box["out"] = step_fn()
            c0 = be.launch_count()
            run()
            per_step = be.launch_count() - c0
            return run, (lambda: box["out"]), (lambda: (sa.grad, sb.grad)), per_step, None
        graph = torch.cuda.CUDAGraph()
        sa.grad = sb.grad = None
        c0 = be.launch_count()
        with torch.cuda.graph(graph, capture_error_mode="thread_local"):
            out = step_fn()
        per_step = be.launch_count() - c0
        grads = (sa.grad, sb.grad)
        return graph.replay, (lambda: out), (lambda: grads), per_step, graph

    # ---- parity, before any timing: the oracle on the same global batch against (a) the row sums of the C-ABI
    # forward above, (b) the module with reduction="none" and non-uniform upstream weights through its captured
    # graph, replayed with inputs X, then with the two views swapped (the reference's loss is symmetric under the
    # swap: rows, weights and gradients swap too), then X again; (c) the timed "mean" graph itself.
    parity = None
    if not args.no_parity:
        import numpy as np
        po = ParityOracle(world, rank, n_local, host_cores() // world)
        w_loc = torch.tensor(po.weights_local(), dtype=torch.bfloat16)
        errs = {"rowsum": rel_max(stats[1].float().cpu().numpy(), po.rowsum)}
        run_none, out_none, grads_none, _, graph_none = build(none_step)
        pairs = torch.as_tensor(po.pairs, device=dev)
        want0, want1 = po.grads("none")
        loss_err, grad_err = [], []
        with torch.no_grad():
            x0, x1 = sa.detach().clone(), sb.detach().clone()
        for swap in (False, True, False):
            with torch.no_grad():
                sa.copy_(x1 if swap else x0)
                sb.copy_(x0 if swap else x1)
                gw.copy_(w_loc.flip(0) if swap else w_loc)
            run_none()
            torch.cuda.synchronize()
            rows_got = out_none().detach().float().cpu().numpy()
            g0, g1 = (g.detach()[pairs].float().cpu().numpy() for g in grads_none())
            if swap:
                rows_got, g0, g1 = rows_got[::-1], g1, g0
            loss_err.append(rel_max(rows_got, po.loss_rows))
            grad_err.append(max(rel_max(g0, want0), rel_max(g1, want1)))
        errs["row_loss"], errs["grad_none"] = max(loss_err), max(grad_err)
        del run_none, out_none, grads_none, graph_none
    step, out_mean, grads_mean, kernels_per_step, graph_mean = build(mean_step)
    static_loss = out_mean()
    if not args.no_parity:
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        want0, want1 = po.grads("mean")
        g0, g1 = (g.detach()[pairs].float().cpu().numpy() for g in grads_mean())
        errs["grad_mean"] = max(rel_max(g0, want0), rel_max(g1, want1))
        errs["loss_mean"] = rel_max(float(out_mean().detach()), float(po.loss_rows.mean()))
        ok = all(np.isfinite(v) and v < PARITY_TOL for v in errs.values())
        flag = torch.tensor([1.0 if ok else 0.0, errs["row_loss"], errs["rowsum"],
                             max(errs["grad_none"], errs["grad_mean"]), errs["loss_mean"]], dtype=torch.float64)
        if world > 1:
            lo = flag[:1].clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=ctl)
            dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=ctl)       # worst rank
            flag[0] = lo[0]
        parity = {"ok": bool(flag[0] > 0.5), "tolerance": PARITY_TOL, "max_rel_loss": float(flag[1]),
                  "max_rel_rowsum": float(flag[2]), "max_rel_grad": float(flag[3]),
                  "max_rel_mean_loss": float(flag[4]),
                  "rows_checked": 2 * n_local * world, "grad_rows_checked": 2 * len(po.pairs) * world,
                  "replays": 3, "oracle": "oracle.ntxent_sampled_fp64 (fp64, all-row lse on the global batch)",
                  "how": "max over ranks of max|got - want| / max|want|; per-row losses and sampled gradient rows "
                         "through the captured graph of the module (reduction='none', non-uniform upstream "
                         "weights; inputs X, views swapped, X), the timed 'mean' graph's loss and gradients after "
                         "3 more replays, row sums through the C ABI"}
        if not parity["ok"]:
            if rank == 0:
                print(json.dumps({"parity": parity, "error": "CUDA path disagrees with the oracle"}), flush=True)
            sys.stdout.flush()
            os._exit(3)
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
        return float(out_mean().detach())
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
            loss_host[k & 1].copy_(out_mean().detach(), non_blocking=True)
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
            "gpu_launches": kernels_per_step * K,
            "gpu_launches_how": f"{kernels_per_step} kernels of libntxent_b200.so per step (counted by the library while "
                                "the step was captured / launched: ntx_launch_count) x timed steps",
            "parity": parity,
            "roofline": {"bound": "tensor", "kernel": f"tc_bwd_kernel<{DIM // 64}, {'true' if world > 1 and GATHER_BACKEND == 'peer' else 'false'}, true, false>",
                         "achieved": achieved, "peak": peak,
                         "unit": "TFLOP/s", "frac": achieved / peak,
                         "traffic": BWD_TRAFFIC_BYTES_C4_1GPU if (world == 1 and N_GLOBAL == 8192 and DIM == 128) else None,
                         "traffic_source": "ncu --set full, profiles/r02_all_kernels_ncu_full_summary.csv (dram bytes read + written per launch of this kernel, captured once; not measured by this run)",
                         "peak_source": peak_src, "algorithmic_flops_per_launch": bwd_flops,
                         "kernel_ms": bwd_ms,
                         "executed_flops_per_launch": bwd_executed,
                         "executed_tflops": bwd_executed / (bwd_ms * 1e-3) / 1e12,
                         "peak_sustained": peak_sustained,
                         "executed_frac_of_sustained": (bwd_executed / (bwd_ms * 1e-3) / 1e12 / peak_sustained
                                                        if peak_sustained else None),
                         "fwd_kernel": {"kernel": (f"tc_fwd_sym_kernel<{DIM // 64}, true>" if world == 1 and DIM <= 128
                                                   else f"tc_fwd_kernel<{DIM // 64}, true, false>"), "kernel_ms": fwd_ms,
                                        "achieved": fwd_flops / (fwd_ms * 1e-3) / 1e12},
                         "whole_step_frac": 18.0 * N_GLOBAL * N_GLOBAL * DIM / world / (ms_per_step * 1e-3) / 1e12 / peak},
            "clocks": clocks.summary(),
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_block()
        print(json.dumps(line), flush=True)
    # ---- orderly teardown: graphs first (they hold the symmetric-memory buffers), then the exchange objects, then
    # the process groups.  A watchdog names the stage if anything stalls instead of hiding it.
    stage = {"name": "start"}

    def watchdog():
        time.sleep(60)
        print(f"[bench] rank {rank}: teardown stalled in stage '{stage['name']}' for 60 s; exiting", file=sys.stderr,
              flush=True)
        os._exit(0)
    threading.Thread(target=watchdog, daemon=True).start()
    sys.stdout.flush()
    stage["name"] = "synchronize"
    torch.cuda.synchronize()
    stage["name"] = "drop graphs"
    del step, out_mean, grads_mean, graph_mean, static_loss
    if world > 1:
        stage["name"] = "drop peer exchange"
        from simclr_b200 import peer_exchange
        peer_exchange.close_all()
        torch.cuda.synchronize()
        stage["name"] = "control barrier"
        dist.barrier(group=ctl)
        stage["name"] = "destroy_process_group"
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    a = parse()
    if a.workload == "c5":
        N_GLOBAL, DIM = 32768, 256
    sys.exit(run_reference(a) if a.impl == "reference" else run_ours(a))
