"""Drop-in ``NT_Xent`` for nzw0301/SimCLR's ``loss.py`` on B200.

Same class name, constructor (``temperature``, ``reduction``, ``device``) and call
signature as the reference module (loss.py:4-65), so ``main.py:83`` /
``main.py:114-115`` run unchanged after ``from simclr_b200.loss import NT_Xent``.
The arithmetic runs in ``libntxent_b200.so`` (hand-written sm_100a kernels) through
one ``torch.autograd.Function``; nothing of size (2N)^2 is ever allocated and the
host is never synchronised (the reference syncs twice per call: the CPU-built mask
at loss.py:38 and ``nonzero`` behind the boolean index at loss.py:46-47).

New, keyword-only and off by default: ``gather=True`` makes the negatives span every
rank of the process group (rows stay sharded by rank, keys are all-gathered) -- the
reference computes the loss on each rank's local batch only.
"""
from __future__ import annotations

import contextlib
import os

import torch
import torch.distributed as dist

from . import _cabi, peer_exchange

MIN_TC_TEMPERATURE = 0.0235         # NTX_MIN_TEMPERATURE of the library: below it only fp32 rows are served
FP32_TC_MIN_KEYS = 2048             # fp32 inputs with at least this many keys use the two-term tensor-core path,
FP32_TC_MIN_TEMPERATURE = 0.35      # from this temperature upwards.  Its 2^-17 rounding grows with 1/temperature:
                                    # gradient error against the fp64 oracle 2e-7 at 0.5, 1e-6 at 0.1 on random rows, but
                                    # on tightly clustered rows (cosines near 1, gradients are small differences of
                                    # large sums) 3e-6 at 0.5, 3.4e-6 at 0.35, 1e-5 at 0.25, 1.4e-5 at 0.1: the 1e-5
                                    # contract keeps a factor 3 in hand from 0.35 upwards
_backend = _cabi.Backend()          # tests may swap in a stand-in with the same methods
_workspaces: dict = {}
_shapes: dict = {}


def _shape_info(be, q_rows: int, k_rows: int, d: int, rows_dtype: torch.dtype, dev: torch.device):
    """(row pitch, scratch bytes) for a problem shape, asked from the library once per shape."""
    key = (id(be), q_rows, k_rows, d, rows_dtype, dev)
    info = _shapes.get(key)
    if info is None:
        ld = be.row_pitch(d, rows_dtype)
        plan = be.plan(q_rows, k_rows, d, rows_dtype)
        info = _shapes[key] = (ld, max(plan.fwd_ws_bytes, plan.bwd_ws_bytes))
    return info


def _workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    """Scratch for the partial row sums / dZ slabs, cached per (device, stream)."""
    key = (device, _cabi.current_stream_handle(device) if device.type == "cuda" else 0)
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < nbytes:
        if len(_workspaces) >= 16:          # streams come and go (their handles are reused): do not grow without bound
            _workspaces.clear()
        buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = buf
    return buf


def _require_cuda(t: torch.Tensor) -> None:
    if t.device.type != "cuda":
        raise RuntimeError(f"NT_Xent (B200) has no CPU path: inputs must be CUDA tensors (got {t.device})")


def _device_of(t: torch.Tensor):
    """Context that makes the tensor's GPU current: the library plans, opts kernels in and launches on the
    current device, while the stream is taken from the tensor's device."""
    return torch.cuda.device(t.device) if t.device.type == "cuda" else contextlib.nullcontext()


def _fp32_tc_enabled() -> bool:
    return os.environ.get("NTXENT_FP32_TC", "1").lower() not in ("0", "false", "no", "off")


def _env_flag(name: str) -> bool:
    return os.environ.get(name, "0").lower() in ("1", "true", "yes", "on")


class NTXentFunction(torch.autograd.Function):
    """forward: loss of this rank's rows; backward: d loss / d view0, d view1.

    Saves only O(N*D) tensors: the inputs, the normalised rows (keys when gathering),
    the inverse norms and the per-row sums of exponentials.
    """

    @staticmethod
    def forward(ctx, view0, view1, temperature, reduction, rows_dtype, group, world, rank, peer=False, cat=False):
        be = _backend
        if cat:                      # one (2N, D) tensor [view0 ; view1]: its halves are used in place, no copy
            half = view0.shape[0] // 2
            view0, view1 = view0[:half], view0[half:]
        n, d = view0.shape
        dev = view0.device
        q_rows = 2 * n
        ld, ws_bytes = _shape_info(be, q_rows, world * q_rows, d, rows_dtype, dev)
        stats = torch.empty((4, q_rows), dtype=torch.float32, device=dev)   # inv_norm, rowsum, pos, row loss
        loss_sum = torch.empty(2, dtype=view0.dtype, device=dev)            # [sum, mean] in the dtype the loss is returned in
        inv_norm, rowsum, pos_logit, row_loss = stats[0], stats[1], stats[2], stats[3]
        exchange = None
        ws = _workspace(dev, ws_bytes)
        # bf16 rows are stored times sqrt(log2 e / temperature): a similarity then is its logit's exponent and the
        # tensor-core epilogues skip one packed FMA per pair of logits (1 for fp32 rows)
        rs = be.row_scale(temperature, rows_dtype)
        if world > 1 and peer:
            # kernels store rows, row sums and weights straight into every rank's exchange block (NVLink peer
            # memory) and order them with flag words: no collective call, no barrier kernel
            exchange = peer_exchange.exchange_for(be, group, q_rows, ld, rows_dtype, dev)
            exchange.begin_forward(any(ctx.needs_input_grad[:2]))
            keys = torch.empty(1, dtype=torch.int32, device=dev)           # the step number, set on the device
            be.peer_l2norm(view0, view1, exchange.blocks, ld, rows_dtype, inv_norm, keys, row_scale=rs)
            be.peer_fwd(exchange.blocks, n, d, ld, rows_dtype, temperature, keys, rowsum, pos_logit, row_loss,
                        loss_sum, ws, row_scale=rs)
        else:
            # two-term rows (fp32 values as [hi | lo] bf16 terms) live in a bf16 tensor: name the layout to the library
            store = _cabi.storage_dtype(rows_dtype)
            kind = {} if store is rows_dtype else {"rows_dtype": rows_dtype}
            rows = torch.empty((q_rows, ld), dtype=store, device=dev)
            be.l2norm(view0, view1, rows, inv_norm, row_scale=rs, **kind)
            if world > 1:
                keys = torch.empty((world * q_rows, ld), dtype=store, device=dev)
                dist.all_gather_into_tensor(keys, rows, group=group)      # rank-major key buffer
                q = keys[rank * q_rows:(rank + 1) * q_rows]
            else:
                keys = q = rows
            be.fwd(q, keys, d, (n, world, rank, rs), temperature, rowsum, pos_logit, row_loss, loss_sum, ws, **kind)

        ctx.save_for_backward(view0, view1, keys, stats)
        ctx.meta = (temperature, reduction, group, world, rank, d, exchange, rows_dtype, cat, rs)
        if reduction == "none":
            return row_loss.view(2, n).to(view0.dtype, copy=True)          # loss.py:55-58
        return loss_sum[0] if reduction == "sum" else loss_sum[1]          # loss.py:60-65

    @staticmethod
    def backward(ctx, grad_out):
        be = _backend
        view0, view1, keys, stats = ctx.saved_tensors
        temperature, reduction, group, world, rank, d, exchange, rows_dtype, cat, rs = ctx.meta
        n = view0.shape[0]
        q_rows = 2 * n
        dev = view0.device
        inv_norm, rowsum = stats[0], stats[1]
        g = grad_out.detach().contiguous()            # read by the kernels in its own dtype (fp32 or bf16)
        if g.dtype not in (torch.float32, torch.bfloat16):
            g = g.to(torch.float32)
        # row weights w_r of  sum_r w_r * row_loss_r : one upstream gradient (mean / sum) or one per row (none)
        w_stride = 1 if reduction == "none" else 0
        w_scale = 1.0 / q_rows if reduction == "mean" else 1.0
        w_q = g.view(-1)
        if cat:                      # one gradient tensor for the one input; the kernels write its halves
            dcat = torch.empty((q_rows, d), dtype=view0.dtype, device=dev)
            dview0, dview1 = dcat[:n], dcat[n:]
        else:
            dview0 = torch.empty_like(view0)
            dview1 = torch.empty_like(view1)
        grads = (dcat, None) if cat else (dview0, dview1)
        ld, ws_bytes = _shape_info(be, q_rows, world * q_rows, d, rows_dtype, dev)
        ws = _workspace(dev, ws_bytes)
        if exchange is not None:
            # `keys` is the step number here: the library sends this rank's row weights to every rank and reads
            # every key's row sum and weight from its own exchange block
            be.peer_post_weights(exchange.blocks, n, ld, rows_dtype, keys, w_q, w_stride)
            be.peer_bwd(exchange.blocks, n, d, ld, rows_dtype, temperature, keys, rowsum, w_stride, w_scale, view0,
                        view1, inv_norm, dview0, dview1, ws, row_scale=rs)
            exchange.end_backward()
            return grads + (None,) * 8
        if world > 1:
            # one collective for both per-key vectors: every rank contributes [rowsum ; row weights]
            mine = torch.stack((rowsum, w_q.to(torch.float32).expand(q_rows)))     # (2, q_rows)
            both = torch.empty((world * 2, q_rows), dtype=torch.float32, device=dev)
            dist.all_gather_into_tensor(both, mine, group=group)
            both = both.view(world, 2, q_rows)
            rowsum_k = both[:, 0, :].reshape(-1)                                   # rank-major, like the keys
            w_k = both[:, 1, :].reshape(-1)
            w_q, w_stride = w_k[rank * q_rows:(rank + 1) * q_rows], 1              # ranks may hold different upstream gradients
            q = keys[rank * q_rows:(rank + 1) * q_rows]
        else:
            rowsum_k, w_k, q = rowsum, w_q, keys
        kind = {} if _cabi.storage_dtype(rows_dtype) is rows_dtype else {"rows_dtype": rows_dtype}
        be.bwd(q, keys, d, (n, world, rank, rs), temperature, rowsum, rowsum_k, w_q, w_k, w_stride, w_scale,
               view0, view1, inv_norm, dview0, dview1, ws, **kind)
        return grads + (None,) * 8


class NTXentHiddenFunction(torch.autograd.Function):
    """NT-Xent taken one layer earlier: inputs are the activations in front of the head's last
    ``Linear(h, d, bias=False)`` (model.py:69) and that layer's weight.  The projections z = h @ W.T exist only
    inside a tcgen05 GEMM whose epilogue normalises them (``ntx_project_l2norm``); backward gets dz from the
    normalised rows (``ntx_bwd_rows``) and finishes the layer with two library GEMMs."""

    @staticmethod
    def forward(ctx, h0, h1, weight, temperature, reduction, cat=False):
        be = _backend
        if cat:                      # one (2N, H) tensor [h0 ; h1]: its halves are used in place
            half = h0.shape[0] // 2
            h0, h1 = h0[:half], h0[half:]
        n = h0.shape[0]
        d = weight.shape[0]
        dev = h0.device
        q_rows = 2 * n
        ld, ws_bytes = _shape_info(be, q_rows, q_rows, d, torch.bfloat16, dev)
        rows = torch.empty((q_rows, ld), dtype=torch.bfloat16, device=dev)
        stats = torch.empty((4, q_rows), dtype=torch.float32, device=dev)   # inv_norm, rowsum, pos, row loss
        loss_sum = torch.empty(2, dtype=h0.dtype, device=dev)
        rs = be.row_scale(temperature, torch.bfloat16)
        be.project_l2norm(h0, h1, weight, rows, stats[0], row_scale=rs)
        be.fwd(rows, rows, d, (n, 1, 0, rs), temperature, stats[1], stats[2], stats[3], loss_sum,
               _workspace(dev, ws_bytes))
        ctx.save_for_backward(h0, h1, weight, rows, stats)
        ctx.meta = (temperature, reduction, d, ws_bytes, cat, rs)
        if reduction == "none":
            return stats[3].view(2, n).to(h0.dtype, copy=True)
        return loss_sum[0] if reduction == "sum" else loss_sum[1]

    @staticmethod
    def backward(ctx, grad_out):
        be = _backend
        h0, h1, weight, rows, stats = ctx.saved_tensors
        temperature, reduction, d, ws_bytes, cat, rs = ctx.meta
        n = h0.shape[0]
        q_rows = 2 * n
        g = grad_out.detach().contiguous().view(-1)
        if g.dtype not in (torch.float32, torch.bfloat16):
            g = g.to(torch.float32)
        w_stride = 1 if reduction == "none" else 0
        w_scale = 1.0 / q_rows if reduction == "mean" else 1.0
        dz = torch.empty((q_rows, d), dtype=torch.bfloat16, device=h0.device)
        be.bwd_rows(rows, rows, d, (n, 1, 0, rs), temperature, stats[1], stats[1], g, g, w_stride, w_scale, stats[0],
                    dz, _workspace(h0.device, ws_bytes))
        hdim = h0.shape[1]
        dh = torch.empty((q_rows, hdim), dtype=torch.bfloat16, device=h0.device)
        if d <= 128 and d % 8 == 0 and (n * d) % 8 == 0:
            # both GEMMs of the layer from one tcgen05 kernel: each dz tile feeds dh = dz . W and dW += dz^T . h
            dw = torch.empty_like(weight)
            ws2 = _workspace(h0.device, max(ws_bytes, be.project_bwd_workspace_bytes(n, hdim)))
            be.project_bwd(dz, h0, h1, weight, dh[:n], dh[n:], dw, ws2)
        else:
            # shapes the kernel is not built for (d > 128, odd d): library GEMMs
            torch.matmul(dz, weight, out=dh)                                 # (2n, d) . (d, h)
            dw = torch.addmm(dz[n:].t() @ h1, dz[:n].t(), h0)                # dz^T . [h0 ; h1]
        if cat:
            return dh, None, dw, None, None, None
        return dh[:n], dh[n:], dw, None, None, None


class NT_Xent(torch.nn.Module):
    """
    Normalised temperature-scaled cross entropy loss (B200 implementation).

    Constructor and ``forward`` follow the reference module (loss.py:9-31); ``device``
    is accepted for compatibility (main.py:83 passes an int) -- the kernels run on the
    device of the inputs, which must be CUDA tensors.
    """

    def __init__(self, temperature: float = 0.1, reduction: str = "mean",
                 device: torch.device = torch.device("cpu"), *, gather: bool | None = None,
                 process_group=None, compute_dtype: torch.dtype | None = None,
                 gather_backend: str | None = None) -> None:
        """
        :param temperature: Temperature parameter. The value must be positive.  Below 0.0235 the kernels' fixed-shift
            logsumexp would leave fp32 range (exp((cos - 1) / temperature)): such temperatures run on fp32 rows with
            a max-shifted logsumexp, slower but exact for any positive value, like the reference.
        :param reduction: Same to PyTorch's `reduction` in losses.
        :param device: PyTorch's device instance (kept for signature compatibility).
        :param gather: keyword-only. True = negatives from every rank of ``process_group``;
            None = take it from the environment variable NTXENT_GATHER (default off, which
            is the reference's behaviour: each rank's loss sees only its local batch).
        :param process_group: keyword-only, group used when gathering (default: world).
        :param gather_backend: keyword-only. "nccl" (all-gather collectives, any topology) or "peer" (the
            kernels store into every rank's buffers over NVLink peer memory; one node, at most 8 ranks,
            forward and backward calls alternating). None = NTXENT_GATHER_BACKEND, default "nccl".
        :param compute_dtype: keyword-only. dtype of the normalised rows: torch.float32 (CUDA-core
            kernels, 1e-5 parity), torch.bfloat16 (tcgen05 kernels) or None = the input dtype
            (NTXENT_COMPUTE_DTYPE=bf16 overrides None).
        """
        assert temperature > 0.
        assert reduction in {"none", "mean", "sum"}

        super(NT_Xent, self).__init__()
        self.temperature = temperature
        self.reduction = reduction
        self.device = device
        self.gather = _env_flag("NTXENT_GATHER") if gather is None else bool(gather)
        self.process_group = process_group
        gather_backend = gather_backend or os.environ.get("NTXENT_GATHER_BACKEND", "nccl").lower()
        if gather_backend not in ("nccl", "peer"):
            raise ValueError(f"gather_backend must be 'nccl' or 'peer', got {gather_backend!r}")
        self.gather_backend = gather_backend
        if compute_dtype is None and os.environ.get("NTXENT_COMPUTE_DTYPE", "").lower() in ("bf16", "bfloat16"):
            compute_dtype = torch.bfloat16
        if compute_dtype not in (None, torch.float32, torch.bfloat16):
            raise TypeError(f"compute_dtype must be torch.float32 or torch.bfloat16, got {compute_dtype}")
        self.compute_dtype = compute_dtype

    def forward(self, view0: torch.Tensor, view1: torch.Tensor) -> torch.Tensor:
        """
        :param view0: Feature representation. The shape is (N, D).
        :param view1: The other feature representations. The shape is same to `view0`'s shape.
        :return: Loss value. The shape depends on `reduction`: (2, N) or a scaler.
        """
        if view0.dim() != 2 or view1.dim() != 2:
            raise RuntimeError(f"NT_Xent expects two (N, D) tensors, got {tuple(view0.shape)} and "
                               f"{tuple(view1.shape)}")
        if view0.shape[0] != view1.shape[0]:
            raise IndexError(f"view0 has {view0.shape[0]} rows but view1 has {view1.shape[0]}")
        if view0.shape[1] != view1.shape[1]:
            raise RuntimeError(f"feature sizes differ: {view0.shape[1]} vs {view1.shape[1]}")
        if view0.dtype != view1.dtype or view0.device != view1.device:
            raise RuntimeError("view0 and view1 must share dtype and device")
        if view0.dtype in (torch.float16, torch.float64):
            # the reference takes these too (its result has the input dtype); here they are computed through the fp32
            # path -- more accurate than the reference's fp16 arithmetic, fp32-accurate for fp64 inputs -- with two
            # differentiable casts around it
            return self.forward(view0.to(torch.float32), view1.to(torch.float32)).to(view0.dtype)
        if view0.dtype not in (torch.float32, torch.bfloat16):
            raise TypeError(f"NT_Xent (B200) supports floating-point inputs (float32, bfloat16; float16 and float64 "
                            f"through float32), got {view0.dtype}")
        _require_cuda(view0)
        if view0.shape[0] < 1 or view0.shape[1] < 1:
            raise RuntimeError("empty batch or feature dimension")

        world, rank, group = 1, 0, None
        if self.gather and dist.is_available() and dist.is_initialized():
            group = self.process_group
            world, rank = dist.get_world_size(group), dist.get_rank(group)
        rows_dtype = self._rows_dtype(view0, 2 * view0.shape[0] * world)
        with _device_of(view0):
            return NTXentFunction.apply(view0.contiguous(), view1.contiguous(), float(self.temperature),
                                        self.reduction, rows_dtype, group, world, rank,
                                        world > 1 and self.gather_backend == "peer")

    def _rows_dtype(self, x: torch.Tensor, keys: int = 0):
        """Layout of the normalised rows: ``compute_dtype``, else the input dtype -- except that
        * bf16 rows wider than the tensor-core kernels' 256 features per tile, or at a temperature below their
          fixed-shift range, fall back to fp32 rows on the CUDA-core kernels (any D, any temperature > 0), so the
          module accepts every feature size and temperature the reference does;
        * fp32 rows of a large batch (``keys`` >= FP32_TC_MIN_KEYS gathered rows, D <= 128, temperature >=
          FP32_TC_MIN_TEMPERATURE; either gather back end) are held as two bf16 terms and go through the tcgen05 kernels
          (``_cabi.BF16X2``): fp32-grade results, an order of magnitude faster than the CUDA-core kernels at
          N = 8192.  ``NTXENT_FP32_TC=0`` keeps fp32 on the CUDA cores."""
        dt = self.compute_dtype or x.dtype
        if dt == torch.bfloat16 and (x.shape[1] > 256 or self.temperature < MIN_TC_TEMPERATURE):
            return torch.float32
        if (dt == torch.float32 and keys >= FP32_TC_MIN_KEYS and x.shape[1] <= 128
                and self.temperature >= FP32_TC_MIN_TEMPERATURE and _fp32_tc_enabled()):
            return _cabi.BF16X2
        return dt

    def forward_cat(self, z: torch.Tensor) -> torch.Tensor:
        """
        The loss of ``view0 = z[:N]``, ``view1 = z[N:]`` for one ``(2N, D)`` tensor ``z = [view0 ; view1]``, as a
        training loop gets it from ONE batched call of the encoder on both views (``z = model(torch.cat([x0, x1]))``
        instead of the two calls at main.py:112-113).  The halves are used in place -- no split or concatenation
        copies in forward, one ``(2N, D)`` gradient in backward -- and the result equals ``forward(z[:N], z[N:])``
        bit for bit.  Opt-in: batching the two views through the encoder changes its BatchNorm statistics (one
        joint batch instead of one batch per view), which is a change of the training recipe, not of this loss.

        :param z: Feature representations of both views, shape (2N, D): rows [0, N) are the first view.
        :return: same as ``forward``.
        """
        if z.dim() != 2 or z.shape[0] % 2 != 0 or z.shape[0] < 2 or z.shape[1] < 1:
            raise RuntimeError(f"forward_cat expects one (2N, D) tensor, got {tuple(z.shape)}")
        if z.dtype in (torch.float16, torch.float64):          # as in forward: through the fp32 path
            return self.forward_cat(z.to(torch.float32)).to(z.dtype)
        if z.dtype not in (torch.float32, torch.bfloat16):
            raise TypeError(f"NT_Xent (B200) supports floating-point inputs (float32, bfloat16; float16 and float64 "
                            f"through float32), got {z.dtype}")
        _require_cuda(z)
        world, rank, group = 1, 0, None
        if self.gather and dist.is_available() and dist.is_initialized():
            group = self.process_group
            world, rank = dist.get_world_size(group), dist.get_rank(group)
        rows_dtype = self._rows_dtype(z, z.shape[0] * world)
        with _device_of(z):
            return NTXentFunction.apply(z.contiguous(), None, float(self.temperature), self.reduction, rows_dtype,
                                        group, world, rank, world > 1 and self.gather_backend == "peer", True)

    def forward_hidden(self, h0: torch.Tensor, h1: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
        """
        The loss of ``z0 = h0 @ weight.T``, ``z1 = h1 @ weight.T`` without materialising ``z0``, ``z1``: the last
        layer of the projection head (``Linear(h, d, bias=False)``, model.py:69) runs fused with the normalisation.
        bfloat16 only, single rank, hidden size a multiple of 64, d <= 256.

        :param h0: activations in front of the head's last layer for the first view, shape (N, H).
        :param h1: the same for the other view.
        :param weight: that layer's ``weight``, shape (D, H).
        :return: same as ``forward``.
        """
        if h0.dim() != 2 or h1.shape != h0.shape or weight.dim() != 2 or weight.shape[1] != h0.shape[1]:
            raise RuntimeError(f"forward_hidden expects (N, H), (N, H), (D, H); got {tuple(h0.shape)}, "
                               f"{tuple(h1.shape)}, {tuple(weight.shape)}")
        if not (h0.dtype == h1.dtype == weight.dtype == torch.bfloat16):
            raise TypeError("forward_hidden runs on the bf16 tensor-core path: h0, h1 and weight must be bfloat16")
        _require_cuda(h0)
        with _device_of(h0):
            return NTXentHiddenFunction.apply(h0.contiguous(), h1.contiguous(), weight.contiguous(),
                                              float(self.temperature), self.reduction)

    def forward_hidden_cat(self, h: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
        """``forward_hidden`` for one ``(2N, H)`` tensor ``h = [h0 ; h1]`` (both views batched through the encoder,
        see ``forward_cat``): halves used in place, one ``(2N, H)`` gradient.

        :param h: activations in front of the head's last layer for both views, shape (2N, H).
        :param weight: that layer's ``weight``, shape (D, H).
        :return: same as ``forward``.
        """
        if h.dim() != 2 or h.shape[0] % 2 != 0 or h.shape[0] < 2 or weight.dim() != 2 or weight.shape[1] != h.shape[1]:
            raise RuntimeError(f"forward_hidden_cat expects (2N, H), (D, H); got {tuple(h.shape)}, {tuple(weight.shape)}")
        if not (h.dtype == weight.dtype == torch.bfloat16):
            raise TypeError("forward_hidden_cat runs on the bf16 tensor-core path: h and weight must be bfloat16")
        _require_cuda(h)
        with _device_of(h):
            return NTXentHiddenFunction.apply(h.contiguous(), None, weight.contiguous(), float(self.temperature),
                                              self.reduction, True)

    def extra_repr(self) -> str:
        return (f"temperature={self.temperature}, reduction={self.reduction!r}, gather={self.gather}, "
                f"compute_dtype={self.compute_dtype}")
