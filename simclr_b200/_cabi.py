"""ctypes binding of libntxent_b200.so (include/ntxent_b200.h).

The library is the product; there is no CPU or pure-PyTorch fallback.  If it
has not been built, loading raises with the command that builds it.
"""
from __future__ import annotations

import ctypes
import os
import threading

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NTXENT_B200_LIB") or os.path.join(_PKG, "libntxent_b200.so")

NTX_F32, NTX_BF16, NTX_BF16X2 = 0, 1, 2


class _RowsKind:
    """A rows layout that is not a torch dtype: stands where a ``rows_dtype`` is expected."""

    def __init__(self, name: str, storage: torch.dtype):
        self.name, self.storage = name, storage

    def __repr__(self):
        return f"rows:{self.name}"


# fp32 rows held as two bf16 terms [hi | lo] (NTX_BF16X2 of include/ntxent_b200.h): stored in a bf16 tensor of
# ntx_row_pitch(d, NTX_BF16X2) columns, computed on the tensor cores to fp32-grade accuracy
BF16X2 = _RowsKind("bf16x2", torch.bfloat16)
DTYPES = {torch.float32: NTX_F32, torch.bfloat16: NTX_BF16, BF16X2: NTX_BF16X2}


def storage_dtype(rows_dtype) -> torch.dtype:
    """torch dtype of the tensor that holds rows of `rows_dtype`."""
    return rows_dtype.storage if isinstance(rows_dtype, _RowsKind) else rows_dtype


EXPORTS = (
    "ntx_version", "ntx_last_error", "ntx_row_pitch", "ntx_row_scale", "ntx_plan", "ntx_l2norm", "ntx_fwd",
    "ntx_bwd", "ntx_debug_indices", "ntx_profile_next", "ntx_fwd_static", "ntx_peer_block_bytes",
    "ntx_peer_l2norm", "ntx_peer_fwd", "ntx_peer_post_weights", "ntx_peer_bwd", "ntx_project_l2norm", "ntx_bwd_rows", "ntx_launch_count",
    "ntx_larc_workspace_bytes", "ntx_larc_sgd_step", "ntx_project_bwd_workspace_bytes", "ntx_project_bwd",
    "ntx_simclr_augment",
)

MAX_PEERS = 8


class NtxLayout(ctypes.Structure):
    _fields_ = [("n_local", ctypes.c_int32), ("world", ctypes.c_int32), ("rank", ctypes.c_int32),
                ("row_scale", ctypes.c_float), ("loss_dtype", ctypes.c_int32)]


class NtxPeerBlocks(ctypes.Structure):
    _fields_ = [("base", ctypes.c_void_p * MAX_PEERS), ("multicast", ctypes.c_void_p), ("world", ctypes.c_int32),
                ("rank", ctypes.c_int32)]


class NtxLarcParam(ctypes.Structure):
    _fields_ = [("param", ctypes.c_void_p), ("grad", ctypes.c_void_p), ("momentum", ctypes.c_void_p),
                ("numel", ctypes.c_int64), ("weight_decay", ctypes.c_float), ("first_chunk", ctypes.c_int32),
                ("n_chunks", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class NtxPlan(ctypes.Structure):
    _fields_ = [
        ("tile_rows", ctypes.c_int32), ("tile_keys", ctypes.c_int32), ("row_blocks", ctypes.c_int32),
        ("key_tiles", ctypes.c_int32), ("grid", ctypes.c_int32), ("max_slots", ctypes.c_int32),
        ("d_pad", ctypes.c_int32), ("tensor_cores", ctypes.c_int32),
        ("fwd_ws_bytes", ctypes.c_uint64), ("bwd_ws_bytes", ctypes.c_uint64),
        ("fwd_ws_bytes_min", ctypes.c_uint64), ("fwd_symmetric", ctypes.c_int32),
        ("fwd_grid", ctypes.c_int32), ("fwd_max_slots", ctypes.c_int32), ("segments", ctypes.c_int32),
    ]


class NtxError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libntxent_b200 error {code}: {message}")
        self.code = code


_lock = threading.Lock()
_lib = None


def _declare(lib):
    vp, i, f, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_size_t
    lib.ntx_version.restype = i
    lib.ntx_last_error.restype = ctypes.c_char_p
    lib.ntx_row_pitch.argtypes = [i, i]
    lib.ntx_row_scale.argtypes = [f, i]
    lib.ntx_plan.argtypes = [i, i, i, i, i, ctypes.POINTER(NtxPlan)]
    lib.ntx_l2norm.argtypes = [vp, vp, i, i, i, vp, i, i, vp, f, vp]
    lib.ntx_fwd.argtypes = [vp, vp, i, i, i, i, i, ctypes.POINTER(NtxLayout), f, vp, vp, vp, vp, vp, sz, vp]
    lib.ntx_bwd.argtypes = [vp, vp, i, i, i, i, i, ctypes.POINTER(NtxLayout), f, vp, vp, vp, vp, i, i, f,
                            vp, vp, i, vp, vp, vp, vp, sz, vp]
    lib.ntx_fwd_static.argtypes = [vp, vp, i, i, i, i, i, ctypes.POINTER(NtxLayout), f, vp, vp, vp, vp, vp]
    pb = ctypes.POINTER(NtxPeerBlocks)
    lib.ntx_peer_block_bytes.argtypes = [i, i, i, i]
    lib.ntx_peer_l2norm.argtypes = [vp, vp, i, i, i, pb, i, i, vp, f, vp, vp]
    lib.ntx_peer_fwd.argtypes = [pb, i, i, i, i, f, f, i, vp, vp, vp, vp, vp, vp, sz, vp]
    lib.ntx_peer_post_weights.argtypes = [pb, i, i, i, vp, vp, i, i, vp]
    lib.ntx_peer_bwd.argtypes = [pb, i, i, i, i, f, f, vp, vp, i, f, vp, vp, i, vp, vp, vp, vp, sz, vp]
    lib.ntx_project_l2norm.argtypes = [vp, vp, vp, i, i, i, vp, i, vp, f, vp]
    lib.ntx_bwd_rows.argtypes = [vp, vp, i, i, i, i, i, ctypes.POINTER(NtxLayout), f, vp, vp, vp, vp, i, i, f, vp, vp,
                                 vp, sz, vp]
    lib.ntx_debug_indices.argtypes = [ctypes.POINTER(NtxLayout), vp, vp, vp]
    lib.ntx_profile_next.argtypes = [i, vp, vp]
    for name in EXPORTS:
        if name not in ("ntx_last_error", "ntx_row_scale"):
            getattr(lib, name).restype = i
    lib.ntx_project_bwd_workspace_bytes.argtypes = [i, i, i]
    lib.ntx_project_bwd_workspace_bytes.restype = ctypes.c_size_t
    lib.ntx_project_bwd.argtypes = [vp, vp, vp, vp, i, i, i, vp, vp, vp, vp, sz, vp]
    lib.ntx_simclr_augment.argtypes = [vp, i, i, i, f, ctypes.c_ulonglong, ctypes.c_ulonglong, i, vp, vp, vp, vp]
    lib.ntx_larc_workspace_bytes.argtypes = [i, i]
    lib.ntx_larc_sgd_step.argtypes = [vp, i, vp, i, i, f, f, f, f, i, vp, sz, vp]
    lib.ntx_launch_count.restype = ctypes.c_longlong
    lib.ntx_row_scale.restype = ctypes.c_float
    lib.ntx_larc_workspace_bytes.restype = ctypes.c_size_t
    lib.ntx_peer_block_bytes.restype = ctypes.c_size_t


def load():
    """The loaded library (built by ``python -m simclr_b200.build`` / ``__graft_entry__.build()``)."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise RuntimeError(
                        f"{LIB_PATH} is missing: the CUDA library is the only implementation of this "
                        "path (no CPU fallback). Build it with `python -m simclr_b200.build`.")
                lib = ctypes.CDLL(LIB_PATH)
                _declare(lib)
                _lib = lib
    return _lib


def _check(rc: int):
    if rc != 0:
        raise NtxError(rc, load().ntx_last_error().decode("utf-8", "replace"))


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def current_stream_handle(device: torch.device) -> int:
    """cudaStream_t of the device's current stream as an integer (the raw getter is ~10x cheaper than
    building a torch.cuda.Stream object; five of these per step used to cost ~20 us of host time)."""
    if _raw_stream is not None:
        return _raw_stream(device.index if device.index is not None else torch.cuda.current_device())
    return torch.cuda.current_stream(device).cuda_stream


def _stream(t: torch.Tensor) -> int:
    return current_stream_handle(t.device)


class Backend:
    """Tensor-level view of the C ABI.  Every method enqueues on the current stream of the
    tensors' device and returns immediately."""

    name = "libntxent_b200"

    def launch_count(self) -> int:
        """Kernels the library has launched (or recorded under stream capture) so far in this process."""
        return int(load().ntx_launch_count())

    def row_scale(self, temperature: float, dtype: torch.dtype) -> float:
        """Factor to store into normalised rows of `dtype` so that the tensor-core kernels skip the per-logit scaling
        (sqrt(log2 e / temperature) for bf16 rows, 1 otherwise); goes to ``l2norm`` and into the layout tuple."""
        return float(load().ntx_row_scale(float(temperature), DTYPES[dtype]))

    def row_pitch(self, d: int, dtype: torch.dtype) -> int:
        return load().ntx_row_pitch(d, DTYPES[dtype])

    def plan(self, q_rows: int, k_rows: int, d: int, dtype: torch.dtype, sm_count: int = 0) -> NtxPlan:
        p = NtxPlan()
        _check(load().ntx_plan(q_rows, k_rows, d, DTYPES[dtype], sm_count, ctypes.byref(p)))
        return p

    def l2norm(self, view0, view1, rows, inv_norm, row_scale: float = 1.0, rows_dtype=None):
        """``rows_dtype``: what the rows tensor holds when that is not its torch dtype (``BF16X2``)."""
        n, d = view0.shape
        _check(load().ntx_l2norm(view0.data_ptr(), view1.data_ptr(), n, d, DTYPES[view0.dtype],
                                 rows.data_ptr(), rows.shape[1], DTYPES[rows_dtype or rows.dtype],
                                 inv_norm.data_ptr(), float(row_scale), _stream(rows)))

    def project_l2norm(self, h0, h1, weight, rows, inv_norm, row_scale: float = 1.0):
        """ntx_project_l2norm: rows = normalize([h0 ; h1] @ weight.T) (bf16), the projections never stored."""
        n, hdim = h0.shape
        _check(load().ntx_project_l2norm(h0.data_ptr(), h1.data_ptr(), weight.data_ptr(), n, hdim, weight.shape[0],
                                         rows.data_ptr(), rows.shape[1], inv_norm.data_ptr(), float(row_scale),
                                         _stream(rows)))

    # ---- peer exchange (include/ntxent_b200.h, "Multi-rank exchange"): `blocks` is an NtxPeerBlocks
    def peer_block_bytes(self, q_rows: int, ld: int, rows_dtype: torch.dtype, world: int) -> int:
        return int(load().ntx_peer_block_bytes(q_rows, ld, DTYPES[rows_dtype], world))

    @staticmethod
    def peer_blocks(ptrs, rank: int, multicast: int = 0) -> NtxPeerBlocks:
        pb = NtxPeerBlocks()
        pb.multicast = multicast or None
        if not 1 <= len(ptrs) <= MAX_PEERS:
            raise ValueError(f"peer exchange: world {len(ptrs)} outside [1, {MAX_PEERS}] (one node)")
        for p, addr in enumerate(ptrs):
            pb.base[p] = addr
        pb.world, pb.rank = len(ptrs), rank
        return pb

    def peer_l2norm(self, view0, view1, blocks, ld, rows_dtype, inv_norm, step, row_scale: float = 1.0):
        """ntx_peer_l2norm: normalised rows into the step's key buffer of every rank's block; ``step`` (one int32
        on the device) receives the step number the later calls of this step take."""
        n, d = view0.shape
        _check(load().ntx_peer_l2norm(view0.data_ptr(), view1.data_ptr(), n, d, DTYPES[view0.dtype],
                                      ctypes.byref(blocks), ld, DTYPES[rows_dtype], inv_norm.data_ptr(),
                                      float(row_scale), step.data_ptr(), _stream(view0)))

    def peer_fwd(self, blocks, n_local, d, ld, rows_dtype, temperature, step, rowsum, pos_logit, row_loss, loss_sum,
                 workspace, row_scale: float = 1.0):
        _check(load().ntx_peer_fwd(ctypes.byref(blocks), n_local, d, ld, DTYPES[rows_dtype], float(temperature),
                                   float(row_scale), DTYPES[loss_sum.dtype], step.data_ptr(), rowsum.data_ptr(), pos_logit.data_ptr(), row_loss.data_ptr(),
                                   loss_sum.data_ptr(), workspace.data_ptr(),
                                   workspace.numel() * workspace.element_size(), _stream(rowsum)))

    def peer_post_weights(self, blocks, n_local, ld, rows_dtype, step, w, w_stride):
        """ntx_peer_post_weights: this rank's row weights (upstream gradient) into every rank's block."""
        _check(load().ntx_peer_post_weights(ctypes.byref(blocks), n_local, ld, DTYPES[rows_dtype], step.data_ptr(),
                                            w.data_ptr(), int(w_stride), DTYPES[w.dtype], _stream(w)))

    def peer_bwd(self, blocks, n_local, d, ld, rows_dtype, temperature, step, rowsum_q, w_stride, w_scale, view0,
                 view1, inv_norm, dview0, dview1, workspace, row_scale: float = 1.0):
        _check(load().ntx_peer_bwd(ctypes.byref(blocks), n_local, d, ld, DTYPES[rows_dtype], float(temperature),
                                   float(row_scale), step.data_ptr(), rowsum_q.data_ptr(), int(w_stride), float(w_scale),
                                   view0.data_ptr(), view1.data_ptr(), DTYPES[view0.dtype], inv_norm.data_ptr(),
                                   dview0.data_ptr(), dview1.data_ptr(), workspace.data_ptr(),
                                   workspace.numel() * workspace.element_size(), _stream(view0)))

    def fwd(self, q, k, d, layout, temperature, rowsum, pos_logit, row_loss, loss_sum, workspace, rows_dtype=None):
        lay = NtxLayout(*layout)
        lay.loss_dtype = DTYPES[loss_sum.dtype]               # two floats, or two bf16 values written ready to return
        _check(load().ntx_fwd(q.data_ptr(), k.data_ptr(), q.shape[0], k.shape[0], d, q.shape[1],
                              DTYPES[rows_dtype or q.dtype], ctypes.byref(lay), float(temperature), rowsum.data_ptr(),
                              pos_logit.data_ptr(), row_loss.data_ptr(), loss_sum.data_ptr(),
                              workspace.data_ptr(), workspace.numel() * workspace.element_size(),
                              _stream(q)))

    def fwd_static(self, q, k, d, layout, temperature, rowsum, pos_logit, row_loss, loss_sum):
        """ntx_fwd_static: the forward with the library's own fixed scratch array, for callers that
        cannot provide scratch memory.  Limited to problems needing at most 1 MiB of scratch; calls
        that may overlap on the GPU race on the shared array."""
        lay = NtxLayout(*layout)
        lay.loss_dtype = DTYPES[loss_sum.dtype]
        _check(load().ntx_fwd_static(q.data_ptr(), k.data_ptr(), q.shape[0], k.shape[0], d, q.shape[1],
                                     DTYPES[q.dtype], ctypes.byref(lay), float(temperature),
                                     rowsum.data_ptr(), pos_logit.data_ptr(), row_loss.data_ptr(),
                                     loss_sum.data_ptr(), _stream(q)))

    @staticmethod
    def _weight_dtype(w_q, w_k) -> int:
        if w_q.dtype != w_k.dtype:
            raise TypeError(f"row weights of the rows and of the keys differ in dtype: {w_q.dtype} vs {w_k.dtype}")
        return DTYPES[w_q.dtype]

    def bwd(self, q, k, d, layout, temperature, rowsum_q, rowsum_k, w_q, w_k, w_stride, w_scale, view0,
            view1, inv_norm, dview0, dview1, workspace, rows_dtype=None):
        lay = NtxLayout(*layout)
        _check(load().ntx_bwd(q.data_ptr(), k.data_ptr(), q.shape[0], k.shape[0], d, q.shape[1],
                              DTYPES[rows_dtype or q.dtype], ctypes.byref(lay), float(temperature), rowsum_q.data_ptr(),
                              rowsum_k.data_ptr(), w_q.data_ptr(), w_k.data_ptr(), int(w_stride),
                              self._weight_dtype(w_q, w_k), float(w_scale), view0.data_ptr(), view1.data_ptr(),
                              DTYPES[view0.dtype],
                              inv_norm.data_ptr(), dview0.data_ptr(), dview1.data_ptr(),
                              workspace.data_ptr(), workspace.numel() * workspace.element_size(),
                              _stream(q)))

    def bwd_rows(self, q, k, d, layout, temperature, rowsum_q, rowsum_k, w_q, w_k, w_stride, w_scale, inv_norm, dz,
                 workspace):
        """ntx_bwd_rows: gradient with respect to the un-normalised projections, from the normalised rows."""
        lay = NtxLayout(*layout)
        _check(load().ntx_bwd_rows(q.data_ptr(), k.data_ptr(), q.shape[0], k.shape[0], d, q.shape[1],
                                   DTYPES[q.dtype], ctypes.byref(lay), float(temperature), rowsum_q.data_ptr(),
                                   rowsum_k.data_ptr(), w_q.data_ptr(), w_k.data_ptr(), int(w_stride),
                                   self._weight_dtype(w_q, w_k), float(w_scale), inv_norm.data_ptr(), dz.data_ptr(),
                                   workspace.data_ptr(),
                                   workspace.numel() * workspace.element_size(), _stream(q)))

    def project_bwd_workspace_bytes(self, n: int, hdim: int) -> int:
        return int(load().ntx_project_bwd_workspace_bytes(n, hdim, 0))

    def project_bwd(self, dz, h0, h1, weight, dh0, dh1, dweight, workspace):
        """ntx_project_bwd: dh = dz @ W and dW = dz.T @ [h0 ; h1] from one tcgen05 kernel (d <= 128)."""
        n, hdim = h0.shape
        _check(load().ntx_project_bwd(dz.data_ptr(), h0.data_ptr(), h1.data_ptr(), weight.data_ptr(), n, hdim,
                                      weight.shape[0], dh0.data_ptr(), dh1.data_ptr(), dweight.data_ptr(),
                                      workspace.data_ptr(), workspace.numel() * workspace.element_size(), _stream(dz)))

    def simclr_augment(self, images, strength, seed, offset, out_size, view0, view1, params_out=None):
        """ntx_simclr_augment: two augmented float views of a (B, H, W, 3) uint8 batch (dataset.py:19-50 on the GPU)."""
        b, h, w, c = images.shape
        if c != 3 or images.dtype != torch.uint8 or not images.is_contiguous():
            raise TypeError("images must be a contiguous (B, H, W, 3) uint8 tensor")
        _check(load().ntx_simclr_augment(images.data_ptr(), b, h, w, float(strength), int(seed), int(offset), int(out_size),
                                         view0.data_ptr(), view1.data_ptr(),
                                         params_out.data_ptr() if params_out is not None else None, _stream(images)))

    def larc_sgd_step(self, table_dev, n_tensors, chunks_dev, n_chunks, chunk_elems, lr, momentum, trust, eps, clip,
                      workspace):
        """ntx_larc_sgd_step: fused LARC + SGD-momentum update of every tensor in the device table."""
        _check(load().ntx_larc_sgd_step(table_dev.data_ptr(), n_tensors, chunks_dev.data_ptr(), n_chunks, chunk_elems,
                                        float(lr), float(momentum), float(trust), float(eps), 1 if clip else 0,
                                        workspace.data_ptr(), workspace.numel(), _stream(workspace)))

    def profile_next(self, which: int, ev_start: "torch.cuda.Event", ev_stop: "torch.cuda.Event"):
        """Arm the measurement hook: the next fwd (0) / bwd (1) call on this thread brackets its
        main kernel with the two events (which must already exist: record them once first)."""
        _check(load().ntx_profile_next(which, ev_start.cuda_event, ev_stop.cuda_event))

    def debug_indices(self, layout, self_col, pos_col):
        lay = NtxLayout(*layout)
        _check(load().ntx_debug_indices(ctypes.byref(lay), self_col.data_ptr(), pos_col.data_ptr(),
                                        _stream(self_col)))
