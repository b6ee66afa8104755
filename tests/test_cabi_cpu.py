"""The C-ABI library without a GPU: it loads, exports every declared symbol, and its
integer planning / argument validation behave.  No compute calls here."""
import ctypes
import os
import re

import pytest
import torch

from simclr_b200 import _cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ntxent_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ntx_[a-z0-9_]+)\s*\(", text)))


def test_library_loads_and_exports_every_declared_symbol():
    lib = _cabi.load()
    names = declared_symbols()
    assert len(names) >= 8
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/ntxent_b200.h but not exported"
    assert set(names) == set(_cabi.EXPORTS)
    assert lib.ntx_version() == 100


def test_row_pitch():
    be = _cabi.Backend()
    assert be.row_pitch(128, torch.float32) == 128
    assert be.row_pitch(20, torch.float32) == 20
    assert be.row_pitch(21, torch.float32) == 24
    assert be.row_pitch(128, torch.bfloat16) == 128
    assert be.row_pitch(96, torch.bfloat16) == 128
    assert be.row_pitch(1, torch.bfloat16) == 64
    # two-term rows (fp32 values as [hi | lo] bf16 terms): twice the bf16 pitch
    assert be.row_pitch(128, _cabi.BF16X2) == 256
    assert be.row_pitch(33, _cabi.BF16X2) == 128


def owner(T, G, t):
    return ((t + 1) * G - 1) // T


@pytest.mark.parametrize("q_rows,k_rows,d,dtype,sms", [
    (16384, 16384, 128, torch.bfloat16, 148),     # BASELINE config 4 on one GPU
    (2048, 16384, 128, torch.bfloat16, 148),      # config 4, one of 8 ranks
    (8192, 65536, 256, torch.bfloat16, 148),      # config 5, one of 8 ranks
    (512, 512, 128, torch.float32, 148),          # config 1
    (1024, 1024, 128, torch.float32, 148),
    (66, 66, 20, torch.float32, 148),
    (2, 2, 8, torch.float32, 148),
    (600, 1200, 100, torch.bfloat16, 7),
    (16384, 16384, 128, _cabi.BF16X2, 148),       # config 4 with fp32 inputs on the tensor cores
    (300, 600, 40, _cabi.BF16X2, 148),
])
def test_plan_partition_is_exact(q_rows, k_rows, d, dtype, sms):
    p = _cabi.Backend().plan(q_rows, k_rows, d, dtype, sms)
    assert p.row_blocks == -(-q_rows // p.tile_rows) and p.key_tiles == -(-k_rows // p.tile_keys)
    T, G = p.row_blocks * p.key_tiles, p.grid
    assert 1 <= G <= T
    begins = [i * T // G for i in range(G + 1)]
    assert begins[0] == 0 and begins[-1] == T and all(b1 > b0 for b0, b1 in zip(begins, begins[1:]))
    # the closed-form owner used by the finalize kernels agrees with the ranges
    step = max(1, T // 5000)
    for t in list(range(0, T, step)) + [T - 1]:
        i = owner(T, G, t)
        assert begins[i] <= t < begins[i + 1]
    slots = 1
    for rb in range(p.row_blocks):
        first, last = owner(T, G, rb * p.key_tiles), owner(T, G, rb * p.key_tiles + p.key_tiles - 1)
        slots = max(slots, last - first + 1)
    assert p.max_slots == slots
    rows_pad = p.row_blocks * p.tile_rows
    assert p.fwd_ws_bytes >= slots * rows_pad * 4 + 16      # + the finalize kernel's ticket and block sums
    assert p.bwd_ws_bytes >= slots * rows_pad * p.d_pad * 4       # + the fold counters of the tensor-core kernel
    assert p.tensor_cores == {torch.bfloat16: 1, _cabi.BF16X2: 2}.get(dtype, 0)
    if dtype is _cabi.BF16X2:                     # d_pad is the width of one term = of the dZ slots; never symmetric
        assert p.d_pad == _cabi.Backend().row_pitch(d, dtype) // 2 and p.fwd_symmetric == 0


def test_argument_errors_are_reported_without_a_device():
    lib = _cabi.load()
    lay = _cabi.NtxLayout(4, 1, 0)
    # temperature <= 0 is rejected like the reference's assert (loss.py:16)
    rc = lib.ntx_fwd(16, 16, 8, 8, 8, 8, 0, ctypes.byref(lay), 0.0, 16, 16, 16, 16, 16, 64, None)
    assert rc == -1 and b"temperature" in lib.ntx_last_error()
    # rows must be 2 * n_local
    rc = lib.ntx_fwd(16, 16, 10, 10, 8, 8, 0, ctypes.byref(lay), 0.5, 16, 16, 16, 16, 16, 64, None)
    assert rc == -1 and b"q_rows" in lib.ntx_last_error()
    rc = lib.ntx_bwd(16, 16, 8, 8, 8, 8, 0, ctypes.byref(lay), 0.5, 16, 16, 16, 16, 3, 0, 1.0, 16, 16, 0, 16, 16,
                     16, 16, 64, None)
    assert rc == -1 and b"w_stride" in lib.ntx_last_error()
    p = _cabi.NtxPlan()
    assert lib.ntx_plan(16, 16, 512, 1, 148, ctypes.byref(p)) == -4     # bf16 rows wider than a tile
    with pytest.raises(_cabi.NtxError):
        _cabi.Backend().plan(16, 8, 8, torch.float32, 148)               # fewer keys than rows


def test_newer_entry_points_validate_before_touching_the_device():
    """Peer exchange / projection / rows-backward: argument errors come back as NTX_E_ARG with a message and no
    CUDA call is made (fake, aligned, non-NULL addresses)."""
    lib = _cabi.load()
    # peer exchange: at most 8 ranks (one node), no NULL / misaligned block, rank inside the world, a step word
    be = _cabi.Backend()
    assert be.peer_block_bytes(2048, 128, torch.bfloat16, 8) >= 256 + 2 * 8 * 2048 * 4 + 2 * 8 * 2048 * 128 * 2
    assert be.peer_block_bytes(2048, 128, torch.bfloat16, 9) == 0
    bad_rank = be.peer_blocks([4096, 8192], 2)
    assert lib.ntx_peer_l2norm(4096, 4096, 4, 8, 0, ctypes.byref(bad_rank), 8, 0, 4096, 1.0, 4096, None) == -1
    assert b"rank=2" in lib.ntx_last_error()
    holed = be.peer_blocks([4096, 0], 0)
    assert lib.ntx_peer_fwd(ctypes.byref(holed), 4, 8, 8, 0, 0.5, 1.0, 0, 4096, 4096, 4096, 4096, 4096, 4096, 1 << 20, None) == -1
    assert b"rank 1" in lib.ntx_last_error()
    skew = be.peer_blocks([4096, 8192 + 16], 0)
    assert lib.ntx_peer_fwd(ctypes.byref(skew), 4, 8, 8, 0, 0.5, 1.0, 0, 4096, 4096, 4096, 4096, 4096, 4096, 1 << 20, None) == -1
    assert b"aligned" in lib.ntx_last_error()
    good = be.peer_blocks([4096, 8192], 0)
    assert lib.ntx_peer_fwd(ctypes.byref(good), 4, 8, 8, 0, 0.5, 1.0, 0, None, 4096, 4096, 4096, 4096, 4096, 1 << 20, None) == -1
    assert b"step" in lib.ntx_last_error()
    assert lib.ntx_peer_post_weights(ctypes.byref(good), 4, 8, 0, 4096, 4096, 2, 0, None) == -1
    assert b"w_stride" in lib.ntx_last_error()
    # augmentation: image staged in shared memory (<= 64 x 64), output side, strength like main.py:49
    assert lib.ntx_simclr_augment(4096, 4, 80, 32, 0.5, 0, 0, 32, 4096, 4096, None, None) == -1
    assert b"shared memory" in lib.ntx_last_error()
    assert lib.ntx_simclr_augment(4096, 4, 32, 32, 0.0, 0, 0, 32, 4096, 4096, None, None) == -1
    assert lib.ntx_simclr_augment(4096, 4, 32, 32, 0.5, 0, 0, 100, 4096, 4096, None, None) == -1
    # projection: hidden size in whole TMA blocks, row pitch of bf16 rows, d <= 256
    assert lib.ntx_project_l2norm(4096, 4096, 4096, 4, 100, 64, 4096, 64, 4096, 1.0, None) == -1
    assert b"multiple of 64" in lib.ntx_last_error()
    assert lib.ntx_project_l2norm(4096, 4096, 4096, 4, 128, 96, 4096, 96, 4096, 1.0, None) == -1
    assert b"ntx_row_pitch" in lib.ntx_last_error()
    assert lib.ntx_project_l2norm(4096, 4096, 4096, 4, 128, 300, 4096, 320, 4096, 1.0, None) == -1
    assert lib.ntx_project_l2norm(4096, None, 4096, 4, 128, 64, 4096, 64, 4096, 1.0, None) == -1
    # rows-backward: output required (odd sizes go to the unaligned finalize: no size restriction)
    lay = _cabi.NtxLayout(4, 1, 0)
    rc = lib.ntx_bwd_rows(4096, 4096, 8, 8, 8, 64, 1, ctypes.byref(lay), 0.5, 4096, 4096, 4096, 4096, 0, 0, 1.0, 4096,
                          None, 4096, 64, None)
    assert rc == -1 and b"NULL" in lib.ntx_last_error()


def test_row_scale_is_accepted_only_as_handed_out():
    """bf16 rows may carry ntx_row_scale(temperature); any other factor, or a factor on fp32 rows, is an argument
    error (the kernels would silently compute another loss)."""
    import math
    lib = _cabi.load()
    be = _cabi.Backend()
    assert be.row_scale(0.5, torch.float32) == 1.0
    assert abs(be.row_scale(0.5, torch.bfloat16) - math.sqrt(math.log2(math.e) / 0.5)) < 1e-6
    assert be.row_scale(0.01, torch.bfloat16) == 1.0                 # below the tensor-core kernels' temperature range
    lay = _cabi.NtxLayout(4, 1, 0, 1.5)
    rc = lib.ntx_fwd(4096, 4096, 8, 8, 8, 64, 1, ctypes.byref(lay), 0.5, 4096, 4096, 4096, 4096, 4096, 1 << 20, None)
    assert rc == -1 and b"row_scale" in lib.ntx_last_error()
    lay = _cabi.NtxLayout(4, 1, 0, be.row_scale(0.5, torch.bfloat16))
    rc = lib.ntx_fwd(4096, 4096, 8, 8, 8, 8, 0, ctypes.byref(lay), 0.5, 4096, 4096, 4096, 4096, 4096, 1 << 20, None)
    assert rc == -1 and b"row_scale" in lib.ntx_last_error()          # fp32 rows


def test_two_term_rows_arguments_are_validated_without_a_device():
    """NTX_BF16X2 (fp32 values as [hi | lo] bf16 terms): pitch, feature limit, temperature floor and the entry points
    that do not take the layout answer with an error code and a message before any CUDA call."""
    lib = _cabi.load()
    X2 = _cabi.NTX_BF16X2
    assert lib.ntx_row_pitch(128, X2) == 256 and lib.ntx_row_pitch(1, X2) == 128 and lib.ntx_row_pitch(0, X2) == 0
    assert abs(lib.ntx_row_scale(0.5, X2) - lib.ntx_row_scale(0.5, _cabi.NTX_BF16)) < 1e-7
    assert lib.ntx_row_scale(0.01, X2) == 1.0                              # below the fixed-shift range: unit rows
    # normalise: one term must fit a 128-wide tile, ld is exactly the two-term pitch, rows 16-byte aligned
    assert lib.ntx_l2norm(4096, 4096, 4, 200, 0, 4096, 512, X2, 4096, 1.0, None) == -1
    assert b"128" in lib.ntx_last_error()
    assert lib.ntx_l2norm(4096, 4096, 4, 100, 0, 4096, 128, X2, 4096, 1.0, None) == -1
    assert b"ntx_row_pitch" in lib.ntx_last_error()
    assert lib.ntx_l2norm(4096, 4096, 4, 100, 0, 4096 + 8, 256, X2, 4096, 1.0, None) == -1
    assert b"aligned" in lib.ntx_last_error()
    p = _cabi.NtxPlan()
    assert lib.ntx_plan(16, 16, 129, X2, 148, ctypes.byref(p)) == -4
    assert lib.ntx_plan(16, 16, 128, X2, 148, ctypes.byref(p)) == 0 and p.tensor_cores == 2 and p.d_pad == 128
    assert lib.ntx_plan(16, 16, 128, 3, 148, ctypes.byref(p)) == -1        # no such rows dtype
    lay = _cabi.NtxLayout(4, 1, 0)
    # forward: wrong pitch for the layout, feature limit, temperature below the tensor-core floor
    rc = lib.ntx_fwd(4096, 4096, 8, 8, 8, 64, X2, ctypes.byref(lay), 0.5, 4096, 4096, 4096, 4096, 4096, 1 << 20, None)
    assert rc == -1 and b"ntx_row_pitch" in lib.ntx_last_error()
    rc = lib.ntx_fwd(4096, 4096, 8, 8, 200, 512, X2, ctypes.byref(lay), 0.5, 4096, 4096, 4096, 4096, 4096, 1 << 20, None)
    assert rc == -4 and b"128" in lib.ntx_last_error()
    rc = lib.ntx_fwd(4096, 4096, 8, 8, 8, 128, X2, ctypes.byref(lay), 0.01, 4096, 4096, 4096, 4096, 4096, 1 << 20, None)
    assert rc == -4 and b"temperature" in lib.ntx_last_error()
    # the rows-backward (Jacobian from the normalised rows) takes fp32 or bf16 rows only
    rc = lib.ntx_bwd_rows(4096, 4096, 8, 8, 8, 128, X2, ctypes.byref(lay), 0.5, 4096, 4096, 4096, 4096, 0, 0, 1.0, 4096,
                          4096, 4096, 1 << 20, None)
    assert rc == -1 and b"dtype" in lib.ntx_last_error()


def test_plan_invariants_fuzzed():
    """ntx_plan over random shapes, rank counts, SM counts and all three rows layouts: the tile partition covers every
    tile once, slot counts bound the CTAs sharing a row block, the scratch sizes cover the slots."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=150, deadline=None)
    @given(n_local=st.integers(1, 3000), world=st.integers(1, 8), d=st.integers(1, 256), kind=st.integers(0, 2),
           sms=st.integers(1, 160))
    def check(n_local, world, d, kind, sms):
        dtype = (torch.float32, torch.bfloat16, _cabi.BF16X2)[kind]
        if dtype is _cabi.BF16X2 and d > 128:
            with pytest.raises(_cabi.NtxError):
                _cabi.Backend().plan(2 * n_local, 2 * n_local * world, d, dtype, sms)
            return
        q_rows, k_rows = 2 * n_local, 2 * n_local * world
        p = _cabi.Backend().plan(q_rows, k_rows, d, dtype, sms)
        assert p.row_blocks == -(-q_rows // p.tile_rows) and p.key_tiles == -(-k_rows // p.tile_keys)
        T, G = p.row_blocks * p.key_tiles, p.grid
        assert 1 <= G <= min(T, sms * (1 if p.tensor_cores else 2))
        slots = max(owner(T, G, rb * p.key_tiles + p.key_tiles - 1) - owner(T, G, rb * p.key_tiles) + 1
                    for rb in range(p.row_blocks))
        assert p.max_slots == slots
        rows_pad = p.row_blocks * p.tile_rows
        assert p.fwd_ws_bytes >= p.fwd_ws_bytes_min >= slots * rows_pad * 4
        assert p.bwd_ws_bytes >= slots * rows_pad * p.d_pad * 4
        assert p.d_pad >= d and (p.fwd_symmetric == 0 or (p.tensor_cores == 1 and world == 1 and p.d_pad <= 128))

    check()
