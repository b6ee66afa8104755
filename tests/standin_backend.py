"""Test-only stand-in for simclr_b200._cabi.Backend on CPU tensors (torch, float64 inside).

It exists so the AUTOGRAD AND COLLECTIVE WIRING of simclr_b200.loss (gather of keys, gather of
row sums / row weights, per-rank scaling) can be exercised with world_size > 1 on the gloo
backend in a container without GPUs.  It is monkeypatched in by the tests; the product never
imports it and has no CPU path.  Same method names and argument meaning as the real Backend.
"""
import types

import torch


class StandInBackend:
    name = "standin"

    @staticmethod
    def _two_term(dtype):
        from simclr_b200 import _cabi
        return dtype is _cabi.BF16X2

    def row_pitch(self, d, dtype):
        if self._two_term(dtype):                      # fp32 values as [hi | lo] bf16 terms: twice the bf16 pitch
            return 2 * ((d + 63) // 64 * 64)
        g = 64 if dtype == torch.bfloat16 else 4
        return (d + g - 1) // g * g

    @staticmethod
    def _values(rows, rows_dtype):
        """The fp64 values a row buffer stands for (two-term rows: hi + lo)."""
        if rows_dtype is None:
            return rows.double()
        dh = rows.shape[1] // 2
        return rows[:, :dh].double() + rows[:, dh:].double()

    def plan(self, q_rows, k_rows, d, dtype, sm_count=0):
        return types.SimpleNamespace(fwd_ws_bytes=16, bwd_ws_bytes=16)

    def row_scale(self, temperature, dtype):
        return 1.0              # unit rows: the scale only matters to the tensor-core kernels

    def l2norm(self, view0, view1, rows, inv_norm, row_scale=1.0, rows_dtype=None):
        assert row_scale == 1.0
        z = torch.cat([view0, view1]).double()
        denom = z.norm(dim=1).clamp_min(1e-12)
        rows.zero_()
        x = (z / denom[:, None]).float()
        if rows_dtype is None:
            rows[:, :z.shape[1]] = x.to(rows.dtype)
        else:                                          # two bf16 terms of the fp32 value, as the split kernel writes them
            assert self._two_term(rows_dtype) and rows.dtype == torch.bfloat16
            dh = rows.shape[1] // 2
            hi = x.bfloat16()
            rows[:, :z.shape[1]] = hi
            rows[:, dh:dh + z.shape[1]] = (x - hi.float()).bfloat16()
        inv_norm.copy_((1.0 / denom).float())

    @staticmethod
    def _cols(layout, q_rows):
        n, world, rank = layout[:3]
        r = torch.arange(q_rows)
        off = rank * q_rows
        return off + r, off + (r + n) % q_rows

    def fwd(self, q, k, d, layout, temperature, rowsum, pos_logit, row_loss, loss_sum, workspace, rows_dtype=None):
        q, k = self._values(q, rows_dtype), self._values(k, rows_dtype)
        s = (q @ k.t()) / temperature
        self_col, pos_col = self._cols(layout, q.shape[0])
        r = torch.arange(q.shape[0])
        pos = s[r, pos_col]
        e = torch.exp(s - 1.0 / temperature)
        e[r, self_col] = 0.0
        R = e.sum(dim=1)
        loss = torch.log(R) + 1.0 / temperature - pos
        rowsum.copy_(R.float()); pos_logit.copy_(pos.float()); row_loss.copy_(loss.float())
        loss_sum[0] = float(loss.sum()); loss_sum[1] = float(loss.mean())

    def bwd(self, q, k, d, layout, temperature, rowsum_q, rowsum_k, w_q, w_k, w_stride, w_scale, view0,
            view1, inv_norm, dview0, dview1, workspace, rows_dtype=None):
        q, k = self._values(q, rows_dtype), self._values(k, rows_dtype)
        n = view0.shape[0]
        q_rows = 2 * n
        wq = (w_q.double() * w_scale).expand(q_rows) if w_stride == 0 else w_q.double() * w_scale
        wk = (w_k.double() * w_scale).expand(k.shape[0]) if w_stride == 0 else w_k.double() * w_scale
        alpha, beta = wq / rowsum_q.double(), wk / rowsum_k.double()
        s = (q.double() @ k.double().t()) / temperature
        self_col, pos_col = self._cols(layout, q_rows)
        r = torch.arange(q_rows)
        e = torch.exp(s - 1.0 / temperature)
        e[r, self_col] = 0.0
        c = e * (alpha[:, None] + beta[None, :])
        z = torch.cat([view0, view1]).double()
        zh = z * inv_norm.double()[:, None]
        p = (r + n) % q_rows
        dzh = (c @ k.double()[:, :z.shape[1]] - (wq + wq[p])[:, None] * zh[p]) / temperature
        proj = (zh * dzh).sum(dim=1, keepdim=True)
        dz = (dzh - zh * proj) * inv_norm.double()[:, None]
        dview0.copy_(dz[:n].to(dview0.dtype)); dview1.copy_(dz[n:].to(dview1.dtype))
