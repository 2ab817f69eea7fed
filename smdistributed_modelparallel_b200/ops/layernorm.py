"""Fused LayerNorm (optionally fused with the residual add) on the sm_100a kernels.

Same role as the reference's Apex ``FusedLayerNorm`` / ``MixedFusedLayerNorm`` (reference
smdistributed/modelparallel/torch/apex/normalization/fused_layer_norm.py:26-78) and, through
``layer_norm_with_stats`` / the ``*_local_sums`` helpers, its hidden-sharded
``DistributedLayerNormForTP`` (reference torch/nn/layer_norm.py:24-102).
"""

from __future__ import annotations

import torch
import torch.nn.functional as F

from . import _native as N
from .workspace import get_workspace


def _ok(*ts):
    return all(t is None or (t.is_cuda and t.is_contiguous() and t.dtype in N._DTYPE_CODE) for t in ts)


class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, weight, bias, eps, out_dtype):
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        r2 = residual.reshape(-1, H) if residual is not None else None
        rows = x2.shape[0]
        y = torch.empty(rows, H, dtype=out_dtype or x.dtype, device=x.device)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        s = torch.empty_like(x2) if r2 is not None else None
        N.count_launch()
        N.call("smpk_layernorm_fwd", N.stream_ptr(x.device), x2.data_ptr(), N.ptr(r2), N.ptr(s),
               N.dtype_code(x.dtype), N.ptr(weight), N.ptr(bias),
               N.dtype_code(weight.dtype) if weight is not None else 0, y.data_ptr(), N.dtype_code(y.dtype),
               mean.data_ptr(), rstd.data_ptr(), rows, H, float(eps))
        xin = s if s is not None else x2
        ctx.save_for_backward(xin, weight, mean, rstd)
        ctx.has_bias = bias is not None
        ctx.has_residual = residual is not None
        ctx.shape = x.shape
        y = y.reshape(x.shape)
        if residual is not None:
            return y, s.reshape(x.shape)
        return y

    @staticmethod
    def backward(ctx, dy, dsum=None):
        xin, weight, mean, rstd = ctx.saved_tensors
        H = xin.shape[-1]
        rows = xin.shape[0]
        dy2 = dy.reshape(-1, H)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = torch.empty_like(xin)
        want_w = weight is not None and (ctx.needs_input_grad[2] or ctx.needs_input_grad[3])
        dgamma = torch.empty_like(weight) if want_w else None
        dbeta = torch.empty_like(weight) if (want_w and ctx.has_bias) else None
        scratch = None
        if want_w:
            grid = N.query("smpk_ln_bwd_grid", rows, H)
            scratch = get_workspace(xin.device, "ln_bwd", 2 * grid * H * 4)
        add = None
        if ctx.has_residual and dsum is not None:
            add = dsum.reshape(-1, H)
            if not add.is_contiguous() or add.dtype != dx.dtype:
                add = add.contiguous().to(dx.dtype)
        N.count_launch(2 if want_w else 1)
        N.call("smpk_layernorm_bwd", N.stream_ptr(xin.device), dy2.data_ptr(), N.dtype_code(dy2.dtype), xin.data_ptr(),
               N.dtype_code(xin.dtype), mean.data_ptr(), rstd.data_ptr(), N.ptr(weight),
               N.dtype_code(weight.dtype) if weight is not None else 0, dx.data_ptr(), N.dtype_code(dx.dtype),
               N.ptr(add), N.ptr(dgamma), N.ptr(dbeta), 0, N.ptr(scratch), rows, H)
        dx = dx.reshape(ctx.shape)
        return dx, (dx if ctx.has_residual else None), dgamma, dbeta, None, None


def layer_norm(x, weight=None, bias=None, eps: float = 1e-5, residual=None, out_dtype=None):
    """LayerNorm over the last dimension.  With ``residual`` returns ``(LN(x+residual), x+residual)``."""
    if x.is_cuda and _ok(x, residual, weight, bias):
        return _LayerNormFn.apply(x, residual, weight, bias, eps, out_dtype)
    if x.is_cuda and not (x.is_contiguous() and (residual is None or residual.is_contiguous())):
        return layer_norm(x.contiguous(), weight, bias, eps, None if residual is None else residual.contiguous(),
                          out_dtype)
    s = x if residual is None else x + residual
    w = weight.float() if weight is not None else None
    b = bias.float() if bias is not None else None
    y = F.layer_norm(s.float(), (x.shape[-1],), w, b, eps).to(out_dtype or x.dtype)
    return (y, s) if residual is not None else y


# ---- statistics-supplied phases (hidden-sharded LayerNorm across tensor-parallel ranks) --------
def partial_stats(x2: torch.Tensor) -> torch.Tensor:
    """[rows, 2] fp32: per-row (sum, sum of squares) of the local hidden shard."""
    rows, H = x2.shape
    if not (x2.is_cuda and x2.is_contiguous()):
        xf = x2.float()
        return torch.stack([xf.sum(-1), (xf * xf).sum(-1)], dim=-1)
    stats = torch.empty(rows, 2, dtype=torch.float32, device=x2.device)
    N.count_launch()
    N.call("smpk_ln_partial_stats", N.stream_ptr(x2.device), x2.data_ptr(), N.dtype_code(x2.dtype), stats.data_ptr(),
           rows, H)
    return stats


def apply_stats(x2, mean, rstd, weight, bias, out_dtype=None):
    rows, H = x2.shape
    if not (x2.is_cuda and x2.is_contiguous()):
        y = (x2.float() - mean[:, None]) * rstd[:, None]
        if weight is not None:
            y = y * weight.float()
        if bias is not None:
            y = y + bias.float()
        return y.to(out_dtype or x2.dtype)
    y = torch.empty(rows, H, dtype=out_dtype or x2.dtype, device=x2.device)
    N.count_launch()
    N.call("smpk_ln_apply_stats", N.stream_ptr(x2.device), x2.data_ptr(), N.dtype_code(x2.dtype), mean.data_ptr(),
           rstd.data_ptr(), N.ptr(weight), N.ptr(bias), N.dtype_code(weight.dtype) if weight is not None else 0,
           y.data_ptr(), N.dtype_code(y.dtype), rows, H)
    return y


def bwd_local_sums(dy2, x2, mean, rstd, weight):
    rows, H = x2.shape
    if not (x2.is_cuda and x2.is_contiguous() and dy2.is_contiguous()):
        xh = (x2.float() - mean[:, None]) * rstd[:, None]
        dg = dy2.float() * (weight.float() if weight is not None else 1.0)
        return torch.stack([dg.sum(-1), (dg * xh).sum(-1)], dim=-1)
    sums = torch.empty(rows, 2, dtype=torch.float32, device=x2.device)
    N.count_launch()
    N.call("smpk_ln_bwd_local_sums", N.stream_ptr(x2.device), dy2.data_ptr(), N.dtype_code(dy2.dtype), x2.data_ptr(),
           N.dtype_code(x2.dtype), mean.data_ptr(), rstd.data_ptr(), N.ptr(weight),
           N.dtype_code(weight.dtype) if weight is not None else 0, sums.data_ptr(), rows, H)
    return sums


def bwd_finish(dy2, x2, mean, rstd, weight, sums, h_total):
    rows, H = x2.shape
    if not (x2.is_cuda and x2.is_contiguous() and dy2.is_contiguous()):
        xh = (x2.float() - mean[:, None]) * rstd[:, None]
        dg = dy2.float() * (weight.float() if weight is not None else 1.0)
        dx = rstd[:, None] * (dg - (sums[:, :1] + xh * sums[:, 1:2]) / h_total)
        return dx.to(x2.dtype)
    dx = torch.empty_like(x2)
    N.count_launch()
    N.call("smpk_ln_bwd_finish", N.stream_ptr(x2.device), dy2.data_ptr(), N.dtype_code(dy2.dtype), x2.data_ptr(),
           N.dtype_code(x2.dtype), mean.data_ptr(), rstd.data_ptr(), N.ptr(weight),
           N.dtype_code(weight.dtype) if weight is not None else 0, sums.data_ptr(), dx.data_ptr(),
           N.dtype_code(dx.dtype), rows, H, int(h_total))
    return dx
