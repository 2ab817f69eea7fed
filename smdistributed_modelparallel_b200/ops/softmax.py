"""Scaled causal / masked softmax on the sm_100a kernels.

Counterparts of the reference's ``ScaledUpperTriangMaskedSoftmax`` and ``ScaledMaskedSoftmax``
autograd functions over its CUDA library (reference smdistributed/modelparallel/torch/nn/
softmax.py:40-93); the 16 < sk <= 2048 restriction of the reference kernels (reference
torch/nn/transformer.py:101) does not apply.
"""

from __future__ import annotations

import torch

from . import _native as N

PLAIN, CAUSAL, MASKED = 0, 1, 2


def _ref_forward(x, mask, mode, scale):
    xf = x.float() * scale
    sq, sk = x.shape[-2], x.shape[-1]
    if mode == CAUSAL:
        keep = torch.ones(sq, sk, dtype=torch.bool, device=x.device).tril(diagonal=sk - sq)
        xf = xf.masked_fill(~keep, float("-inf"))
    elif mode == MASKED:
        xf = xf.masked_fill(mask.bool(), float("-inf"))
    y = torch.softmax(xf, dim=-1)
    return torch.nan_to_num(y, nan=0.0).to(x.dtype)


class _SoftmaxFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, mask, mode, scale):
        ctx.mode, ctx.scale = mode, scale
        if x.is_cuda:
            x = x.contiguous()
            sq, sk = x.shape[-2], x.shape[-1]
            bh = x.numel() // (sq * sk)
            heads = x.shape[1] if (mode == MASKED and x.dim() == 4) else 1
            y = torch.empty_like(x)
            m = mask.contiguous() if mask is not None else None
            N.count_launch()
            N.call("smpk_softmax_fwd", N.stream_ptr(x.device), x.data_ptr(), y.data_ptr(), N.dtype_code(x.dtype),
                   N.ptr(m), mode, float(scale), bh, heads, sq, sk)
        else:
            y = _ref_forward(x, mask, mode, scale)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        if y.is_cuda:
            dy = dy.contiguous()
            sq, sk = y.shape[-2], y.shape[-1]
            bh = y.numel() // (sq * sk)
            dx = torch.empty_like(y)
            N.count_launch()
            N.call("smpk_softmax_bwd", N.stream_ptr(y.device), dy.data_ptr(), y.data_ptr(), dx.data_ptr(),
                   N.dtype_code(y.dtype), int(ctx.mode == CAUSAL), float(ctx.scale), bh, sq, sk)
        else:
            yf, df = y.float(), dy.float()
            dx = (ctx.scale * yf * (df - (df * yf).sum(-1, keepdim=True))).to(y.dtype)
        return dx, None, None, None


def scaled_upper_triang_masked_softmax(x: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
    """softmax(scale * x + causal mask) over the last dim; x is [batch*heads, sq, sk] (or 4-D)."""
    return _SoftmaxFn.apply(x, None, CAUSAL, scale)


def scaled_masked_softmax(x: torch.Tensor, mask: torch.Tensor | None, scale: float = 1.0) -> torch.Tensor:
    """softmax(scale * x) with positions where ``mask`` is non-zero excluded; x [b, heads, sq, sk], mask [b,1,sq,sk]."""
    if mask is None:
        return _SoftmaxFn.apply(x, None, PLAIN, scale)
    if mask.dtype != torch.uint8:
        mask = mask.to(torch.uint8)
    if mask.dim() == 4 and tuple(mask.shape) != (x.shape[0], 1, x.shape[2], x.shape[3]):
        mask = mask.expand(x.shape[0], 1, x.shape[2], x.shape[3])
    return _SoftmaxFn.apply(x, mask, MASKED, scale)


def get_batch_per_block(sq: int, sk: int, b: int, np_: int) -> int:
    return N.query("smpk_get_batch_per_block", sq, sk, b, np_)
