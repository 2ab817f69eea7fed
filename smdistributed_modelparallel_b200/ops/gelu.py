"""bias + GELU (tanh or erf form) as a standalone op; inside the MLP it is a GEMM epilogue instead.

Parity: reference smdistributed/modelparallel/torch/nn/gelu.py:28-64 (TorchScript ``bias_gelu`` /
``bias_gelu_back`` and the ``GeLUFunction`` wrapper).
"""

from __future__ import annotations

import torch
import torch.nn.functional as F

from . import _native as N

_MODE = {"tanh": 0, "erf": 1}


def _run(x, bias, dy, mode, backward):
    n = x.shape[-1]
    rows = x.numel() // n
    out = torch.empty_like(x)
    N.count_launch()
    N.call("smpk_bias_gelu", N.stream_ptr(x.device), x.data_ptr(), N.ptr(bias),
           N.dtype_code(bias.dtype) if bias is not None else 0, N.ptr(dy), out.data_ptr(), N.dtype_code(x.dtype),
           rows, n, _MODE[mode], int(backward))
    return out


def gelu_backward(dy: torch.Tensor, pre: torch.Tensor, mode: str = "tanh", bias=None) -> torch.Tensor:
    """dy * gelu'(pre + bias)."""
    if pre.is_cuda and pre.is_contiguous() and dy.is_contiguous() and pre.dtype in N._DTYPE_CODE:
        return _run(pre, bias, dy.to(pre.dtype), mode, True)
    z = pre.detach().float() + (bias.float() if bias is not None else 0.0)
    z.requires_grad_(True)
    with torch.enable_grad():
        y = F.gelu(z, approximate="tanh" if mode == "tanh" else "none")
        (g,) = torch.autograd.grad(y.sum(), z)
    return (dy.float() * g).to(pre.dtype)


class _BiasGeluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bias, mode):
        ctx.mode = mode
        ctx.save_for_backward(x, bias)
        if x.is_cuda and x.is_contiguous() and x.dtype in N._DTYPE_CODE:
            return _run(x, bias, None, mode, False)
        z = x.float() + (bias.float() if bias is not None else 0.0)
        return F.gelu(z, approximate="tanh" if mode == "tanh" else "none").to(x.dtype)

    @staticmethod
    def backward(ctx, dy):
        x, bias = ctx.saved_tensors
        dx = gelu_backward(dy.contiguous(), x, ctx.mode, bias)
        db = None
        if bias is not None and ctx.needs_input_grad[1]:
            from .gemm import colsum
            db = colsum(dx.reshape(-1, dx.shape[-1]), out_dtype=bias.dtype)
        return dx, db, None


def bias_gelu(x: torch.Tensor, bias: torch.Tensor | None = None, mode: str = "tanh") -> torch.Tensor:
    return _BiasGeluFn.apply(x, bias, mode)
