"""LayerNorm modules: ``FusedLayerNorm`` (token-local, native kernel) and ``DistributedLayerNorm``
(hidden dimension sharded across tensor-parallel ranks, statistics reduced across the group).

Reference: Apex ``FusedLayerNorm`` / ``MixedFusedLayerNorm`` vendored at smdistributed/modelparallel/
torch/apex/normalization/fused_layer_norm.py, and ``DistributedLayerNormForTP`` at
torch/nn/layer_norm.py:24-152 (two all-reduces in forward, two in backward).  The
distributed variant here packs each pair of all-reduces into one ([rows, 2] tensor).
"""

from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn

from ..ops import layernorm as LN
from .dist_module import DistributedModule
from .utils import get_local_channels, initialize_with_output_partition, tp_group, tp_size


class FusedLayerNorm(nn.Module):
    """Drop-in LayerNorm over the last dimension running the fused sm_100a kernel on CUDA tensors.

    ``mixed_dtype=True`` mirrors Apex's ``MixedFusedLayerNorm``: the input may be fp32 (fp32 residual
    stream) while the output takes the parameter dtype.
    """

    def __init__(self, normalized_shape, eps: float = 1e-5, elementwise_affine: bool = True, mixed_dtype: bool = False):
        super().__init__()
        if isinstance(normalized_shape, int):
            normalized_shape = (normalized_shape,)
        assert len(normalized_shape) == 1, "only last-dimension normalisation is supported"
        self.normalized_shape = tuple(normalized_shape)
        self.eps = eps
        self.mixed_dtype = mixed_dtype
        if elementwise_affine:
            self.weight = nn.Parameter(torch.ones(*normalized_shape))
            self.bias = nn.Parameter(torch.zeros(*normalized_shape))
        else:
            self.register_parameter("weight", None)
            self.register_parameter("bias", None)

    def reset_parameters(self):
        if self.weight is not None:
            nn.init.ones_(self.weight)
            nn.init.zeros_(self.bias)

    def forward(self, x, residual=None):
        out_dtype = self.weight.dtype if (self.mixed_dtype and self.weight is not None) else None
        return LN.layer_norm(x, self.weight, self.bias, self.eps, residual=residual, out_dtype=out_dtype)

    def extra_repr(self):
        return f"{self.normalized_shape}, eps={self.eps}"


MixedFusedLayerNorm = FusedLayerNorm


class _DistLN(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps, h_total, group):
        H = x.shape[-1]
        x2 = x.reshape(-1, H).contiguous()
        stats = LN.partial_stats(x2)                                  # [rows, 2] local sum, sumsq
        if group is not None and dist.get_world_size(group) > 1:
            dist.all_reduce(stats, group=group)
        mean = (stats[:, 0] / h_total).contiguous()
        var = (stats[:, 1] / h_total - mean * mean).clamp_min(0.0)
        rstd = torch.rsqrt(var + eps).contiguous()
        y = LN.apply_stats(x2, mean, rstd, weight, bias)
        ctx.save_for_backward(x2, weight, mean, rstd)
        ctx.h_total, ctx.group, ctx.shape, ctx.has_bias = h_total, group, x.shape, bias is not None
        return y.reshape(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, weight, mean, rstd = ctx.saved_tensors
        dy2 = dy.reshape(x2.shape).contiguous()
        sums = LN.bwd_local_sums(dy2, x2, mean, rstd, weight)          # [rows, 2] local s1, s2
        if ctx.group is not None and dist.get_world_size(ctx.group) > 1:
            dist.all_reduce(sums, group=ctx.group)
        dx = LN.bwd_finish(dy2, x2, mean, rstd, weight, sums, ctx.h_total)
        dw = db = None
        if weight is not None:
            xh = (x2.float() - mean[:, None]) * rstd[:, None]
            dw = (dy2.float() * xh).sum(0).to(weight.dtype)
            if ctx.has_bias:
                db = dy2.float().sum(0).to(weight.dtype)
        return dx.reshape(ctx.shape), dw, db, None, None, None


class DistributedLayerNorm(DistributedModule):
    """LayerNorm whose hidden dimension is split across the tensor-parallel group (``optimize=memory``
    activations ``[B, S, H/t]``).  Parameters hold the local ``H/t`` slice."""

    def __init__(self, hidden_size=None, eps: float = 1e-5, normalized_shape=None):
        super().__init__()
        if hidden_size is None:                       # the reference's argument name (nn/layer_norm.py:108)
            hidden_size = normalized_shape
        if isinstance(hidden_size, (tuple, list, torch.Size)):
            if len(hidden_size) != 1:
                raise ValueError("DistributedLayerNorm normalises over the last (hidden) dimension only")
            hidden_size = int(hidden_size[0])
        self.hidden_size = hidden_size
        self.normalized_shape = torch.Size([get_local_channels(hidden_size)])
        self.orig_last_dim = hidden_size
        self.local_hidden_size = get_local_channels(hidden_size)
        self.eps = eps
        with initialize_with_output_partition(self):
            self.weight = nn.Parameter(torch.ones(self.local_hidden_size))
            self.bias = nn.Parameter(torch.zeros(self.local_hidden_size))

    def reset_parameters(self):
        nn.init.ones_(self.weight)
        nn.init.zeros_(self.bias)

    def forward(self, x):
        group = tp_group() if tp_size() > 1 else None
        return _DistLN.apply(x, self.weight, self.bias, self.eps, self.hidden_size, group)
