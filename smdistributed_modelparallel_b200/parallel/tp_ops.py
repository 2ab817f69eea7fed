"""Column- and row-parallel linear layers in the token-parallel ("sequence-parallel") layout.

Between the GEMMs of a transformer block every tp rank holds only its own tokens ``[T/t, H]``:

* column-parallel (QKV, MLP dense1, LM head): all-gather tokens -> ``[T, H]`` x ``W_shard[N/t, H]^T``
* row-parallel (attention dense, MLP dense2): ``[T, K/t]`` x ``W_shard[N, K/t]^T`` -> partial sums,
  reduce-scatter over tokens -> ``[T/t, N]``

This is the layout the reference only reaches at its edges - it all-gathers the batch at layer-stack
entry, all-reduces after every row-parallel GEMM and narrows at exit (reference
smdistributed/modelparallel/torch/nn/transformer.py:1436,1474,1522,1141,1544).  Here
each rank's local batch *is* its token shard, LayerNorm / dropout / residual run on 1/t of the tokens,
and every all-reduce becomes an all-gather before or a reduce-scatter after a GEMM - the two shapes
the fused kernels (``parallel/fused_tp.py``) implement over NVSwitch peer memory.  ``comm="nccl"`` is
the plain NCCL + GEMM composition (the timed baseline); ``comm="fused"`` uses the fused kernels.
"""

from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops.gelu import gelu_backward
from ..ops.gemm import colsum, gemm

_comm_mode = "nccl"


def set_comm_mode(mode: str) -> None:
    global _comm_mode
    assert mode in ("nccl", "fused")
    _comm_mode = mode


def get_comm_mode() -> str:
    return _comm_mode


def select_comm_mode(cfg, cuda_available: bool) -> str:
    """The mode ``smp.init`` installs: 'fused' unless the config key ``fused_tp_collectives`` is false, the
    environment says ``SMP_TP_COMM=nccl`` or there is no CUDA device."""
    import os

    want = bool(getattr(cfg, "fused_tp_collectives", True)) and os.environ.get("SMP_TP_COMM", "fused") != "nccl"
    return "fused" if (want and cuda_available) else "nccl"


def _group_size(group) -> int:
    return 1 if group is None else dist.get_world_size(group)


def gather_tokens(x: torch.Tensor, group) -> torch.Tensor:
    """[T/t, ...] -> [T, ...] (rank-major order)."""
    n = _group_size(group)
    if n == 1:
        return x
    x = x.contiguous()
    out = torch.empty((x.shape[0] * n,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, x, group=group)
    return out


def scatter_sum_tokens(x: torch.Tensor, group) -> torch.Tensor:
    """[T, ...] partial sums -> [T/t, ...] summed over ranks."""
    n = _group_size(group)
    if n == 1:
        return x
    x = x.contiguous()
    assert x.shape[0] % n == 0, "token count must be divisible by the tensor-parallel degree"
    out = torch.empty((x.shape[0] // n,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.reduce_scatter_tensor(out, x, group=group)
    return out


class GatherTokens(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return gather_tokens(x, group)

    @staticmethod
    def backward(ctx, g):
        return scatter_sum_tokens(g, ctx.group), None


class ScatterSumTokens(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return scatter_sum_tokens(x, group)

    @staticmethod
    def backward(ctx, g):
        return gather_tokens(g, ctx.group), None


class NarrowTokens(torch.autograd.Function):
    """Keep the local token shard of a tensor that is identical on all tp ranks; backward all-gathers."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        n, r = _group_size(group), (dist.get_rank(group) if group is not None else 0)
        per = x.shape[0] // n
        return x.narrow(0, r * per, per).contiguous()

    @staticmethod
    def backward(ctx, g):
        return gather_tokens(g, ctx.group), None


def _wgrad(dy, x, weight):
    """dW = dy^T x, accumulated into weight.main_grad when the reducer installed one."""
    mg = getattr(weight, "main_grad", None)
    if mg is not None:
        gemm(dy, x, a_mn_major=True, b_mn_major=True, out=mg, residual=mg, beta=1.0)
        return None
    return gemm(dy, x, a_mn_major=True, b_mn_major=True)


def _bgrad(dy, bias):
    mg = getattr(bias, "main_grad", None)
    if mg is not None:
        colsum(dy, out=mg, accumulate=True)
        return None
    return colsum(dy, out_dtype=bias.dtype)


class ColumnParallelLinear(torch.autograd.Function):
    """x_local [T/t, K] --all-gather--> [T, K] @ W[N_local, K]^T (+ bias, + optional GELU) -> [T, N_local]."""

    @staticmethod
    def forward(ctx, x_local, weight, bias, activation, group, gathered_input):
        ctx.group, ctx.activation = group, activation
        ctx.gathered_input = gathered_input
        X = x_local if gathered_input else _ag(x_local, group)
        need_grad = any(ctx.needs_input_grad[:3])
        if activation is None:
            y = gemm(X, weight, bias=bias)
            pre = None
        else:
            y = torch.empty(X.shape[0], weight.shape[0], dtype=X.dtype, device=X.device)
            pre = torch.empty_like(y) if need_grad else None
            gemm(X, weight, bias=bias, out=y, epilogue=f"gelu_{activation}", aux_out=pre)
        ctx.save_for_backward(x_local, weight, bias, pre)
        return y

    @staticmethod
    def backward(ctx, dy):
        x_local, weight, bias, pre = ctx.saved_tensors
        dy = dy if dy.is_contiguous() else dy.contiguous()
        if ctx.activation is not None:
            dy = gelu_backward(dy, pre, ctx.activation)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dX = gemm(dy, weight, b_mn_major=True)
            dx = dX if ctx.gathered_input else _rs(dX, ctx.group)
        if ctx.needs_input_grad[1]:
            X = x_local if ctx.gathered_input else _ag(x_local, ctx.group)
            dw = _wgrad(dy, X, weight)
        if bias is not None and ctx.needs_input_grad[2]:
            db = _bgrad(dy, bias)
        return dx, dw, db, None, None, None


class RowParallelLinear(torch.autograd.Function):
    """x [T, K_local] @ W[N, K_local]^T (+ bias on the bias-owning rank) --reduce-scatter--> [T/t, N]."""

    @staticmethod
    def forward(ctx, x, weight, bias, group, scatter_output):
        ctx.group, ctx.scatter_output = group, scatter_output
        y = gemm(x, weight, bias=bias)
        ctx.save_for_backward(x, weight, bias)
        return _rs(y, group) if scatter_output else y

    @staticmethod
    def backward(ctx, dy_local):
        x, weight, bias = ctx.saved_tensors
        dy_local = dy_local if dy_local.is_contiguous() else dy_local.contiguous()
        dY = _ag(dy_local, ctx.group) if ctx.scatter_output else dy_local
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dY, weight, b_mn_major=True)
        if ctx.needs_input_grad[1]:
            dw = _wgrad(dY, x, weight)
        if bias is not None and ctx.needs_input_grad[2]:
            db = _bgrad(dY, bias)
        return dx, dw, db, None, None


class FusedMLP(torch.autograd.Function):
    """y = gelu(x W1^T + b1) W2^T + b2 on one rank (tensor-parallel degree 1), four GEMM launches forward + backward
    plus wgrads: GELU rides the first GEMM's epilogue, an optional ``residual`` [T, H] is added by the second GEMM's
    epilogue, and in backward GELU' rides the epilogue of the dgrad GEMM through W2 (``(dy W2) * gelu'(pre)``), so no elementwise pass over the [T, 4H] intermediate remains."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, activation, residual=None):
        ctx.activation = activation
        ctx.has_residual = residual is not None
        inter = torch.empty(x.shape[0], w1.shape[0], dtype=x.dtype, device=x.device)
        need = any(ctx.needs_input_grad[:5])
        pre = torch.empty_like(inter) if need else None
        gemm(x, w1, bias=b1, out=inter, epilogue=f"gelu_{activation}", aux_out=pre)
        y = gemm(inter, w2, bias=b2, residual=residual, beta=1.0)      # + residual in the epilogue when given
        ctx.save_for_backward(x, w1, b1, w2, b2, pre, inter)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w1, b1, w2, b2, pre, inter = ctx.saved_tensors
        dy = dy if dy.is_contiguous() else dy.contiguous()
        dz = gemm(dy, w2, b_mn_major=True, residual=pre, epilogue=f"dgelu_{ctx.activation}")   # d(pre-activation)
        dw2 = _wgrad(dy, inter, w2) if ctx.needs_input_grad[3] else None
        db2 = _bgrad(dy, b2) if (b2 is not None and ctx.needs_input_grad[4]) else None
        dx = gemm(dz, w1, b_mn_major=True) if ctx.needs_input_grad[0] else None
        dw1 = _wgrad(dz, x, w1) if ctx.needs_input_grad[1] else None
        db1 = _bgrad(dz, b1) if (b1 is not None and ctx.needs_input_grad[2]) else None
        return dx, dw1, db1, dw2, db2, None, (dy if ctx.has_residual else None)


def fused_mlp_applicable(x, w1, group) -> bool:
    return _group_size(group) == 1 and x.is_cuda and x.dtype in (torch.bfloat16, torch.float16) and w1.dtype == x.dtype


def _ag(x, group):
    if _group_size(group) == 1:
        return x
    if _comm_mode == "fused" and x.is_cuda:
        from . import fused_tp
        if fused_tp.available(group):
            return fused_tp.all_gather_tokens(x, group)
    return gather_tokens(x, group)


def _rs(x, group):
    if _group_size(group) == 1:
        return x
    if _comm_mode == "fused" and x.is_cuda:
        from . import fused_tp
        if fused_tp.available(group):
            return fused_tp.reduce_scatter_tokens(x, group)
    return scatter_sum_tokens(x, group)


def column_parallel_linear(x_local, weight, bias=None, activation=None, group=None, gathered_input=False):
    return ColumnParallelLinear.apply(x_local, weight, bias, activation, group, gathered_input)


def row_parallel_linear(x, weight, bias=None, group=None, scatter_output=True):
    return RowParallelLinear.apply(x, weight, bias, group, scatter_output)


# ------------------------------------------------------------------------------------------------
# Layer-level entry points: pick the fused peer-memory kernels when they apply, else NCCL + GEMM.
def _fused_applicable(rows_local, *tensors):
    if _comm_mode != "fused":
        return False
    from . import fused_tp
    return fused_tp.available() and fused_tp.usable(rows_local, *tensors)


def ag_linear(x_local, weight, bias=None, activation=None, group=None, want_gathered=False):
    """Column-parallel linear on token-sharded input: y[T, N/t] = act(allgather(x_local) W^T + b).
    With ``want_gathered`` also returns the gathered input (for a second GEMM on the same activations)."""
    if _group_size(group) == 1:
        y = ColumnParallelLinear.apply(x_local, weight, bias, activation, None, True)
        return (y, x_local) if want_gathered else y
    if _fused_applicable(x_local.shape[0], x_local, weight):
        from . import fused_tp
        return fused_tp.FusedAGLinear.apply(x_local, weight, bias, activation, want_gathered)
    gathered = GatherTokens.apply(x_local, group)
    y = ColumnParallelLinear.apply(gathered, weight, bias, activation, group, True)
    return (y, gathered) if want_gathered else y


def linear_rs(x, weight, bias=None, group=None, residual_full=None):
    """Row-parallel linear with token-sharded output: y[T/t, N] = reduce_scatter(x W^T + b + residual_full)."""
    n = _group_size(group)
    if n == 1:
        y = RowParallelLinear.apply(x, weight, bias, None, False)
        return y if residual_full is None else y + residual_full
    if _fused_applicable(x.shape[0] // n, x, weight, residual_full) and x.shape[0] % n == 0:
        from . import fused_tp
        return fused_tp.FusedLinearRS.apply(x, weight, bias, residual_full)
    partial = RowParallelLinear.apply(x, weight, bias, group, False)
    if residual_full is not None:
        partial = partial + residual_full
    return ScatterSumTokens.apply(partial, group)
