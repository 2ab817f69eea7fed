"""Tensor-parallel helpers: uneven channel splits, parameter tagging scopes, and the TP collectives as
autograd functions.

Public names follow the reference (reference smdistributed/modelparallel/torch/nn/utils.py:120-214
parameter scopes, :250-270 split arithmetic, :317-445 collective API, :465-662 autograd functions,
:701-821 implementations).  Collectives here run on ``torch.distributed`` (NCCL / gloo); the fused
compute+collective kernels for the transformer hot path live in ``parallel/fused_tp.py`` and are
selected by the layers, not here.  Tags are stored on the parameter objects themselves:

* ``_smp_dist_axis``    - axis along which the full weight is split across tp ranks
* ``_smp_scaled_batch`` - the parameter's gradient already covers the whole tp group's batch
* ``_smp_one_rank``     - only tp_rank 0 holds a meaningful copy (bias of a row-parallel layer)
"""

from __future__ import annotations

from contextlib import contextmanager

import torch
import torch.distributed as dist
import torch.nn as nn
from torch.nn import init

from ..backend.exceptions import SMPInvalidArgumentError, SplitShapeLenError, SplitShapeMismatchError


def _state():
    from ..state_mod import state
    return state


def tp_size() -> int:
    s = _state()
    return s.core.tp_size() if getattr(s, "initialized", False) else 1


def tp_rank() -> int:
    s = _state()
    return s.core.tp_rank() if getattr(s, "initialized", False) else 0


def tp_group():
    return _state().tp_process_group


# ----------------------------------------------------------------------------- split arithmetic
def get_local_channels(num_channels: int, rank: int | None = None, size: int | None = None) -> int:
    """Channels owned by ``rank`` when ``num_channels`` is split as evenly as possible (first ranks get
    the remainder)."""
    rank = tp_rank() if rank is None else rank
    size = tp_size() if size is None else size
    base, rem = divmod(num_channels, size)
    return base + (1 if rank < rem else 0)


def get_start_pos_for_slicing(num_channels: int, rank: int | None = None, size: int | None = None) -> int:
    rank = tp_rank() if rank is None else rank
    size = tp_size() if size is None else size
    base, rem = divmod(num_channels, size)
    return rank * base + min(rank, rem)


def get_split_shapes(num_channels: int, size: int | None = None) -> list[int]:
    size = tp_size() if size is None else size
    return [get_local_channels(num_channels, r, size) for r in range(size)]


def _check_split_shapes(split_shapes, total=None):
    if split_shapes is None:
        return
    if len(split_shapes) != tp_size():
        raise SplitShapeLenError(f"split_shapes has {len(split_shapes)} entries for tensor-parallel degree {tp_size()}")
    if total is not None and sum(split_shapes) != total:
        raise SplitShapeMismatchError(f"split_shapes {split_shapes} do not add up to {total}")


# ----------------------------------------------------------------------------- parameter tagging
def mark_distributed(p, axis):
    p._smp_dist_axis = axis
    p._smp_scaled_batch = True


def mark_scaled_batch(p):
    p._smp_scaled_batch = True


def mark_one_rank(p):
    p._smp_one_rank = True


def is_distributed_parameter(p) -> bool:
    return getattr(p, "_smp_dist_axis", None) is not None


def is_scaled_batch_parameter(p) -> bool:
    return bool(getattr(p, "_smp_scaled_batch", False))


def is_one_rank_parameter(p) -> bool:
    return bool(getattr(p, "_smp_one_rank", False))


@contextmanager
def parameter_creation_scope(module, scaled_batch=True, dtype=None, use_normal=False, initializer_range=0.02):
    """Parameters created inside see the whole tp group's batch -> tagged scaled-batch (reduced over the
    RDP group only).  Optionally creates them in ``dtype`` and re-initialises Linear/Embedding weights
    with N(0, initializer_range)."""
    prev = torch.get_default_dtype()
    if dtype is not None:
        torch.set_default_dtype(dtype)
    before = {id(p) for p in module.parameters()} | {id(b) for b in module.buffers()}
    try:
        yield
    finally:
        torch.set_default_dtype(prev)
    if scaled_batch:
        for t in list(module.parameters()) + list(module.buffers()):
            if id(t) not in before:
                mark_scaled_batch(t)
    if use_normal:
        def normal_init(m, std=initializer_range):
            if isinstance(m, nn.Linear):
                m.weight.data.normal_(0.0, std)
                if m.bias is not None:
                    m.bias.data.zero_()
            elif isinstance(m, nn.Embedding):
                m.weight.data.normal_(0.0, std)
                if m.padding_idx is not None:
                    m.weight.data[m.padding_idx].zero_()

        from ..parameter import is_delayed_init_active, register_custom_initializer
        for m in module.modules():
            if isinstance(m, (nn.Linear, nn.Embedding)):
                if is_delayed_init_active() or m.weight.is_meta:
                    # no storage yet: replay the same initialiser when the module is materialised
                    if getattr(m, "_smp_custom_init", None) is None:
                        register_custom_initializer(m, normal_init)
                else:
                    normal_init(m)


@contextmanager
def _fan_patch(which, fan_in_scale, fan_out_scale):
    """Make torch.nn.init see the fan-in / fan-out of the full (unsplit) weight."""
    orig = init._calculate_fan_in_and_fan_out
    t = tp_size()

    def patched(tensor):
        fi, fo = orig(tensor)
        fi, fo = fi * (fan_in_scale or 1), fo * (fan_out_scale or 1)
        return (fi * t, fo) if which == "input" else (fi, fo * t)

    init._calculate_fan_in_and_fan_out = patched
    try:
        yield
    finally:
        init._calculate_fan_in_and_fan_out = orig


@contextmanager
def _partition_scope(module, which, fan_in_scale, fan_out_scale, exclude_from_dist, axis):
    exclude = set(exclude_from_dist or [])
    before = {id(p) for p in module.parameters()} | {id(b) for b in module.buffers()}
    with _fan_patch(which, fan_in_scale, fan_out_scale):
        yield
    named = list(module.named_parameters()) + list(module.named_buffers())
    for n, t in named:
        if n in exclude:
            mark_one_rank(t)
        elif id(t) not in before:
            mark_distributed(t, axis)


def initialize_with_input_partition(module, fan_in_scale=None, fan_out_scale=None, exclude_from_dist=None, axis=1):
    """Scope for creating weights whose INPUT channels are split across tp ranks (row-parallel)."""
    return _partition_scope(module, "input", fan_in_scale, fan_out_scale, exclude_from_dist, axis)


def initialize_with_output_partition(module, fan_in_scale=None, fan_out_scale=None, exclude_from_dist=None, axis=0):
    """Scope for creating weights whose OUTPUT channels are split across tp ranks (column-parallel)."""
    return _partition_scope(module, "output", fan_in_scale, fan_out_scale, exclude_from_dist, axis)


# ----------------------------------------------------------------------------- raw collectives
def _throttled(fn, *a, **k):
    from ..utils.throttler import throttler
    return throttler.run(fn, *a, **k)


def _allgather_impl(x: torch.Tensor, dim: int, split_shapes=None, group=None) -> torch.Tensor:
    """Concatenate every rank's ``x`` along ``dim`` (sizes along ``dim`` may differ per rank)."""
    group = tp_group() if group is None else group
    n = dist.get_world_size(group)
    if n == 1:
        return x
    x = x.contiguous()
    if split_shapes is None or len(set(split_shapes)) == 1:
        if dim == 0:
            out = torch.empty((x.shape[0] * n,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
            _throttled(dist.all_gather_into_tensor, out, x, group=group)
            return out
        parts = torch.empty((n,) + tuple(x.shape), dtype=x.dtype, device=x.device)
        _throttled(dist.all_gather_into_tensor, parts.view((n * x.shape[0],) + tuple(x.shape[1:])), x, group=group)
        return torch.cat(list(parts.unbind(0)), dim=dim)
    # uneven: pad to the largest shard, gather, trim
    mx = max(split_shapes)
    pad_shape = list(x.shape)
    pad_shape[dim] = mx
    padded = x.new_zeros(pad_shape)
    padded.narrow(dim, 0, x.shape[dim]).copy_(x)
    parts = torch.empty((n,) + tuple(pad_shape), dtype=x.dtype, device=x.device)
    _throttled(dist.all_gather_into_tensor, parts.view((n * pad_shape[0],) + tuple(pad_shape[1:])), padded, group=group)
    return torch.cat([parts[r].narrow(dim, 0, split_shapes[r]) for r in range(n)], dim=dim)


def _reduce_scatter_impl(x: torch.Tensor, dim: int, split_shapes=None, group=None) -> torch.Tensor:
    """Sum over ranks, each rank keeping its slice along ``dim``."""
    group = tp_group() if group is None else group
    n = dist.get_world_size(group)
    if n == 1:
        return x
    r = dist.get_rank(group)
    total = x.shape[dim]
    shapes = split_shapes or get_split_shapes(total, n)
    if sum(shapes) != total:
        raise SplitShapeMismatchError(f"split shapes {shapes} do not add up to {total}")
    if len(set(shapes)) == 1 and dim == 0:
        x = x.contiguous()
        out = torch.empty((shapes[0],) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        _throttled(dist.reduce_scatter_tensor, out, x, group=group)
        return out
    mx = max(shapes)
    pieces = []
    start = 0
    for k in range(n):
        p = x.narrow(dim, start, shapes[k])
        start += shapes[k]
        if shapes[k] != mx:
            pad_shape = list(p.shape)
            pad_shape[dim] = mx
            q = x.new_zeros(pad_shape)
            q.narrow(dim, 0, shapes[k]).copy_(p)
            p = q
        pieces.append(p.contiguous())
    stacked = torch.cat(pieces, 0)            # rank k's block is rows [k*rows, (k+1)*rows)
    out = torch.empty_like(pieces[0])
    _throttled(dist.reduce_scatter_tensor, out, stacked, group=group)
    return out.narrow(dim, 0, shapes[r]).contiguous() if shapes[r] != mx else out


def _allreduce_impl(x: torch.Tensor, group=None) -> torch.Tensor:
    group = tp_group() if group is None else group
    if dist.get_world_size(group) == 1:
        return x
    x = x.contiguous().clone()
    _throttled(dist.all_reduce, x, group=group)
    return x


def _narrow_impl(x: torch.Tensor, dim: int, split_shapes=None, group=None) -> torch.Tensor:
    group = tp_group() if group is None else group
    n = dist.get_world_size(group)
    if n == 1:
        return x
    r = dist.get_rank(group)
    shapes = split_shapes or get_split_shapes(x.shape[dim], n)
    return x.narrow(dim, sum(shapes[:r]), shapes[r]).contiguous()


def _all_to_all_impl(x, split_dim, merge_dim, split_shapes=None, merge_shapes=None, group=None):
    """Split ``x`` along split_dim (piece k goes to rank k) and concatenate what arrives along merge_dim."""
    group = tp_group() if group is None else group
    n = dist.get_world_size(group)
    if n == 1:
        return x
    r = dist.get_rank(group)
    s_shapes = split_shapes or get_split_shapes(x.shape[split_dim], n)
    if merge_shapes is None:
        # every rank contributes a piece whose merge_dim extent equals that rank's local extent
        mine = torch.tensor([x.shape[merge_dim]], device=x.device, dtype=torch.int64)
        allm = torch.empty(n, device=x.device, dtype=torch.int64)
        dist.all_gather_into_tensor(allm, mine, group=group)
        merge_shapes = [int(v) for v in allm.tolist()]
    from ..backend import native_torch
    if native_torch.enabled("a2a"):
        # flat send / receive buffers and one grouped all-to-all in the native runtime (csrc/torch_ext/comm_runtime.cpp)
        return _throttled(native_torch.load().scatter_and_merge, group, x, split_dim, merge_dim,
                          [int(v) for v in s_shapes], [int(v) for v in merge_shapes])
    sends = [p.contiguous() for p in torch.split(x, s_shapes, dim=split_dim)]
    recvs = []
    for k in range(n):
        shp = list(x.shape)
        shp[split_dim] = s_shapes[r]
        shp[merge_dim] = merge_shapes[k]
        recvs.append(torch.empty(shp, dtype=x.dtype, device=x.device))
    if x.is_cuda:
        _throttled(dist.all_to_all, recvs, sends, group=group)
    else:
        # gloo has no list all-to-all: pairwise exchange (the reference's CPU path is point-to-point too,
        # reference torch/collectives.py:247-318)
        recvs[r].copy_(sends[r])
        ops = []
        for k in range(n):
            if k != r:
                peer = dist.get_global_rank(group, k)
                ops.append(dist.P2POp(dist.isend, sends[k], peer, group))
                ops.append(dist.P2POp(dist.irecv, recvs[k], peer, group))
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return torch.cat(recvs, dim=merge_dim)


# ----------------------------------------------------------------------------- autograd functions
class AllgatherForTP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dim, split_shapes):
        ctx.dim, ctx.shapes = dim, split_shapes
        return _allgather_impl(x, dim, split_shapes)

    @staticmethod
    def backward(ctx, g):
        return _reduce_scatter_impl(g, ctx.dim, ctx.shapes), None, None


class ReduceScatterForTP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dim, split_shapes):
        ctx.dim, ctx.shapes = dim, split_shapes
        return _reduce_scatter_impl(x, dim, split_shapes)

    @staticmethod
    def backward(ctx, g):
        return _allgather_impl(g, ctx.dim, ctx.shapes), None, None


class NarrowForTP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dim, split_shapes):
        ctx.dim, ctx.shapes = dim, split_shapes
        return _narrow_impl(x, dim, split_shapes)

    @staticmethod
    def backward(ctx, g):
        return _allgather_impl(g, ctx.dim, ctx.shapes), None, None


class FwdAllreduceForTP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return _allreduce_impl(x)

    @staticmethod
    def backward(ctx, g):
        return g


class BwdAllreduceForTP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return _allreduce_impl(g)


class ScatterAndMergeForTP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, split_dim, merge_dim, split_shapes, merge_shapes):
        ctx.args = (split_dim, merge_dim, split_shapes, merge_shapes, x.shape[merge_dim])
        return _all_to_all_impl(x, split_dim, merge_dim, split_shapes, merge_shapes)

    @staticmethod
    def backward(ctx, g):
        split_dim, merge_dim, split_shapes, merge_shapes, _ = ctx.args
        return _all_to_all_impl(g, merge_dim, split_dim, merge_shapes, split_shapes), None, None, None, None


def allgather_for_tp(x, dim=0, split_shapes=None):
    """All-gather along ``dim``; reduce-scatter in backward."""
    if tp_size() == 1:
        return x
    _check_split_shapes(split_shapes)
    return AllgatherForTP.apply(x, dim, split_shapes)


def fused_allgather_for_tp(tensors, dim=0):
    """All-gather several same-shaped tensors with one collective."""
    if tp_size() == 1:
        return list(tensors)
    stacked = torch.stack(list(tensors), 0)
    out = allgather_for_tp(stacked, dim + 1)
    return list(out.unbind(0))


def reduce_scatter_for_tp(x, dim=0, split_shapes=None):
    """Sum across tp ranks and keep the local slice along ``dim``; all-gather in backward."""
    if tp_size() == 1:
        return x
    _check_split_shapes(split_shapes, x.shape[dim])
    return ReduceScatterForTP.apply(x, dim, split_shapes)


def narrow_for_tp(x, dim=0, split_shapes=None):
    """Keep the local slice along ``dim`` (no communication); all-gather in backward."""
    if tp_size() == 1:
        return x
    _check_split_shapes(split_shapes, x.shape[dim])
    return NarrowForTP.apply(x, dim, split_shapes)


def fwd_allreduce_for_tp(x):
    return x if tp_size() == 1 else FwdAllreduceForTP.apply(x)


def bwd_allreduce_for_tp(x):
    return x if tp_size() == 1 else BwdAllreduceForTP.apply(x)


def scatter_and_merge_for_tp(x, split_dim, merge_dim, split_shapes=None, merge_shapes=None):
    """All-to-all: split along ``split_dim``, concatenate arrivals along ``merge_dim``."""
    if tp_size() == 1:
        return x
    if split_dim == merge_dim:
        raise SMPInvalidArgumentError("split_dim and merge_dim must differ")
    return ScatterAndMergeForTP.apply(x, split_dim, merge_dim, split_shapes, merge_shapes)


def shard_sequence(x, seq_dim=1):
    """prescaled_batch: keep this tp rank's slice of the sequence (reference nn/utils.py:45-56)."""
    return narrow_for_tp(x, seq_dim)


def unshard_sequence(x, seq_dim=1):
    return allgather_for_tp(x, seq_dim)
