"""Tensor collectives with uneven shapes: variable-count all-gather and the split/merge all-to-all.

Same observable semantics as the reference's native NCCL helpers (reference smdistributed/modelparallel/torch/
collectives.py:214-245, expected results pinned by test/torch/mpi_4ps/test_collectives.py:13-114):

* ``allgatherv_tensor(tensor, counts, group)`` - ``tensor`` is a flat buffer of ``sum(counts)`` elements in which
  rank r has filled its own segment ``[sum(counts[:r]), +counts[r])``; after the call every segment is filled on
  every rank (in place).
* ``scatter_and_merge_tensor(tensor, split_axis, merge_axis, group, merge_shapes)`` - slice along ``split_axis``
  (remainder spread over the first ranks), exchange slices all-to-all, concatenate the received slices along
  ``merge_axis``: rank i ends up with the i-th slice of every rank.

Both run in the native tensor-communication runtime (``csrc/torch_ext/comm_runtime.cpp``: flat send / receive buffers and one grouped all-to-all /
padded all-gather through the c10d process group - NCCL on GPUs, gloo on CPU); ``SMP_NATIVE_COMM=0``
selects the ``torch.distributed`` formulation below instead.
"""

from __future__ import annotations

import torch
import torch.distributed as dist

from .backend.collectives import CommGroup
from .nn.utils import _all_to_all_impl


def _pg(group):
    from .state_mod import state

    if not isinstance(group, CommGroup):
        return group
    return {CommGroup.WORLD: state.world_process_group, CommGroup.DP_GROUP: state.dp_process_group,
            CommGroup.TP_GROUP: state.tp_process_group, CommGroup.PP_GROUP: state.pp_process_group,
            CommGroup.RDP_GROUP: state.rdp_process_group, CommGroup.MP_GROUP: state.mp_process_group}[group]


def allgatherv_tensor(tensor: torch.Tensor, counts, group=CommGroup.DP_GROUP) -> torch.Tensor:
    pg = _pg(group)
    n = 1 if pg is None else dist.get_world_size(pg)
    counts = [int(c) for c in counts]
    if len(counts) != n:
        raise ValueError(f"allgatherv: {len(counts)} counts for a group of {n}")
    flat = tensor.view(-1)
    if flat.numel() != sum(counts):
        raise ValueError(f"allgatherv: buffer holds {flat.numel()} elements, counts sum to {sum(counts)}")
    if n == 1:
        return tensor
    from .backend import native_torch
    if native_torch.enabled("a2a") and flat.is_contiguous():
        native_torch.load().allgatherv(pg, flat, counts)          # csrc/torch_ext/comm_runtime.cpp
        return tensor
    r = dist.get_rank(pg)
    offs = [sum(counts[:i]) for i in range(n)]
    if len(set(counts)) == 1:                       # even split: one in-place all-gather
        dist.all_gather_into_tensor(flat, flat[offs[r]:offs[r] + counts[r]].clone(), group=pg)
        return tensor
    mine = flat[offs[r]:offs[r] + counts[r]].clone()
    outs = [flat[offs[i]:offs[i] + counts[i]] if i != r else mine for i in range(n)]
    pieces = [torch.empty(counts[i], dtype=flat.dtype, device=flat.device) for i in range(n)]
    _uneven_allgather(pieces, mine, counts, pg)
    for i in range(n):
        if i != r:
            outs[i].copy_(pieces[i])
    return tensor


def _pad(t, n):
    out = t.new_zeros(n)
    out[:t.numel()] = t
    return out


def _uneven_allgather(pieces, mine, counts, pg):
    """all-gather of unequal segments: pad to the longest, gather, trim."""
    n, longest = len(counts), max(counts)
    buf = torch.empty(n * longest, dtype=mine.dtype, device=mine.device)
    dist.all_gather_into_tensor(buf, _pad(mine, longest), group=pg)
    for i in range(n):
        pieces[i].copy_(buf[i * longest:i * longest + counts[i]])


def scatter_and_merge_tensor(tensor: torch.Tensor, split_axis: int, merge_axis: int, group=CommGroup.TP_GROUP,
                             merge_shapes=None) -> torch.Tensor:
    pg = _pg(group)
    if pg is None or dist.get_world_size(pg) == 1:
        return tensor
    nd = tensor.dim()
    split_axis = split_axis + nd if split_axis < 0 else split_axis
    merge_axis = merge_axis + nd if merge_axis < 0 else merge_axis
    return _all_to_all_impl(tensor, split_axis, merge_axis, None, list(merge_shapes) if merge_shapes else None, pg)
