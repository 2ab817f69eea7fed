"""State-dict plumbing shared by ``DistributedModel`` and the checkpoint API: gathering a full state dict
from tensor / pipeline shards, and loading full or partial state dicts into a sharded model (re-slicing
tensor-parallel weights when the degree differs).

Reference: smdistributed/modelparallel/torch/utils.py:395-478 (``slice_tp_tensors_in_state_dict`` with
distribution axes, uneven ``weight_split_shapes`` and one-rank parameters) and :481-658
(``collect_and_merge_in_group``: concatenate across TP, union across PP).
"""

from __future__ import annotations

from collections import OrderedDict

import torch

from .backend.collectives import CommGroup
from .backend.exceptions import MissingKeysInCheckpointError
from .backend.logger import get_logger
from .nn.utils import get_split_shapes, is_distributed_parameter, is_one_rank_parameter
from .state_mod import state

_META_KEYS = ("_smp_is_partial", "_smp_load_info")
ONE_RANK = "one_rank"


def collect_tp_layout(model) -> dict:
    """state-dict key -> (axis | ONE_RANK, split_shapes | None) for every tensor-parallel-sharded entry."""
    layout = {}
    for mname, m in model.module.named_modules():
        prefix = "module." + (mname + "." if mname else "")
        declared = getattr(m, "state_dict_layout", None)
        if callable(declared):
            for suffix, spec in declared().items():
                layout[prefix + suffix] = spec
            continue
        for pname, p in list(m.named_parameters(recurse=False)) + list(m.named_buffers(recurse=False)):
            if is_one_rank_parameter(p):
                layout[prefix + pname] = (ONE_RANK, None)
            elif is_distributed_parameter(p):
                shapes = getattr(m, "weight_split_shapes", None)
                if isinstance(shapes, dict):
                    shapes = shapes.get(pname)
                layout[prefix + pname] = (p._smp_dist_axis, shapes)
    return layout


def _strip_meta(sd):
    return OrderedDict((k, v) for k, v in sd.items() if k not in _META_KEYS)


def gather_full_state_dict(model, prefix="", cast_to_cpu=True, gather_to_rank0=False):
    core, comm = state.core, state.comm
    local = _strip_meta(model.local_state_dict())
    if cast_to_cpu:
        local = OrderedDict((k, v.detach().cpu() if torch.is_tensor(v) else v) for k, v in local.items())
    layout = collect_tp_layout(model)
    merged = local
    if core.tp_size() > 1:
        parts = comm.allgather(local, CommGroup.TP_GROUP)
        merged = OrderedDict()
        keys = list(OrderedDict.fromkeys(k for p in parts for k in p))
        for k in keys:
            spec = layout.get(k)
            have = [p[k] for p in parts if k in p]
            if spec is None or not torch.is_tensor(have[0]):
                merged[k] = have[0]
            elif spec[0] == ONE_RANK:
                merged[k] = parts[0].get(k, have[0])
            else:
                merged[k] = torch.cat(have, dim=spec[0])
    if core.pp_size() > 1:
        parts = comm.allgather(merged, CommGroup.PP_GROUP)
        merged = OrderedDict()
        for p in parts:
            for k, v in p.items():
                merged.setdefault(k, v)
    out = OrderedDict()
    for k, v in merged.items():
        out[prefix + k[len("module."):] if k.startswith("module.") else prefix + k] = v
    if gather_to_rank0 and core.rank() != 0:
        return OrderedDict()
    return out


def slice_for_rank(t: torch.Tensor, spec, tp_rank: int, tp_size: int) -> torch.Tensor:
    axis, shapes = spec
    if axis == ONE_RANK:
        return t
    shapes = shapes or get_split_shapes(t.shape[axis], tp_size)
    if sum(shapes) != t.shape[axis]:
        shapes = get_split_shapes(t.shape[axis], tp_size)
    start = sum(shapes[:tp_rank])
    return t.narrow(axis, start, shapes[tp_rank]).clone()


def load_into_model(model, state_dict, strict=True, translate_function=None):
    """Load a partial (this rank's) or full state dict into the local shards of ``model``."""
    from .checkpoint import _LazyModelCheckpoint, load_model_from_checkpoint

    if isinstance(state_dict, _LazyModelCheckpoint):
        path, tag, partial, tf = state_dict.args
        return load_model_from_checkpoint(model, path, tag, partial, strict, translate_function or tf)
    core = state.core
    info = state_dict.get("_smp_load_info") if isinstance(state_dict, dict) else None
    is_partial = bool(state_dict.get("_smp_is_partial", False)) if isinstance(state_dict, dict) else False
    sd = _strip_meta(state_dict)
    if translate_function is not None:
        sd = translate_function(sd)
    elif not is_partial:
        for _, hf_to_smp in state.tp_registry.active_translate_functions:
            if hf_to_smp is not None:
                sd = hf_to_smp(sd)
    sd = OrderedDict((k if k.startswith("module.") else "module." + k, v) for k, v in sd.items())
    layout = collect_tp_layout(model)
    saved_tp = info["tensor_parallel_degree"] if info else 1
    need_slice = (not is_partial) or saved_tp != core.tp_size()
    if need_slice and is_partial and saved_tp != 1:
        raise MissingKeysInCheckpointError(
            "a partial checkpoint saved with a different tensor_parallel_degree must be converted to a full "
            "checkpoint first (load it under its own degree and save with partial=False)")
    if need_slice and core.tp_size() > 1:
        for k, spec in layout.items():
            if k in sd and torch.is_tensor(sd[k]):
                if spec[0] == ONE_RANK:
                    if core.tp_rank() != 0:
                        del sd[k]
                else:
                    sd[k] = slice_for_rank(sd[k], spec, core.tp_rank(), core.tp_size())
    # restrict to what this rank actually holds
    expected = set(_strip_meta(model.local_state_dict()).keys())
    filtered = OrderedDict((k[len("module."):], v) for k, v in sd.items() if k in expected)
    missing = sorted(expected - set(sd.keys()))
    unexpected = [] if core.pp_size() > 1 else sorted(set(sd.keys()) - expected)
    res = model.module.load_state_dict(filtered, strict=False)
    if strict and (missing or unexpected):
        raise MissingKeysInCheckpointError(f"missing keys: {missing[:8]}{'...' if len(missing) > 8 else ''}; "
                                           f"unexpected keys: {unexpected[:8]}{'...' if len(unexpected) > 8 else ''}")
    opt = state.optimizer
    if opt is not None and getattr(opt, "fp16_optimizer", None) is not None:
        opt.fp16_optimizer.copy_model_to_master()
    if opt is not None and getattr(opt, "_opt_shard", None) is not None:
        opt._opt_shard.refresh_masters()
    if missing or unexpected:
        get_logger().debug(f"load_state_dict: {len(missing)} missing, {len(unexpected)} unexpected keys")
    return res
