"""Full (gathered) optimizer state dicts - the deprecated path the reference keeps for old checkpoints
(reference smdistributed/modelparallel/torch/optimizers/backcompat_opt.py:44-387: ``state_dict`` gathers every rank's
optimizer state into one dictionary keyed by the parameter's index in the *unpartitioned* model, ``load_state_dict``
slices it back per tensor-parallel rank and keeps the entries of the parameters this pipeline rank owns;
``DistributedOptimizer.load_state_dict`` itself refuses such dictionaries and points here, optimizers/optimizer.py:227).

``full_state_dict(opt)``      every rank of an mp group calls it; returns ``{"state": {global index: {...}},
                              "param_groups": [...], "_smp_param_names": {global index: name}, "_smp_tp_degree": t}``
                              with the per-parameter tensors merged over the tensor-parallel ranks (concatenated along
                              the distribution axis in rank order) and united over the pipeline ranks.
``load_full_state_dict``      the inverse: keeps the local parameters' entries, re-keys them to the local optimizer's
                              indices and takes this tp rank's block of every distributed tensor.

Blocks of a distributed tensor are stored in rank order, so a dictionary has to be loaded under the tensor-parallel
degree it was written with (the reference slices evenly along the axis, which is wrong for its own fused layouts as
well).  Not supported together with optimizer-state sharding or sharded data parallelism (the reference: :163-170).
"""

from __future__ import annotations

import torch

from ..backend.collectives import CommGroup
from ..backend.exceptions import SMPUnsupportedError, SMPValidationError
from ..state_mod import state


def _check(opt):
    cfg = state.cfg
    if cfg.shard_optimizer_state or opt.model.sdp_engine is not None:
        raise SMPUnsupportedError("full optimizer state dicts are not available with optimizer-state sharding or "
                                  "sharded data parallelism; use the partial (local) state dict")
    if not opt._post_partition_done:
        raise SMPValidationError("the model must be partitioned (run one step) before a full optimizer state dict "
                                 "can be built or loaded")


def _local_entries(opt):
    """name -> (state dict of that parameter on CPU, distribution axis or None)."""
    inner = opt.optimizer
    names = {id(p): n for n, p in opt.model.local_named_parameters()}
    out = {}
    for group in inner.param_groups:
        for p in group["params"]:
            st = inner.state.get(p)
            if not st:
                continue
            cpu = {k: (v.detach().cpu().clone() if torch.is_tensor(v) else v) for k, v in st.items()}
            out[names[id(p)]] = (cpu, getattr(p, "_smp_dist_axis", None), bool(getattr(p, "_smp_one_rank", False)))
    return out


def full_state_dict(opt):
    _check(opt)
    core, comm = state.core, state.comm
    local = _local_entries(opt)
    # ---- merge over the tensor-parallel group
    if core.tp_size() > 1:
        per_rank = comm.allgather(local, CommGroup.TP_GROUP)
        merged = {}
        for name in {n for d in per_rank for n in d}:
            holders = [d[name] for d in per_rank if name in d]
            first, axis, one_rank = holders[0]
            if axis is None or one_rank or len(holders) == 1:
                merged[name] = first
                continue
            out = {}
            for key, val in first.items():
                if torch.is_tensor(val) and val.dim() > axis and val.dim() > 0:
                    out[key] = torch.cat([h[0][key] for h in holders], dim=axis)
                else:
                    out[key] = val
            merged[name] = out
    else:
        merged = {n: v[0] for n, v in local.items()}
    # ---- unite over the pipeline group
    if core.pp_size() > 1:
        parts = comm.allgather(merged, CommGroup.PP_GROUP)
        merged = {}
        for d in parts:
            for n, v in d.items():
                merged.setdefault(n, v)
    index_of = {n: i for i, n in opt._global_index_to_name.items()}
    stt = {index_of[n]: v for n, v in merged.items() if n in index_of}
    groups = []
    for g, ids in zip(opt.optimizer.param_groups, opt._global_group_indices):
        d = {k: v for k, v in g.items() if k != "params"}
        d["params"] = list(ids)
        groups.append(d)
    return {"state": dict(sorted(stt.items())), "param_groups": groups,
            "_smp_param_names": dict(opt._global_index_to_name), "_smp_tp_degree": core.tp_size()}


def load_full_state_dict(opt, sd):
    _check(opt)
    core = state.core
    inner_sd = sd.get("optimizer_state_dict", sd)
    t_saved = inner_sd.get("_smp_tp_degree", sd.get("_smp_tp_degree"))
    names = inner_sd.get("_smp_param_names") or sd.get("_smp_param_names") or opt._global_index_to_name
    by_name = {names[i]: st for i, st in inner_sd["state"].items() if i in names}
    inner = opt.optimizer
    local_names = {id(p): n for n, p in opt.model.local_named_parameters()}
    t, r = core.tp_size(), core.tp_rank()
    loaded = 0
    for group in inner.param_groups:
        for p in group["params"]:
            name = local_names.get(id(p))
            st = by_name.get(name)
            if st is None:
                continue
            axis = getattr(p, "_smp_dist_axis", None)
            new = {}
            for key, val in st.items():
                if torch.is_tensor(val) and axis is not None and val.dim() > axis and val.shape != p.shape and \
                        not getattr(p, "_smp_one_rank", False):
                    if t_saved is not None and t_saved != t:
                        raise SMPValidationError(f"full optimizer state written with tensor_parallel_degree {t_saved} "
                                                 f"cannot be sliced for degree {t}")
                    sizes = [val.shape[axis] // t + (1 if i < val.shape[axis] % t else 0) for i in range(t)]
                    if sizes[r] != p.shape[axis]:      # uneven split recorded by the parameter's owner
                        sizes = None
                    start = sum(sizes[:r]) if sizes else r * p.shape[axis]
                    val = val.narrow(axis, start, p.shape[axis]).contiguous()
                new[key] = val.to(p.device) if torch.is_tensor(val) and val.dim() > 0 else val
            inner.state[p] = new
            loaded += 1
    for g, saved in zip(inner.param_groups, inner_sd.get("param_groups", [])):
        for k, v in saved.items():
            if k != "params":
                g[k] = v
    return loaded
