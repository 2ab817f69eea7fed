"""Base class of the distributed (tensor-parallel) modules and the keyword-argument parser they share.

Reference: smdistributed/modelparallel/torch/nn/dist_module.py:1-32 (marker base class with
``can_distribute``) and the ``parse_args`` helper used by every ``_KEYS`` table in
torch/nn/transformer.py.
"""

from __future__ import annotations

from collections import OrderedDict

import torch.nn as nn

from ..backend.exceptions import SMPInvalidArgumentError


class DistributedModule(nn.Module):
    """Marker base: a module whose parameters are already sharded across the tensor-parallel group and
    which must therefore never be replaced or re-distributed by the registry."""

    @staticmethod
    def can_distribute(*args, **kwargs) -> bool:
        return True

    @staticmethod
    def can_shard_activation_offloading() -> bool:
        return False


def parse_args(args, kwargs, keys: OrderedDict, obj=None) -> dict:
    """Resolve positional + keyword arguments against an ordered {name: default} table and (optionally)
    set them as attributes on ``obj``.  Unknown keywords are ignored so one config dict can be handed
    down a tower of modules with different tables (as the reference does)."""
    names = list(keys)
    if len(args) > len(names):
        raise SMPInvalidArgumentError(f"too many positional arguments ({len(args)} > {len(names)})")
    cfg = dict(keys)
    for n, v in zip(names, args):
        cfg[n] = v
    for k, v in kwargs.items():
        if k in cfg:
            if k in names[: len(args)]:
                raise SMPInvalidArgumentError(f"argument {k} given both positionally and by keyword")
            cfg[k] = v
    if obj is not None:
        for k, v in cfg.items():
            setattr(obj, k, v)
    return cfg


def update_copy_dict(d: dict, **updates) -> dict:
    out = dict(d)
    out.update(updates)
    return out
