"""Autograd markers on a module's inputs.

``mark_inputs(args, callback)`` passes every differentiable tensor found in ``args`` (nested tuples / lists /
dicts) through an identity autograd node whose backward runs ``callback()`` after every other node of the
module has run - i.e. when all gradients produced inside the module for this backward pass are final.  Module
full-backward hooks cannot give that signal for modules that take a tuple (they only see top-level tensor
arguments).  Used by the sharded-data-parallel engine (per-unit reduce-scatter) and the data-parallel reducer
(per-layer overlapped all-reduce); the reference gets the same signal from its native GradCounter fed by
autograd hooks (reference torch/model.py:401-403, torch/allreduce/reducer.py:59-92).
"""

from __future__ import annotations

import torch


def map_tensors(obj, fn):
    if torch.is_tensor(obj):
        return fn(obj)
    if isinstance(obj, tuple):
        return tuple(map_tensors(o, fn) for o in obj)
    if isinstance(obj, list):
        return [map_tensors(o, fn) for o in obj]
    if isinstance(obj, dict):
        return {k: map_tensors(v, fn) for k, v in obj.items()}
    return obj


class _Boundary(torch.autograd.Function):
    @staticmethod
    def forward(ctx, callback, *tensors):
        ctx.callback = callback
        return tuple(t.view_as(t) for t in tensors)

    @staticmethod
    def backward(ctx, *grads):
        ctx.callback()
        return (None,) + grads


def mark_inputs(args, callback):
    """Returns ``args`` with its differentiable tensors routed through the marker, or None when there is no
    tensor to hang it on (or autograd is off)."""
    if not torch.is_grad_enabled():
        return None
    found = []
    map_tensors(args, lambda t: found.append(t) if t.requires_grad else None)
    if not found:
        return None
    outs = iter(_Boundary.apply(callback, *found))
    return map_tensors(args, lambda t: next(outs) if t.requires_grad else t)
