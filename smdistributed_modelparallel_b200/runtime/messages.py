"""Control messages of the pipeline engine and the tensor (de)stubbing around them.

Same vocabulary as the reference's request / result dataclasses (reference smdistributed/modelparallel/
torch/messages.py:48-520: ModuleExecutionRequest, ForwardExecutionResult, BackwardExecutionResult,
MicrobatchEndResult, ...), trimmed to what a request/response protocol with segment-wise backward needs.
A message is a small picklable object sent on the bus' server channel; tensors inside arguments are
replaced by ``TensorStub`` (shape / dtype / requires_grad) and travel separately over NCCL or gloo p2p
in the order of their stub indices (reference torch/serialization.py:40-68, 283-363).
"""

from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any

import torch

from ..backend.split import map_structure


@dataclass(frozen=True)
class TensorStub:
    index: int
    shape: tuple
    dtype: torch.dtype
    requires_grad: bool


def _walk(fn, obj, _active=None):
    """``map_structure`` that also descends into the attributes of plain Python objects (an instance of a user class
    carrying tensors - reference torch/serialization.py:283-363 stubs those too, test/torch/test_serialization.py:13-26):
    such an object is shallow-copied with its attributes mapped, the caller's instance is left untouched.  An object
    met again while it is being walked (a reference cycle) is left as it is."""
    import copy
    import types
    from collections.abc import Mapping

    if torch.is_tensor(obj) or isinstance(obj, TensorStub):
        return fn(obj)
    if obj is None or isinstance(obj, (int, float, str, bytes, bool)):
        return fn(obj)
    _active = set() if _active is None else _active
    if id(obj) in _active:
        return obj
    _active.add(id(obj))
    try:
        return _walk_container(fn, obj, _active, copy, types, Mapping)
    finally:
        _active.discard(id(obj))


def _walk_container(fn, obj, _active, copy, types, Mapping):
    if isinstance(obj, (list, tuple)) and not hasattr(obj, "smp_slice"):
        mapped = [_walk(fn, o, _active) for o in obj]
        if isinstance(obj, tuple) and hasattr(obj, "_fields"):
            return type(obj)(*mapped)
        return type(obj)(mapped)
    if isinstance(obj, Mapping):
        try:
            return type(obj)((k, _walk(fn, v, _active)) for k, v in obj.items())
        except TypeError:                       # a mapping class with its own constructor signature
            new = copy.copy(obj)
            for k, v in obj.items():
                new[k] = _walk(fn, v, _active)
            return new
    d = getattr(obj, "__dict__", None)
    if isinstance(d, dict) and d and not isinstance(obj, (torch.nn.Module, type, types.ModuleType, types.FunctionType,
                                                           types.MethodType, torch.dtype, torch.device)):
        mapped = {k: _walk(fn, v, _active) for k, v in d.items()}
        if any(mapped[k] is not d[k] for k in d):
            new = copy.copy(obj)
            new.__dict__.update(mapped)
            return new
        return obj
    return fn(obj)


def stubify(obj):
    """Returns (skeleton, tensors): tensors replaced by stubs, listed in stub order."""
    tensors = []

    def f(x):
        if torch.is_tensor(x):
            tensors.append(x)
            return TensorStub(len(tensors) - 1, tuple(x.shape), x.dtype, bool(x.requires_grad))
        return x

    return _walk(f, obj), tensors


def destubify(skel, tensors):
    return _walk(lambda x: tensors[x.index] if isinstance(x, TensorStub) else x, skel)


def stubs_of(skel):
    out = []
    _walk(lambda x: out.append(x) if isinstance(x, TensorStub) else x, skel)
    return sorted(out, key=lambda s: s.index)


@dataclass
class ForwardRequest:
    """Execute ``module_name`` on its owner for microbatch ``mb`` (reference ModuleExecutionRequest)."""
    mb: int
    call_id: tuple
    module_name: str
    skeleton: Any                 # (args, kwargs) with TensorStubs
    grad_enabled: bool
    requester: int                # global rank
    # fast mode: the chunks [(global rank, (first child, last child + 1))] still to run after this one; the
    # output goes straight to the next of them and only the last one answers the requester
    chain: tuple = ()
    prev_hop: int | None = None   # the chain stage this request came from (None: from the requester)
    autocast: tuple | None = None  # (device type, dtype) when the caller runs under torch.autocast


@dataclass
class ForwardResult:
    mb: int
    call_id: tuple
    skeleton: Any                 # outputs with TensorStubs
    error: str | None = None


@dataclass
class BackwardRequest:
    """Gradients w.r.t. the outputs of a remote call; ``grad_present[i]`` tells which output stubs have one."""
    mb: int
    call_id: tuple
    grad_meta: list               # [(index, shape, dtype)] for the gradients that follow on p2p
    requester: int


@dataclass
class BackwardResult:
    mb: int
    call_id: tuple
    grad_meta: list               # [(input stub index, shape, dtype)] gradients w.r.t. the call's inputs
    error: str | None = None


@dataclass
class StepDone:
    """Leader -> followers: every microbatch finished; carries the step function's (detached) outputs, or
    the text of the exception that ended the step on the leader."""
    step: int
    outputs: Any = None
    error: Any = None


@dataclass
class PartitionInfo:
    """Leader -> everyone on the first step: module name -> pipeline rank."""
    assignment: dict = field(default_factory=dict)
