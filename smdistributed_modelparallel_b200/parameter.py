"""Delayed parameter initialisation.

Large models must not be materialised in full on every rank before partitioning.  Inside
``smp.delay_param_initialization(enabled=True)`` modules are constructed on the ``meta`` device (no
storage); the initialisers that ran during construction are irrelevant because each module's
``reset_parameters`` (or the recorded custom initialiser) is replayed after partitioning, only for the
modules this rank owns, directly on the target device (reference smdistributed/modelparallel/torch/
parameter.py:35-76 records and replays initialisers; :23-32 is the sharded-init variant for SDP: here the
ZeRO-3 engine materialises, initialises, shards and releases one unit at a time - parallel/sdp.py::_build -
so no rank ever holds the whole model).
"""

from __future__ import annotations

from contextlib import contextmanager

import os

import torch
import torch.nn as nn

from .backend.exceptions import DelayedParamDeviceError

_delayed_modules = []          # modules constructed under the context (in construction order)
_active = False


def is_delayed_init_active() -> bool:
    return _active


@contextmanager
def delay_param_initialization(enabled=True):
    global _active
    if not enabled:
        yield
        return
    prev = _active
    _active = True
    try:
        with torch.device("meta"):
            yield
    finally:
        _active = prev


def register_custom_initializer(module: nn.Module, fn):
    """fn(module) initialises the module's own (non-recursive) parameters once they are materialised."""
    module._smp_custom_init = fn


def _materialize_module(m: nn.Module, device):
    has_meta = any(p.is_meta for p in m.parameters(recurse=False)) or any(b.is_meta for b in m.buffers(recurse=False))
    if not has_meta:
        return False
    for name, p in list(m.named_parameters(recurse=False)):
        if p.is_meta:
            new = nn.Parameter(torch.empty(p.shape, dtype=p.dtype, device=device), requires_grad=p.requires_grad)
            for tag in ("_smp_dist_axis", "_smp_scaled_batch", "_smp_one_rank"):
                if hasattr(p, tag):
                    setattr(new, tag, getattr(p, tag))
            m._parameters[name] = new
    for name, b in list(m.named_buffers(recurse=False)):
        if b.is_meta:
            m._buffers[name] = torch.empty(b.shape, dtype=b.dtype, device=device)
    # SMP_USE_FLOAT32_INIT=1 (reference torch/parameter.py:20): run the initialisers on fp32 storage and round
    # once, so 16-bit models start from the same values as their fp32 counterparts
    fp32_init = os.environ.get("SMP_USE_FLOAT32_INIT", "0") not in ("0", "", "false")
    lowp = {}
    if fp32_init:
        for name, p in list(m.named_parameters(recurse=False)):
            if p.dtype in (torch.float16, torch.bfloat16):
                lowp[name] = p.dtype
                p.data = p.data.float()
    custom = getattr(m, "_smp_custom_init", None)
    if custom is not None:
        custom(m)
    elif hasattr(m, "reset_parameters"):
        m.reset_parameters()
    elif hasattr(m, "_init_weights"):
        m._init_weights()
    else:
        for p in m.parameters(recurse=False):
            if p.dim() > 1:
                nn.init.normal_(p, 0.0, 0.02)
            else:
                nn.init.zeros_(p)
    for name, dt in lowp.items():
        m._parameters[name].data = m._parameters[name].data.to(dt)
    return True


def materialize_delayed(root: nn.Module, local_only=None, device=None):
    """Give storage to every meta parameter of the modules selected by ``local_only`` and run their
    initialisers.  Tied parameters (one Parameter object reachable from several modules) are
    materialised once and re-tied."""
    from .state_mod import state

    if device is None:
        device = state.device if getattr(state, "initialized", False) else torch.device("cpu")
    if not any(p.is_meta for p in root.parameters()):
        return 0
    owners = {}
    for m in root.modules():
        for name, p in m.named_parameters(recurse=False):
            owners.setdefault(id(p), []).append((m, name))
    count = 0
    for m in root.modules():
        if local_only is not None and not local_only(m):
            continue
        before = {name: id(p) for name, p in m.named_parameters(recurse=False)}
        if _materialize_module(m, device):
            count += 1
            for name, old_id in before.items():
                new = m._parameters[name]
                if id(new) != old_id:
                    for other, oname in owners.get(old_id, []):
                        if other is not m or oname != name:
                            other._parameters[oname] = new
    return count
