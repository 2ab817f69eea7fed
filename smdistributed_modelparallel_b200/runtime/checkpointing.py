"""Activation checkpointing: ``smp.checkpoint``, ``smp.checkpoint_sequential`` and the machinery behind
``smp.set_activation_checkpointing``.

Reference: smdistributed/modelparallel/torch/patches/checkpoint.py:108-359 (``CheckpointTupledFunction``
handling arbitrary argument structures, RNG + tensor-parallel seed restoration; ``checkpoint`` :312-324;
``checkpoint_sequential`` :327-359) and torch/sequential.py:187-264 (strategies ``each`` /
``contiguous`` / ``group_N`` over an ``nn.Sequential``).  When ``offload_activations`` is on, the saved
inputs of each checkpoint go to pinned host memory through the side-stream offloader (``offload.py``).
"""

from __future__ import annotations

import torch
import torch.nn as nn

from ..backend.exceptions import CheckpointingConfigError, InvalidReturnTypeFromCheckpointedModuleError
from ..backend.split import map_structure


def _flatten(obj, tensors):
    """Replace tensors by placeholders, collecting them in ``tensors``."""
    def f(x):
        if torch.is_tensor(x):
            tensors.append(x)
            return _Slot(len(tensors) - 1)
        return x
    return map_structure(f, obj)


def _unflatten(skel, tensors):
    return map_structure(lambda x: tensors[x.i] if isinstance(x, _Slot) else x, skel)


class _Slot:
    __slots__ = ("i",)

    def __init__(self, i):
        self.i = i


def _rng_snapshot():
    from ..state_mod import state

    snap = {"cpu": torch.get_rng_state()}
    if torch.cuda.is_available():
        snap["cuda"] = torch.cuda.get_rng_state()
    if getattr(state, "initialized", False):
        snap["smp"] = dict(state.rng_manager.get_states())
    return snap


def _rng_restore(snap):
    from ..state_mod import state

    torch.set_rng_state(snap["cpu"])
    if "cuda" in snap:
        torch.cuda.set_rng_state(snap["cuda"])
    if "smp" in snap:
        state.rng_manager.set_states(snap["smp"])


class CheckpointFunction(torch.autograd.Function):
    """Run ``fn(*args, **kwargs)`` without storing intermediate activations; recompute in backward."""

    @staticmethod
    def forward(ctx, fn, preserve_rng_state, skel, holder, *tensors):
        from ..state_mod import state

        ctx.fn, ctx.skel, ctx.preserve = fn, skel, preserve_rng_state
        ctx.rng = _rng_snapshot() if preserve_rng_state else None
        dev_type = next((t.device.type for t in tensors if t.device.type != "cpu"), "cpu")
        ctx.autocast = (dev_type, torch.is_autocast_enabled(dev_type), torch.get_autocast_dtype(dev_type))
        args, kwargs = _unflatten(skel, list(tensors))
        with torch.no_grad():
            out = fn(*args, **kwargs)
        # what to keep for the recomputation: the input tensors (possibly offloaded to pinned host memory)
        offload = getattr(state, "initialized", False) and state.cfg.offload_activations and \
            any(t.is_cuda for t in tensors)
        ctx.requires = [t.requires_grad for t in tensors]
        if offload:
            from ..offload import get_offloader
            off = get_offloader()
            ctx.offload_handles = [off.offload(t) if (t.is_cuda and t.is_floating_point()) else t for t in tensors]
            ctx.saved_direct = None
        else:
            ctx.offload_handles = None
            ctx.save_for_backward(*tensors)
        out_tensors = []
        ctx.out_skel = _flatten(out, out_tensors)
        if not out_tensors:
            raise InvalidReturnTypeFromCheckpointedModuleError()
        ctx.n_out = len(out_tensors)
        holder["out_skel"] = ctx.out_skel
        return tuple(out_tensors)

    @staticmethod
    def backward(ctx, *grads):
        if ctx.offload_handles is not None:
            from ..offload import get_offloader
            off = get_offloader()
            saved = [off.load(h) if not torch.is_tensor(h) else h for h in ctx.offload_handles]
        else:
            saved = list(ctx.saved_tensors)
        inputs = []
        for t, req in zip(saved, ctx.requires):
            d = t.detach()
            d.requires_grad_(req and d.is_floating_point())
            inputs.append(d)
        args, kwargs = _unflatten(ctx.skel, inputs)
        outer = _rng_snapshot() if ctx.preserve else None
        if ctx.preserve:
            _rng_restore(ctx.rng)
        try:
            dev_type, ac_on, ac_dtype = ctx.autocast
            # backward runs outside the caller's autocast region: recompute under the mode of the first pass
            with torch.enable_grad(), torch.autocast(dev_type, dtype=ac_dtype, enabled=ac_on):
                out = ctx.fn(*args, **kwargs)
        finally:
            if ctx.preserve:
                _rng_restore(outer)
        outs = []
        _flatten(out, outs)
        pairs = [(o, g) for o, g in zip(outs, grads) if o.requires_grad and g is not None]
        if pairs:
            torch.autograd.backward([o for o, _ in pairs], [g for _, g in pairs])
        in_grads = [t.grad if (t.requires_grad) else None for t in inputs]
        return (None, None, None, None, *in_grads)


_warned_cache = False


def _is_kv_cache(v):
    return hasattr(v, "update") and hasattr(v, "get_seq_length") and not torch.is_tensor(v)


def _drop_stateful_caches(kwargs, args=()):
    """A key/value cache object (HuggingFace ``Cache``: ``update`` + ``get_seq_length``) handed to a checkpointed
    module is appended to by the forward pass and *again* by the recomputation, which then attends over doubled
    keys and produces wrong gradients.  Training needs no cache: it is dropped (as HuggingFace's own gradient
    checkpointing does with ``use_cache``), with one warning."""
    global _warned_cache
    out = None
    for k, v in kwargs.items():
        if _is_kv_cache(v) or (k == "use_cache" and v is True):
            if out is None:
                out = dict(kwargs)
            out[k] = False if k == "use_cache" else None
    new_args = tuple(None if _is_kv_cache(a) else a for a in args) if any(_is_kv_cache(a) for a in args) else args
    if (out is not None or new_args is not args) and not _warned_cache:
        _warned_cache = True
        from ..backend.logger import get_logger
        get_logger().warning("activation checkpointing: key/value caches passed to a checkpointed module are "
                             "dropped (a cache mutated by the recomputation corrupts gradients); pass use_cache=False")
    return new_args, (kwargs if out is None else out)


def checkpoint(module_or_fn, *args, preserve_rng_state: bool = True, **kwargs):
    """Checkpoint one call of a module (or function): activations inside are recomputed in backward."""
    if torch.is_grad_enabled():
        args, kwargs = _drop_stateful_caches(kwargs, args)
    tensors = []
    skel = _flatten((args, kwargs), tensors)
    if not any(t.requires_grad for t in tensors) and not any(
            p.requires_grad for p in (module_or_fn.parameters() if isinstance(module_or_fn, nn.Module) else [])):
        return module_or_fn(*args, **kwargs)
    # autograd.Function only tracks tensor arguments that require grad; make sure at least one does so the
    # recomputation is reached even when only parameters need gradients
    anchor = None
    if not any(t.requires_grad for t in tensors):
        dev = tensors[0].device if tensors else None
        anchor = torch.zeros((), requires_grad=True, device=dev)
        tensors = tensors + [anchor]
    fn = module_or_fn if anchor is None else _with_anchor(module_or_fn)
    if anchor is not None:
        skel = (skel[0] + (_Slot(len(tensors) - 1),), skel[1])
    holder = {}
    outs = CheckpointFunction.apply(fn, preserve_rng_state, skel, holder, *tensors)
    return _unflatten(holder["out_skel"], list(outs))


def _with_anchor(fn):
    def wrapped(*args, **kwargs):
        return fn(*args[:-1], **kwargs)
    return wrapped


def _chunks(n, strategy):
    if strategy == "each":
        return [(i, i + 1) for i in range(n)]
    if strategy == "contiguous":
        return [(0, n)]
    if isinstance(strategy, str) and strategy.startswith("group_"):
        g = int(strategy[6:])
        if g <= 0:
            raise CheckpointingConfigError("group size must be positive")
        return [(i, min(n, i + g)) for i in range(0, n, g)]
    raise CheckpointingConfigError(f"unknown checkpointing strategy {strategy!r}")


def checkpoint_layers(layers, input, strategy="each", preserve_rng_state=True, pack_args_as_tuple=False):
    """Run ``layers`` one after another, checkpointing them in groups chosen by ``strategy``."""
    def run(a, b):
        def fn(x):
            for layer in layers[a:b]:
                x = layer(*x) if (isinstance(x, tuple) and not pack_args_as_tuple) else layer(x)
            return x
        return fn

    x = input
    for a, b in _chunks(len(layers), strategy):
        x = checkpoint(run(a, b), x, preserve_rng_state=preserve_rng_state)
    return x


def checkpoint_sequential(sequential, input, strategy="each", preserve_rng_state=True, pack_args_as_tuple=False):
    """Checkpoint an ``nn.Sequential`` in groups of layers: ``each`` layer separately, all layers as one
    ``contiguous`` block, or ``group_N`` layers at a time.  When the children live on several pipeline ranks
    the chain is cut at the partition boundaries and every rank checkpoints its own run of layers (reference
    torch/patches/sequential.py: chain splitting across partitions)."""
    if not isinstance(sequential, nn.Sequential):
        raise CheckpointingConfigError("checkpoint_sequential expects an nn.Sequential")
    from ..state_mod import state

    server = getattr(state, "exec_server", None)
    if server is not None and sequential in getattr(server, "_orig_forward", {}) and state.core.pp_size() > 1:
        return server.run_sequential(sequential, input, (strategy, preserve_rng_state, pack_args_as_tuple))
    return checkpoint_layers(list(sequential.children()), input, strategy, preserve_rng_state, pack_args_as_tuple)


def apply_configured_checkpointing(root: nn.Module):
    """Wrap ``forward`` of every module configured through ``smp.set_activation_checkpointing``."""
    from ..state_mod import state

    mm = state.module_manager
    for m in root.modules():
        cfg = mm.get_checkpoint_config(m)
        if cfg is None or not cfg.enabled or getattr(m, "_smp_ckpt_wrapped", False):
            continue
        orig = m.forward
        if isinstance(m, nn.Sequential):
            def fwd(inp, _m=m, _cfg=cfg, _orig=orig):
                if not torch.is_grad_enabled():
                    return _orig(inp)
                return checkpoint_sequential(_m, inp, _cfg.strategy, _cfg.preserve_rng_state, _cfg.pack_args_as_tuple)
        else:
            def fwd(*args, _cfg=cfg, _orig=orig, **kwargs):
                if not torch.is_grad_enabled():
                    return _orig(*args, **kwargs)
                return checkpoint(_orig, *args, preserve_rng_state=_cfg.preserve_rng_state, **kwargs)
        m.forward = fwd
        m._smp_ckpt_wrapped = True
