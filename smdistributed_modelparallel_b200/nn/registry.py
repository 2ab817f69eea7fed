"""Registry mapping ordinary module classes to their tensor-parallel counterparts.

``@smp.tp_register(DistCls, init_hook, forward_hook, return_hook)`` / ``smp.tp_register_with_module``
record: how to turn the original constructor arguments into the distributed module's (init hook), how
to adapt call arguments (forward hook) and results (return hook), and how to translate state dicts in
both directions.  ``distribute(module)`` builds the replacement for a module marked with
``smp.tensor_parallelism`` (reference smdistributed/modelparallel/torch/tp_registry.py:164-310).

Independent implementation notes: constructor arguments are captured through a class-level
``__init__`` wrapper into a weak map; forward / return hooks are attached to the distributed instance
with PyTorch's own pre-forward (with kwargs) and forward hooks instead of a patched ``forward``.
"""

from __future__ import annotations

import inspect
import weakref

import torch.nn as nn

from ..backend.exceptions import TPModuleRegisterError, UnsupportedTPModuleError
from ..backend.logger import get_logger
from .dist_module import DistributedModule


class TensorParallelismRegistry:
    def __init__(self):
        self.tp_modules: dict = {}
        self.init_hooks: dict = {}
        self.translate_functions: dict = {}
        self._forward_hooks: dict = {}
        self._return_hooks: dict = {}
        self._wrapped_inits: set = set()
        self.module_init_args = weakref.WeakKeyDictionary()
        self.active_translate_functions: list = []
        self._builtin_registered = False

    # ------------------------------------------------------------------ registration
    def _track_init(self, module_cls):
        if module_cls in self._wrapped_inits:
            return
        orig = module_cls.__init__
        registry = self

        def tracking_init(obj, *args, **kwargs):
            if type(obj) is module_cls:
                registry.module_init_args[obj] = (args, dict(kwargs))
            return orig(obj, *args, **kwargs)

        tracking_init.__wrapped__ = orig
        module_cls.__init__ = tracking_init
        self._wrapped_inits.add(module_cls)

    def _store(self, module_cls, dist_module, init_hook, forward_hook, return_hook, translate_functions=None):
        if not (inspect.isclass(dist_module) and issubclass(dist_module, DistributedModule)):
            raise TPModuleRegisterError(f"{getattr(dist_module, '__name__', dist_module)} must derive from "
                                        "smp.nn.DistributedModule")
        self._track_init(module_cls)
        self.tp_modules[module_cls] = dist_module
        if init_hook is not None:
            self.init_hooks[module_cls] = init_hook
        if forward_hook is not None:
            self._forward_hooks[(module_cls, dist_module)] = forward_hook
        if return_hook is not None:
            self._return_hooks[(module_cls, dist_module)] = return_hook
        if translate_functions is not None:
            self.translate_functions[module_cls] = translate_functions

    def register(self, dist_module, init_hook=None, forward_hook=None, return_hook=None):
        def deco(module_cls):
            self._store(module_cls, dist_module, init_hook, forward_hook, return_hook)
            return module_cls
        return deco

    def register_with_module(self, module_cls, dist_module, init_hook=None, forward_hook=None, return_hook=None,
                             translate_functions=None):
        self._store(module_cls, dist_module, init_hook, forward_hook, return_hook, translate_functions)

    def register_builtin(self):
        """nn.Linear / nn.Embedding counterparts and the HuggingFace mappings (idempotent)."""
        if self._builtin_registered:
            return
        self._builtin_registered = True
        from .embedding import DistributedEmbedding
        from .linear import DistributedLinear

        def linear_init(in_features, out_features, bias=True, device=None, dtype=None):
            return (in_features, out_features), {"bias": bias}

        def embedding_init(num_embeddings, embedding_dim, padding_idx=None, max_norm=None, norm_type=2.0,
                           scale_grad_by_freq=False, sparse=False, _weight=None, _freeze=False, device=None,
                           dtype=None):
            if max_norm is not None or scale_grad_by_freq or sparse:
                return None
            return (num_embeddings, embedding_dim), {"padding_idx": padding_idx}

        self._store(nn.Linear, DistributedLinear, linear_init, None, None)
        self._store(nn.Embedding, DistributedEmbedding, embedding_init, None, None)
        try:
            from .predefined_hooks import register_predefined
            register_predefined(self)
        except Exception as e:  # transformers missing or an unsupported version
            get_logger().debug(f"HuggingFace tensor-parallel hooks not registered: {e}")

    # ------------------------------------------------------------------ queries
    def is_supported(self, module_cls) -> bool:
        return module_cls in self.tp_modules

    def reset(self):
        self.module_init_args = weakref.WeakKeyDictionary()
        self.active_translate_functions = []

    def clear_module_init_args(self):
        self.module_init_args = weakref.WeakKeyDictionary()

    # ------------------------------------------------------------------ replacement
    def distribute(self, module: nn.Module):
        from ..state_mod import state

        cls = type(module)
        if cls not in self.tp_modules:
            raise UnsupportedTPModuleError(f"no tensor-parallel counterpart registered for {cls.__name__}")
        if module not in self.module_init_args:
            get_logger().warning(f"constructor arguments of {cls.__name__} were not recorded; leaving it as is")
            return module
        args, kwargs = self.module_init_args[module]
        if cls in self.init_hooks:
            converted = self.init_hooks[cls](*args, **kwargs)
            if converted is None:
                get_logger().warning(f"Unable to distribute instance of {cls} since not all of its __init__ "
                                     "arguments are supported.")
                return module
            cargs, ckwargs = converted
            ckwargs = dict(ckwargs)
        else:
            cargs, ckwargs = args, dict(kwargs)
        dist_cls = self.tp_modules[cls]
        accepted = inspect.signature(dist_cls.__init__).parameters
        keys = getattr(dist_cls, "_KEYS", {})
        for k, v in state.module_manager.get_tp_config(module).items():
            if k in accepted or k in keys:
                ckwargs[k] = v
        if not dist_cls.can_distribute(*cargs, **ckwargs):
            return module
        dist_mod = dist_cls(*cargs, **ckwargs)
        fh = self._forward_hooks.get((cls, dist_cls))
        rh = self._return_hooks.get((cls, dist_cls))
        if fh is not None:
            def pre(mod, a, kw, _fh=fh):
                out = _fh(*a, **kw)
                return out if out is not None else (a, kw)
            dist_mod.register_forward_pre_hook(pre, with_kwargs=True)
        if rh is not None:
            dist_mod.register_forward_hook(lambda mod, a, out, _rh=rh: _rh(out))
        tf = self.translate_functions.get(cls)
        if tf is not None and tf not in self.active_translate_functions:
            self.active_translate_functions.append(tf)
        if state.cfg._match_weights:
            match_weights(module, dist_mod, tf[1] if tf is not None else None)
        return dist_mod


def match_weights(module: nn.Module, dist_mod: nn.Module, hf_to_smp=None) -> int:
    """``_match_weights`` (internal config key, reference torch/tp_registry.py:47-204,237-242): give the distributed
    replacement the values of the module it replaces, so a test can compare the two numerically.  The reference copies
    attribute by attribute with one routine per ``optimize`` mode; here the original's state dict goes through the
    registered HuggingFace -> smp key translation and every entry is cut by the layout the distributed modules declare
    (``state_dict_layout`` / the distribution-axis tags), which covers both layouts and uneven splits.  Returns the
    number of tensors copied; modules that still sit on the meta device (delayed initialisation) are left alone."""
    import torch

    from ..checkpoint_utils import ONE_RANK, slice_for_rank
    from .utils import is_distributed_parameter, is_one_rank_parameter, tp_rank, tp_size

    sd = module.state_dict()
    if any(torch.is_tensor(v) and v.is_meta for v in sd.values()):
        return 0
    if hf_to_smp is not None:
        sd = hf_to_smp(dict(sd))
    layout = {}
    for mname, m in dist_mod.named_modules():
        prefix = mname + "." if mname else ""
        declared = getattr(m, "state_dict_layout", None)
        if callable(declared):
            layout.update({prefix + k: v for k, v in declared().items()})
            continue
        for pname, p in m.named_parameters(recurse=False):
            if is_one_rank_parameter(p):
                layout[prefix + pname] = (ONE_RANK, None)
            elif is_distributed_parameter(p):
                shapes = getattr(m, "weight_split_shapes", None)
                layout[prefix + pname] = (p._smp_dist_axis, shapes.get(pname) if isinstance(shapes, dict) else shapes)
    t, r = tp_size(), tp_rank()
    own = dist_mod.state_dict()
    picked = {}
    for k, v in sd.items():
        if k not in own or not torch.is_tensor(v):
            continue
        spec = layout.get(k)
        if spec is not None and t > 1 and spec[0] != ONE_RANK:
            v = slice_for_rank(v, spec, r, t)
        if tuple(v.shape) == tuple(own[k].shape):
            picked[k] = v
    dist_mod.load_state_dict(picked, strict=False)
    return len(picked)
