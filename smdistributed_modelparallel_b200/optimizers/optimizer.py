"""``smp.DistributedOptimizer``.

Wraps a torch optimizer built over ``model.parameters()`` (reference smdistributed/modelparallel/torch/
optimizers/optimizer.py:437-549): keeps only the parameters this rank owns, adds the 16-bit machinery
(``Bit16_Optimizer``: master weights, loss scaling, overflow skip) when fp16/bf16 is configured, hands
the update to the sharded-data-parallel engine or to the optimizer-state-sharding path when those are
enabled, and exposes the partial ("local") state-dict API used by checkpointing
(``local_optimizer_state_dict`` :150-172, ``local_fp16_state_dict`` :175-200, ``load_optimizer_checkpoint``).

The reference monkey-patches the user's optimizer object; this is a thin wrapper class that forwards
unknown attributes to it, which keeps ``isinstance`` checks on the inner optimizer possible through
``.optimizer``.
"""

from __future__ import annotations

import torch

from ..backend.exceptions import DistributedModelNotWrappedError, NotInitializedError
from ..backend.logger import get_logger
from ..precision.bit16 import Bit16_Optimizer, clip_grad_norm_fp32
from ..state_mod import state


class DistributedOptimizer:
    def __init__(self, optimizer, static_loss_scale=1.0, dynamic_loss_scale=False, dynamic_loss_args=None):
        if not state.initialized:
            raise NotInitializedError()
        if state.model is None:
            raise DistributedModelNotWrappedError("create smp.DistributedModel before smp.DistributedOptimizer")
        self.optimizer = optimizer
        self.model = state.model
        cfg = state.cfg
        self._bit16 = None
        self._opt_shard = None
        self._post_partition_done = False
        self._pending_state = None
        self._static_loss_scale = static_loss_scale
        self._dynamic_loss_scale = dynamic_loss_scale
        self._dynamic_loss_args = dynamic_loss_args
        self.redefined_params = False
        # delayed parameter initialisation replaces meta parameters by materialised ones after the partition is
        # known (pp > 1: at the first step, i.e. after this wrapper exists): remember every parameter's name so the
        # inner optimizer can be pointed at the real tensors then
        self._names_at_creation = {id(p): n for n, p in self.model.module.named_parameters(remove_duplicate=False)}
        # index of every parameter in the optimizer as built over the whole (unpartitioned) model: the key space of
        # full optimizer state dicts (optimizers/backcompat_opt.py)
        self._global_index_to_name, self._global_group_indices = {}, []
        idx = 0
        for group in optimizer.param_groups:
            ids = []
            for p in group["params"]:
                self._global_index_to_name[idx] = self._names_at_creation.get(id(p), f"param_{idx}")
                ids.append(idx)
                idx += 1
            self._global_group_indices.append(ids)
        state.optimizer = self
        self.model._call_after_partitioning(self._after_partition)
        if state.loaded_optimizer_state is not None:
            self._pending_state = state.loaded_optimizer_state
            state.loaded_optimizer_state = None

    # ------------------------------------------------------------------ setup once the partition is known
    def _after_partition(self, model, _opt=None):
        if self._post_partition_done:
            return
        self._post_partition_done = True
        cfg = state.cfg
        local = {id(p) for p in model.local_parameters()}
        by_name = dict(model.module.named_parameters(remove_duplicate=False))
        for group in self.optimizer.param_groups:
            kept, seen = [], set()
            for p in group["params"]:
                q = p
                if id(p) not in local:                       # stale (meta) object: look the parameter up by name
                    q = by_name.get(self._names_at_creation.get(id(p)))
                    if q is None or id(q) not in local:
                        continue
                    if p in self.optimizer.state:
                        self.optimizer.state[q] = self.optimizer.state.pop(p)
                    self.redefined_params = True
                if id(q) not in seen:
                    seen.add(id(q))
                    kept.append(q)
            group["params"] = kept
        self._build_name_maps()
        if model.sdp_engine is not None:
            model.sdp_engine.attach_optimizer(self.optimizer, clip=cfg.sdp_gradient_clipping,
                                              static_loss_scale=self._static_loss_scale,
                                              dynamic_loss_scale=self._dynamic_loss_scale,
                                              dynamic_loss_args=self._dynamic_loss_args)
        elif cfg.fp16 or cfg.bf16:
            self._bit16 = Bit16_Optimizer(
                self.optimizer, model=model, static_loss_scale=self._static_loss_scale if cfg.fp16 else 1.0,
                dynamic_loss_scale=self._dynamic_loss_scale and cfg.fp16, dynamic_loss_args=self._dynamic_loss_args,
                shard_optimizer_state=cfg.shard_optimizer_state, bf16=cfg.bf16)
            self._bit16.init_master_params()
        if cfg.shard_optimizer_state and model.sdp_engine is None:
            from ..parallel.optshard import OptimizerStateSharding
            self._opt_shard = OptimizerStateSharding(model, self.optimizer, self._bit16)
        if self._pending_state is not None:
            st, self._pending_state = self._pending_state, None
            self.load_optimizer_checkpoint_state(st)

    def _build_name_maps(self):
        """index <-> name maps for the parameters of the inner optimizer (used by partial checkpoints)."""
        names = {id(p): n for n, p in self.model.local_named_parameters()}
        self.param_index_to_name = {}
        idx = 0
        for group in self.optimizer.param_groups:
            for p in group["params"]:
                self.param_index_to_name[idx] = names.get(id(p), f"param_{idx}")
                idx += 1
        self.param_name_to_index = {n: i for i, n in self.param_index_to_name.items()}

    # ------------------------------------------------------------------ training API
    def zero_grad(self, set_to_none: bool = True):
        self.model.zero_grad_buffer()
        if self._opt_shard is not None or self.model.sdp_engine is not None:
            return
        if self._bit16 is not None:
            self._bit16.zero_grad()
        else:
            self.optimizer.zero_grad(set_to_none=set_to_none)
        for p in self.model.local_parameters():
            p.grad = None

    def update_master_grads(self):
        if self._bit16 is not None:
            return self._bit16.update_master_grads()
        return False

    def clip_master_grads(self, max_norm, norm_type=2):
        if self.model.sdp_engine is not None:
            return self.model.sdp_engine.clip_grad_norm(max_norm)
        if self._opt_shard is not None:
            return self._opt_shard.clip_grad_norm(max_norm, norm_type)
        if self._bit16 is not None:
            return self._bit16.clip_master_grads(max_norm, norm_type)
        pairs = [(p, p.grad) for g in self.optimizer.param_groups for p in g["params"] if p.grad is not None]
        total, coef = clip_grad_norm_fp32(pairs, max_norm, norm_type)
        for _, g in pairs:
            g.mul_(coef.to(g.dtype))
        return total

    def step(self, closure=None):
        if self.model.sdp_engine is not None:
            return self.model.sdp_engine.optimizer_step()
        if self._opt_shard is not None:
            return self._opt_shard.step()
        if self._bit16 is not None:
            return self._bit16.step(closure)
        return self.optimizer.step(closure)

    def backward(self, loss, **kw):
        self.model.backward(loss, **kw)

    # ------------------------------------------------------------------ properties
    @property
    def loss_scale(self):
        if self._bit16 is not None:
            return self._bit16.loss_scale
        if self.model.sdp_engine is not None:
            return self.model.sdp_engine.loss_scale
        return 1.0

    @property
    def overflow(self):
        return self._bit16.overflow if self._bit16 is not None else False

    @property
    def param_groups(self):
        return self.optimizer.param_groups

    @property
    def state(self):
        return self.optimizer.state

    @property
    def fp16_optimizer(self):
        return self._bit16

    def __getattr__(self, name):
        # only reached for attributes not found on the wrapper
        opt = self.__dict__.get("optimizer")
        if opt is None:
            raise AttributeError(name)
        bit16 = self.__dict__.get("_bit16")
        if bit16 is not None and not name.startswith("_") and hasattr(type(bit16), name):
            return getattr(bit16, name)            # master-weight inspection etc. (the 16-bit wrapper's API)
        return getattr(opt, name)

    # ------------------------------------------------------------------ state dicts
    def state_dict(self):
        sd = {"optimizer_state_dict": self.optimizer.state_dict(), "_smp_is_partial": False}
        if self._bit16 is not None:
            sd["fp16"] = self.local_fp16_state_dict()
        return sd

    def load_state_dict(self, sd):
        inner = sd.get("optimizer_state_dict", sd) if isinstance(sd, dict) else sd
        if isinstance(inner, dict) and "_smp_param_names" in inner and "_smp_is_partial" not in inner:
            from ..backend.exceptions import SMPValidationError
            raise SMPValidationError("this is a full (gathered) optimizer state dict; load it with "
                                     "load_state_dict_backcompat")
        self.load_optimizer_checkpoint_state(sd)

    def state_dict_backcompat(self):
        """Optimizer state gathered over the tensor- and pipeline-parallel ranks (deprecated full form; every rank of
        the model-parallel group must call it)."""
        from .backcompat_opt import full_state_dict
        return full_state_dict(self)

    def load_state_dict_backcompat(self, sd):
        from .backcompat_opt import load_full_state_dict
        return load_full_state_dict(self, sd)

    def local_optimizer_state_dict(self):
        """State of the parameters this rank updates, keyed like ``torch.optim.Optimizer.state_dict`` plus
        ``_smp_is_partial`` and the index->name map (so it can be re-keyed under another partition)."""
        if self.model.sdp_engine is not None:
            return self.model.sdp_engine.optimizer_state_dict()
        sd = self.optimizer.state_dict()
        sd["_smp_is_partial"] = True
        sd["_smp_param_index_to_name"] = dict(getattr(self, "param_index_to_name", {}))
        if self._opt_shard is not None:
            sd["_smp_opt_shard"] = self._opt_shard.shard_info()
        return sd

    def local_fp16_state_dict(self):
        if self._bit16 is None:
            return None
        b = self._bit16
        return {
            # the scaler object itself, as the reference writes it (reference optimizers/optimizer.py:185-190)
            "loss_scaler": b.loss_scaler.for_checkpoint(),
            "dynamic_loss_scale": b.dynamic_loss_scale,
            "overflow": b.overflow,
            "first_closure_call_this_step": b.first_closure_call_this_step,
            # under optimizer-state sharding the masters live in the shard slices (saved with the optimizer state)
            "fp32_from_fp16": [] if self._opt_shard is not None
            else [[m.detach().cpu() for m in grp] for grp in b.fp32_from_fp16_groups],
        }

    def load_optimizer_checkpoint_state(self, sd):
        if not self._post_partition_done:
            self._pending_state = sd
            return
        from ..checkpoint import _LazyOptimizerCheckpoint, load_optimizer_from_checkpoint
        if isinstance(sd, _LazyOptimizerCheckpoint):
            return load_optimizer_from_checkpoint(self, sd.args[0], sd.args[1], True, True, tag_is_final=True)
        if self.model.sdp_engine is not None:
            return self.model.sdp_engine.load_optimizer_state_dict(sd)
        inner = sd.get("optimizer_state_dict", sd)
        inner = {k: v for k, v in inner.items() if not k.startswith("_smp_")}
        shard_info = sd.get("optimizer_state_dict", sd).get("_smp_opt_shard")
        if "state" in inner and "param_groups" in inner:
            self.optimizer.load_state_dict(inner)
        if shard_info is not None and self._opt_shard is not None:
            self._opt_shard.load_shard_info(shard_info)
        fp16 = sd.get("fp16")
        if fp16 is not None and self._bit16 is not None:
            b = self._bit16
            b.loss_scaler.load_state_dict(fp16["loss_scaler"])
            b.dynamic_loss_scale = fp16["dynamic_loss_scale"]
            b.overflow = fp16["overflow"]
            b.first_closure_call_this_step = fp16["first_closure_call_this_step"]
            if self._opt_shard is None:
                for cur, saved in zip(b.fp32_from_fp16_groups, fp16["fp32_from_fp16"]):
                    for c, s in zip(cur, saved):
                        c.data.copy_(s.to(c.device))
                b.copy_master_to_model()
            # with optimizer-state sharding the fp32 masters are the per-rank slices restored by load_shard_info;
            # the wrapper's own (unused) master copies must not overwrite the parameters

    def load_optimizer_checkpoint(self, path, tag=None, partial=True, load_sharded_optimizer_state=True):
        from ..checkpoint import load_optimizer_from_checkpoint
        return load_optimizer_from_checkpoint(self, path, tag, partial, load_sharded_optimizer_state)
