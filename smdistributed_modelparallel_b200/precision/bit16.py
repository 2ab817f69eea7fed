"""Mixed-precision wrappers: ``Bit16_Module`` (cast inputs/outputs) and ``Bit16_Optimizer`` (fp32
master weights, loss scaling, overflow skip, model-parallel-aware gradient clipping).

Behavioural parity with the reference (reference smdistributed/modelparallel/torch/fp16/fp16.py:210-239
Bit16_Module, :242-807 Bit16_Optimizer - init_master_params :380-464, step :597-651,
update_master_grads :719-731, clip_grad_norm_fp32 :61-162; state-dict keys at
torch/optimizers/optimizer.py:175-200).

B200-first differences: gradients of 16-bit parameters are accumulated in fp32 directly by the wgrad
GEMM epilogue into ``param.main_grad`` (one flat buffer owned by the DistributedModel), so the master
gradient is that buffer itself - there is no separate fp16 grad tensor, no fp16->fp32 copy pass; for
Adam/AdamW the update, the unscale / clip factor and the write-back of the 16-bit copy are a single
multi-tensor kernel launch sequence (``smpk_mt_adam`` with 5 lists).
"""

from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn

from ..backend.logger import get_logger
from ..ops import multi_tensor as MT
from .loss_scaler import DynamicLossScaler, LossScaler, overflow_sync_group, sync_overflow

_HALF_TYPES = (torch.float16, torch.bfloat16)


def _cast_tree(obj, fn):
    if torch.is_tensor(obj):
        return fn(obj)
    if isinstance(obj, (list, tuple)):
        out = [_cast_tree(o, fn) for o in obj]
        return type(obj)(*out) if hasattr(obj, "_fields") else type(obj)(out)
    if isinstance(obj, dict):
        return {k: _cast_tree(v, fn) for k, v in obj.items()}
    return obj


class Bit16_Module(nn.Module):
    """Holds the wrapped module in fp16/bf16; floating-point inputs are cast down, outputs cast back up."""

    def __init__(self, module: nn.Module, dtype: torch.dtype = torch.float16, cast_outputs: bool = True):
        super().__init__()
        self.dtype = dtype
        self.cast_outputs = cast_outputs
        self.add_module("module", module.to(dtype))

    def forward(self, *args, **kwargs):
        down = lambda t: t.to(self.dtype) if t.is_floating_point() and t.dtype == torch.float32 else t
        up = lambda t: t.float() if t.is_floating_point() and t.dtype == self.dtype else t
        out = self.module(*_cast_tree(args, down), **_cast_tree(kwargs, down))
        return _cast_tree(out, up) if self.cast_outputs else out

    def state_dict(self, *args, **kwargs):
        return self.module.state_dict(*args, **kwargs)

    def load_state_dict(self, state_dict, strict=True):
        return self.module.load_state_dict(state_dict, strict=strict)


def _mp_reduce(t: torch.Tensor, op):
    group = overflow_sync_group()
    if group is not None and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=op, group=group)
    return t


def clip_grad_norm_fp32(params_and_grads, max_norm, norm_type=2.0):
    """Global gradient norm over the model-parallel group and the clip coefficient (device tensor).

    ``params_and_grads``: iterable of (param, fp32 grad).  Tensor-parallel duplicates are counted once:
    replicated parameters contribute only on tp_rank 0 (reference fp16.py:88-111)."""
    from ..nn.utils import is_distributed_parameter, tp_rank

    grads = []
    dup_rank = tp_rank() != 0
    for p, g in params_and_grads:
        if g is None:
            continue
        if dup_rank and not is_distributed_parameter(p):
            continue
        grads.append(g)
    dev = grads[0].device if grads else torch.device("cuda" if torch.cuda.is_available() else "cpu")
    if norm_type == float("inf"):
        total = torch.stack([g.abs().max() for g in grads]).max() if grads else torch.zeros((), device=dev)
        total = _mp_reduce(total.reshape(1).float(), dist.ReduceOp.MAX)
    else:
        if grads:
            norm, _ = MT.multi_tensor_l2norm(MT.DEFAULT_CHUNK, None, [grads], False)
        else:
            norm = torch.zeros(1, device=dev)
        total = _mp_reduce(norm.float() ** 2, dist.ReduceOp.SUM).sqrt()
    coef = (max_norm / (total + 1e-6)).clamp(max=1.0)
    return total, coef


class Bit16_Optimizer:
    def __init__(self, init_optimizer, model=None, static_loss_scale=1.0, dynamic_loss_scale=False,
                 dynamic_loss_args=None, verbose=False, use_smp=True, shard_optimizer_state=False, bf16=False,
                 params_have_main_grad=None):
        # params_have_main_grad: accepted for compatibility; fp32 main_grad buffers are detected per parameter
        self.optimizer = init_optimizer
        self.model = model
        self.bf16 = bf16
        self.verbose = verbose
        self.shard_optimizer_state = shard_optimizer_state
        if dynamic_loss_scale:
            self.dynamic_loss_scale = True
            self.loss_scaler = DynamicLossScaler(**(dynamic_loss_args or {}))
        else:
            self.dynamic_loss_scale = False
            self.loss_scaler = LossScaler(static_loss_scale)
        self.overflow = False
        self.first_closure_call_this_step = True
        self.clip_grad_norm = clip_grad_norm_fp32
        self.fp16_groups = []
        self.fp32_from_fp16_groups = []
        self.fp32_from_fp32_groups = []
        self.master_params_created = False
        self._pending_clip = None          # (coef tensor) applied inside the fused update
        self._flag = None
        self._step_count = 0

    # ------------------------------------------------------------------ master weights
    def init_master_params(self):
        """Create the fp32 master copy of every 16-bit parameter (one flat buffer per param group) and
        point the inner optimizer at the masters."""
        if self.master_params_created:
            return
        opt = self.optimizer
        self.fp16_groups, self.fp32_from_fp16_groups, self.fp32_from_fp32_groups = [], [], []
        self._master_of = {}
        self._half_of = {}
        for group in opt.param_groups:
            halfs = [p for p in group["params"] if p.requires_grad and p.dtype in _HALF_TYPES]
            fulls = [p for p in group["params"] if p.requires_grad and p.dtype == torch.float32]
            frozen = [p for p in group["params"] if not p.requires_grad]
            masters = []
            if halfs:
                total = sum(p.numel() for p in halfs)
                flat = torch.empty(total, dtype=torch.float32, device=halfs[0].device)
                off = 0
                for p in halfs:
                    m = flat[off:off + p.numel()].view_as(p)
                    m.copy_(p.detach())
                    mp = nn.Parameter(m, requires_grad=True)
                    for tag in ("_smp_dist_axis", "_smp_scaled_batch", "_smp_one_rank"):
                        if hasattr(p, tag):
                            setattr(mp, tag, getattr(p, tag))
                    masters.append(mp)
                    self._master_of[p] = mp
                    self._half_of[id(mp)] = p
                    if p in opt.state:                      # migrate pre-existing state (resume)
                        opt.state[mp] = opt.state.pop(p)
                    off += p.numel()
            self.fp16_groups.append(halfs)
            self.fp32_from_fp16_groups.append(masters)
            self.fp32_from_fp32_groups.append(fulls)
            group["params"] = masters + fulls + frozen
        if torch.cuda.is_available():
            self._flag = torch.zeros(1, dtype=torch.int32, device=torch.cuda.current_device())
        self.master_params_created = True

    def _pairs(self):
        for halfs, masters in zip(self.fp16_groups, self.fp32_from_fp16_groups):
            yield from zip(halfs, masters)

    # ------------------------------------------------------------------ gradients
    def zero_grad(self, set_grads_to_None=True):
        if self.model is not None and hasattr(self.model, "zero_grad_buffer"):
            self.model.zero_grad_buffer()
        for group in self.optimizer.param_groups:
            for p in group["params"]:
                if getattr(p, "main_grad", None) is None:
                    p.grad = None
        for halfs in self.fp16_groups:
            for p in halfs:
                p.grad = None

    def backward(self, loss, update_master_grads=True, retain_graph=False):
        self.loss_scaler.backward(loss.float(), retain_graph=retain_graph, model=self.model)
        if update_master_grads:
            self.update_master_grads()

    def _master_grad_lists(self):
        """(masters, fp32 grads) for 16-bit params: the grad is the model param's fp32 ``main_grad`` when
        the DistributedModel installed one, else an fp32 copy of ``.grad``."""
        masters, grads = [], []
        for p, m in self._pairs():
            g = getattr(p, "main_grad", None)
            if g is None:
                if p.grad is None:
                    continue
                g = p.grad.float()
            masters.append(m)
            grads.append(g)
        return masters, grads

    def update_master_grads(self):
        """Detect overflow in this step's gradients and expose them (still scaled) as ``master.grad``."""
        self.init_master_params()
        masters, grads = self._master_grad_lists()
        fulls = [p for grp in self.fp32_from_fp32_groups for p in grp if p.grad is not None]
        all_grads = grads + [p.grad for p in fulls]
        overflow = False
        if self.dynamic_loss_scale and all_grads:
            if all_grads[0].is_cuda:
                norm, _ = MT.multi_tensor_l2norm(MT.DEFAULT_CHUNK, None, [all_grads], False)
                overflow = ~torch.isfinite(norm)
            else:
                overflow = any(not torch.isfinite(g).all() for g in all_grads)
        self.overflow = sync_overflow(overflow) if self.dynamic_loss_scale else False
        for m, g in zip(masters, grads):
            m.grad = g
        return self.overflow

    def clip_master_grads(self, max_norm, norm_type=2):
        """Clip the global gradient norm; returns the (unscaled) norm, or -1 on overflow."""
        if self.overflow:
            return -1
        self.init_master_params()
        pairs = [(m, m.grad) for _, m in self._pairs() if m.grad is not None]
        pairs += [(p, p.grad) for grp in self.fp32_from_fp32_groups for p in grp if p.grad is not None]
        total, coef = self.clip_grad_norm(pairs, max_norm * self.loss_scale, norm_type)
        self._pending_clip = coef.float().reshape(1).contiguous()
        return total / self.loss_scale

    # ------------------------------------------------------------------ update
    def _is_adam_like(self):
        opt = self.optimizer
        name = type(opt).__name__
        return name in ("Adam", "AdamW", "FusedAdam") and not any(g.get("amsgrad", False) for g in opt.param_groups)

    def step(self, closure=None):
        self.init_master_params()
        if closure is not None:
            raise NotImplementedError("closures are not supported by Bit16_Optimizer.step")
        scale = self.loss_scaler.loss_scale
        self.loss_scaler.update_scale(self.overflow)
        if self.overflow:
            get_logger().info(f"OVERFLOW! Skipping step. Attempted loss scale: {scale}, reducing to "
                              f"{self.loss_scaler.loss_scale}")
            self._pending_clip = None
            return None
        inv = 1.0 / scale
        clip = self._pending_clip
        self._pending_clip = None
        if self._is_adam_like():
            self._fused_adam_step(inv, clip)
        else:
            self._generic_step(inv, clip)
        self._step_count += 1
        return None

    def _fused_adam_step(self, inv_scale, clip):
        opt = self.optimizer
        adamw = type(opt).__name__ == "AdamW" or getattr(opt, "adam_w_mode", 0) == 1
        for gi, group in enumerate(opt.param_groups):
            lists16 = [[], [], [], [], []]
            lists32 = [[], [], [], []]
            step = None
            for p in group["params"]:
                if not p.requires_grad or p.grad is None:
                    continue
                st = opt.state[p]
                if "exp_avg" not in st:
                    st["step"] = torch.tensor(0.0) if type(opt).__name__ != "FusedAdam" else 0
                    st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
                    st["exp_avg_sq"] = torch.zeros_like(p, dtype=torch.float32)
                st["step"] = st["step"] + 1
                step = int(st["step"])
                low = self._half_of.get(id(p))
                if low is not None:
                    for lst, t in zip(lists16, (p.grad, p.data, st["exp_avg"], st["exp_avg_sq"], low.data)):
                        lst.append(t)
                else:
                    for lst, t in zip(lists32, (p.grad, p.data, st["exp_avg"], st["exp_avg_sq"])):
                        lst.append(t)
            b1, b2 = group["betas"]
            args = (group["lr"], b1, b2, group["eps"], step or 1, 1 if adamw else 0, group.get("bias_correction", True),
                    group["weight_decay"], inv_scale, clip)
            if lists16[0]:
                MT.multi_tensor_adam(MT.DEFAULT_CHUNK, None, lists16, *args)
            if lists32[0]:
                MT.multi_tensor_adam(MT.DEFAULT_CHUNK, None, lists32, *args)

    def _generic_step(self, inv_scale, clip):
        factor = inv_scale * (float(clip.item()) if clip is not None else 1.0)
        grads = [p.grad for g in self.optimizer.param_groups for p in g["params"] if p.grad is not None]
        owned = []
        for g in grads:
            # main_grad views are shared with the model: scale in place (they are zeroed next step anyway)
            owned.append(g)
        if factor != 1.0 and owned:
            MT.multi_tensor_scale(MT.DEFAULT_CHUNK, None, [owned, owned], factor)
        self.optimizer.step()
        halfs = [h for h, _ in self._pairs()]
        masters = [m.data for _, m in self._pairs()]
        if halfs:
            MT.multi_tensor_scale(MT.DEFAULT_CHUNK, None, [masters, [h.data for h in halfs]], 1.0)

    # ------------------------------------------------------------------ properties / state
    @property
    def loss_scale(self):
        return self.loss_scaler.loss_scale

    @loss_scale.setter
    def loss_scale(self, v):
        self.loss_scaler.cur_scale = v

    @property
    def param_groups(self):
        return self.optimizer.param_groups

    @property
    def state(self):
        return self.optimizer.state

    def state_dict(self):
        self.init_master_params()
        return {
            "loss_scaler": self.loss_scaler.state_dict(),
            "dynamic_loss_scale": self.dynamic_loss_scale,
            "overflow": self.overflow,
            "first_closure_call_this_step": self.first_closure_call_this_step,
            "optimizer_state_dict": self.optimizer.state_dict(),
            "fp32_from_fp16": [[m.detach().clone() for m in grp] for grp in self.fp32_from_fp16_groups],
        }

    def load_state_dict(self, sd):
        self.init_master_params()
        self.loss_scaler.load_state_dict(sd["loss_scaler"])
        self.dynamic_loss_scale = sd["dynamic_loss_scale"]
        self.overflow = sd["overflow"]
        self.first_closure_call_this_step = sd["first_closure_call_this_step"]
        if "optimizer_state_dict" in sd:
            self.optimizer.load_state_dict(sd["optimizer_state_dict"])
        for cur, saved in zip(self.fp32_from_fp16_groups, sd["fp32_from_fp16"]):
            for c, s in zip(cur, saved):
                c.data.copy_(s)
        self.copy_master_to_model()

    # ------------------------------------------------------------------ inspection (reference fp16.py:242-807)
    def get_master_params(self):
        """fp32 master copies, one list per parameter group."""
        self.init_master_params()
        return [list(grp) for grp in self.fp32_from_fp16_groups]

    def get_master_grads(self):
        """fp32 gradients of the masters (still multiplied by the loss scale), aligned with the masters."""
        masters, grads = self._master_grad_lists()
        by_master = {id(m): g for m, g in zip(masters, grads)}
        return [[by_master.get(id(m)) for m in grp] for grp in self.fp32_from_fp16_groups]

    def inspect_master_grad_data(self):
        """Unscaled fp32 gradients, ``None`` where a parameter has none; ``None`` altogether after an overflow
        (the gradients are garbage then)."""
        if self.overflow:
            return None
        inv = 1.0 / float(self.loss_scale)
        return [[None if g is None else g * inv for g in grp] for grp in self.get_master_grads()]

    def reload_model_params(self):
        """The 16-bit model weights were changed from outside (a checkpoint load): refresh the masters."""
        self.init_master_params()
        self.copy_model_to_master()

    def state_dict_for_save_checkpoint(self):
        return self.state_dict()

    def copy_master_to_model(self):
        pairs = list(self._pairs())
        if pairs:
            MT.multi_tensor_scale(MT.DEFAULT_CHUNK, None, [[m.data for _, m in pairs], [h.data for h, _ in pairs]], 1.0)

    def copy_model_to_master(self):
        for h, m in self._pairs():
            m.data.copy_(h.data)
