"""Fused optimizers on the multi-tensor kernels: FusedAdam, FusedLAMB, FusedNovoGrad, FusedSGD,
FusedAdagrad.

Same constructor arguments and update rules as the Apex optimizers the reference vendors (reference
smdistributed/modelparallel/torch/apex/optimizers/fused_adam.py:63-173, fused_sgd.py, fused_adagrad.py,
fused_novograd.py) and its model-parallel aware LAMB, whose global gradient norm is combined across the
pipeline group before clipping (reference torch/optimizers/fused_lamb.py:14-139).
"""

from __future__ import annotations

import torch
import torch.distributed as dist

from ..backend.exceptions import FusedLAMBError
from ..ops import multi_tensor as MT


def _by_dtype(params):
    out = {}
    for p in params:
        out.setdefault((p.dtype, p.grad.dtype), []).append(p)
    return out


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, adam_w_mode=True,
                 weight_decay=0.0, amsgrad=False, set_grad_none=True):
        if amsgrad:
            raise RuntimeError("FusedAdam does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay))
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.set_grad_none = set_grad_none

    def zero_grad(self, set_to_none=None):
        super().zero_grad(set_to_none=self.set_grad_none if set_to_none is None else set_to_none)

    @torch.no_grad()
    def step(self, closure=None, grad_scale=1.0):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            group["step"] = group.get("step", 0) + 1
            ps = [p for p in group["params"] if p.grad is not None]
            for p in ps:
                st = self.state[p]
                if not st:
                    st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
                    st["exp_avg_sq"] = torch.zeros_like(p, dtype=torch.float32)
            b1, b2 = group["betas"]
            for _, plist in _by_dtype(ps).items():
                lists = [[p.grad for p in plist], [p.data for p in plist], [self.state[p]["exp_avg"] for p in plist],
                         [self.state[p]["exp_avg_sq"] for p in plist]]
                MT.multi_tensor_adam(MT.DEFAULT_CHUNK, None, lists, group["lr"], b1, b2, group["eps"], group["step"],
                                     self.adam_w_mode, group["bias_correction"], group["weight_decay"], grad_scale)
        return loss


class FusedSGD(torch.optim.Optimizer):
    def __init__(self, params, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False,
                 wd_after_momentum=False, set_grad_none=False):
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                                      nesterov=nesterov))
        self.wd_after_momentum = wd_after_momentum

    @torch.no_grad()
    def step(self, closure=None, scale=1.0):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            ps = [p for p in group["params"] if p.grad is not None]
            for first, plist in ((True, [p for p in ps if "momentum_buffer" not in self.state[p]]),
                                 (False, [p for p in ps if "momentum_buffer" in self.state[p]])):
                if not plist:
                    continue
                for p in plist:
                    if first:
                        self.state[p]["momentum_buffer"] = torch.zeros_like(p, dtype=torch.float32)
                for _, pl in _by_dtype(plist).items():
                    lists = [[p.grad for p in pl], [p.data for p in pl], [self.state[p]["momentum_buffer"] for p in pl]]
                    MT.multi_tensor_sgd(MT.DEFAULT_CHUNK, None, lists, group["weight_decay"], group["momentum"],
                                        group["dampening"], group["lr"], group["nesterov"], first,
                                        self.wd_after_momentum, scale)
        return loss


class FusedAdagrad(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-2, eps=1e-10, weight_decay=0.0, set_grad_none=True, adagrad_w_mode=False):
        super().__init__(params, dict(lr=lr, eps=eps, weight_decay=weight_decay))
        self.adagrad_w_mode = 1 if adagrad_w_mode else 0

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            ps = [p for p in group["params"] if p.grad is not None]
            for p in ps:
                if not self.state[p]:
                    self.state[p]["sum"] = torch.zeros_like(p, dtype=torch.float32)
            for _, pl in _by_dtype(ps).items():
                lists = [[p.grad for p in pl], [p.data for p in pl], [self.state[p]["sum"] for p in pl]]
                MT.multi_tensor_adagrad(MT.DEFAULT_CHUNK, None, lists, group["lr"], group["eps"], self.adagrad_w_mode,
                                        group["weight_decay"])
        return loss


class FusedLAMB(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.01,
                 amsgrad=False, adam_w_mode=True, grad_averaging=True, set_grad_none=True, max_grad_norm=1.0,
                 use_nvlamb=False):
        if amsgrad:
            raise FusedLAMBError("FusedLAMB does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, grad_averaging=grad_averaging,
                                      max_grad_norm=max_grad_norm))
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.use_nvlamb = use_nvlamb

    def _global_grad_norm(self, grads):
        """L2 norm over every gradient of the model: local norm, then squared-sum all-reduce over the
        model-parallel group so pipeline / tensor partitions clip consistently."""
        from ..state_mod import state

        if not grads:
            return None
        norm, _ = MT.multi_tensor_l2norm(MT.DEFAULT_CHUNK, None, [grads], False)
        if getattr(state, "initialized", False) and state.core.mp_size() > 1:
            sq = norm.float() ** 2
            dist.all_reduce(sq, group=state.mp_process_group)
            norm = sq.sqrt()
        return norm

    @torch.no_grad()
    def step(self, closure=None, grad_scale=1.0):
        loss = closure() if closure is not None else None
        all_grads = [p.grad for g in self.param_groups for p in g["params"] if p.grad is not None]
        gnorm = self._global_grad_norm(all_grads)
        for group in self.param_groups:
            group["step"] = group.get("step", 0) + 1
            ps = [p for p in group["params"] if p.grad is not None]
            for p in ps:
                st = self.state[p]
                if not st:
                    st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
                    st["exp_avg_sq"] = torch.zeros_like(p, dtype=torch.float32)
            b1, b2 = group["betas"]
            for _, pl in _by_dtype(ps).items():
                lists = [[p.grad for p in pl], [p.data for p in pl], [self.state[p]["exp_avg"] for p in pl],
                         [self.state[p]["exp_avg_sq"] for p in pl]]
                MT.multi_tensor_lamb(MT.DEFAULT_CHUNK, None, lists, group["lr"], b1, b2, group["eps"], group["step"],
                                     group["bias_correction"], group["weight_decay"], group["grad_averaging"],
                                     self.adam_w_mode, gnorm, group["max_grad_norm"], self.use_nvlamb, grad_scale)
        return loss


class FusedNovoGrad(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0,
                 amsgrad=False, reg_inside_moment=False, grad_averaging=True, norm_type=2, init_zero=False,
                 set_grad_none=True):
        if amsgrad:
            raise RuntimeError("FusedNovoGrad does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, grad_averaging=grad_averaging, norm_type=norm_type,
                                      init_zero=init_zero))
        self.moment_mode = 0 if reg_inside_moment else 1

    @torch.no_grad()
    def step(self, closure=None, grad_scale=1.0):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            group["step"] = group.get("step", 0) + 1
            ps = [p for p in group["params"] if p.grad is not None]
            for p in ps:
                if not self.state[p]:
                    self.state[p]["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
            b1, b2 = group["betas"]
            for key, pl in _by_dtype(ps).items():
                nk = f"exp_avg_sq_{key}"
                if nk not in group or group[nk].numel() != len(pl):
                    group[nk] = torch.zeros(len(pl), dtype=torch.float32, device=pl[0].device)
                    first = not group["init_zero"]
                else:
                    first = False
                lists = [[p.grad for p in pl], [p.data for p in pl], [self.state[p]["exp_avg"] for p in pl]]
                MT.multi_tensor_novograd(MT.DEFAULT_CHUNK, None, lists, group[nk], group["lr"], b1, b2, group["eps"],
                                         1 if first else max(group["step"], 2), group["bias_correction"],
                                         group["weight_decay"], group["grad_averaging"], self.moment_mode,
                                         group["norm_type"], grad_scale)
        return loss
