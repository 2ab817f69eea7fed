"""Static and dynamic loss scalers for fp16 training.

Parameters and update rule follow the reference (reference smdistributed/modelparallel/torch/fp16/
loss_scaler.py:33-206): the dynamic scaler starts at 2**32, halves on overflow (subject to
``delayed_shift`` hysteresis), doubles after ``scale_window`` overflow-free iterations, never drops
below ``min_scale``.  Overflow detection itself is done on the device by the fused l2norm/scale
kernels and synchronised across the model-parallel group (or the world when optimizer state is
sharded), reference :131-142.
"""

from __future__ import annotations

import torch
import torch.distributed as dist


class LossScalerBase:
    def __init__(self, scale):
        self.cur_scale = float(scale)
        self.dynamic = False
        self.model = None         # the reference's scalers carry the model; kept so pickled objects have the field

    @property
    def loss_scale(self):
        return self.cur_scale

    def update_scale(self, overflow):
        pass

    def has_overflow(self, params):
        return False

    def backward(self, loss, retain_graph=False, model=None):
        from ..state_mod import state
        (model or state.model).backward(loss * self.loss_scale)

    def state_dict(self):
        return {k: v for k, v in self.__dict__.items() if k != "model"}

    def load_state_dict(self, sd):
        """Accepts this library's dict form or a scaler object as the reference pickles it."""
        if not isinstance(sd, dict):
            sd = dict(vars(sd))
        self.__dict__.update({k: v for k, v in sd.items() if k != "model"})

    def for_checkpoint(self):
        """A copy without the model reference, picklable under the reference's class path."""
        import copy

        model, self.model = self.model, None
        try:
            return copy.deepcopy(self)
        finally:
            self.model = model


class LossScaler(LossScalerBase):
    """Fixed loss scale."""

    def __init__(self, scale=1.0):
        super().__init__(scale)


class DynamicLossScaler(LossScalerBase):
    def __init__(self, init_scale=2 ** 32, scale_factor=2.0, scale_window=1000, min_scale=1, delayed_shift=1,
                 consecutive_hysteresis=False):
        super().__init__(init_scale)
        self.dynamic = True
        self.cur_iter = 0
        self.last_overflow_iter = -1
        self.scale_factor = scale_factor
        self.scale_window = scale_window
        self.min_scale = min_scale
        self.delayed_shift = delayed_shift
        self.cur_hysteresis = delayed_shift
        self.consecutive_hysteresis = consecutive_hysteresis

    @staticmethod
    def _has_inf_or_nan(x: torch.Tensor) -> bool:
        s = float(x.float().sum())
        return s != s or s in (float("inf"), -float("inf"))

    def has_overflow_serial(self, params):
        return any(p.grad is not None and self._has_inf_or_nan(p.grad) for p in params)

    def has_overflow(self, params):
        return sync_overflow(self.has_overflow_serial(params))

    def update_scale(self, overflow: bool):
        if overflow:
            if self.delayed_shift == 1 or self.cur_hysteresis == 1:
                self.cur_scale = max(self.cur_scale / self.scale_factor, self.min_scale)
            else:
                self.cur_hysteresis -= 1
            self.last_overflow_iter = self.cur_iter
        else:
            if self.consecutive_hysteresis:
                self.cur_hysteresis = self.delayed_shift
            if (self.cur_iter - self.last_overflow_iter) % self.scale_window == 0:
                if not self.consecutive_hysteresis:
                    self.cur_hysteresis = self.delayed_shift
                self.cur_scale *= self.scale_factor
        self.cur_iter += 1


# fp16_states_*.pt pickles the scaler OBJECT (reference torch/optimizers/optimizer.py:185-190); its class must
# resolve at the reference's module path for either library to load the other's file.
for _cls in (LossScaler, DynamicLossScaler):
    _cls.__module__ = "smdistributed.modelparallel.torch.fp16.loss_scaler"


def overflow_sync_group():
    """Group over which an overflow / grad-norm must agree: the model-parallel group, or the whole
    world when optimizer state is sharded (every rank then updates a slice of every parameter)."""
    from ..state_mod import state

    if not getattr(state, "initialized", False):
        return None
    if state.cfg.shard_optimizer_state or state.cfg.zero2d_enabled():
        return state.world_process_group
    return state.mp_process_group


def sync_overflow(flag) -> bool:
    """MAX-reduce an overflow indicator (python bool or 1-element tensor) over the sync group."""
    from ..state_mod import state

    group = overflow_sync_group()
    if group is None or dist.get_world_size(group) == 1:
        return bool(flag.item() if torch.is_tensor(flag) else flag)
    t = flag.to(torch.float32).reshape(1).clone() if torch.is_tensor(flag) else \
        torch.tensor([float(flag)], device=state.device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return bool(t.item() > 0)
