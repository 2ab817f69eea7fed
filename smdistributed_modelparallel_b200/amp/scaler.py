"""``smp.amp.GradScaler``: torch.amp's GradScaler whose inf/nan verdict is agreed on by every rank that
holds part of the model, so all pipeline / tensor partitions skip (or take) the step together
(reference smdistributed/modelparallel/torch/amp/scaler.py:22-194 - found_inf all-reduced over the
pipeline group before ``optimizer.step``).
"""

from __future__ import annotations

import torch
import torch.distributed as dist

from ..backend.exceptions import SMPAMPError


class GradScaler(torch.amp.GradScaler):
    def __init__(self, init_scale=2.0 ** 16, growth_factor=2.0, backoff_factor=0.5, growth_interval=2000, enabled=True):
        device = "cuda" if torch.cuda.is_available() else "cpu"
        super().__init__(device, init_scale=init_scale, growth_factor=growth_factor, backoff_factor=backoff_factor,
                         growth_interval=growth_interval, enabled=enabled)

    def _sync_found_inf(self, optimizer_state):
        from ..state_mod import state

        if not getattr(state, "initialized", False) or state.core.mp_size() == 1:
            return
        for dev, t in optimizer_state["found_inf_per_device"].items():
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=state.mp_process_group)

    def unscale_(self, optimizer):
        inner = getattr(optimizer, "optimizer", optimizer)
        super().unscale_(inner)
        self._sync_found_inf(self._per_optimizer_states[id(inner)])

    def step(self, optimizer, *args, **kwargs):
        if not self._enabled:
            return optimizer.step(*args, **kwargs)
        if "closure" in kwargs:
            raise SMPAMPError("Closure use is not supported with smp.amp.GradScaler.")
        inner = getattr(optimizer, "optimizer", optimizer)
        st = self._per_optimizer_states[id(inner)]
        from torch.amp.grad_scaler import OptState

        if st["stage"] is OptState.READY:
            self.unscale_(optimizer)
        found = sum(v.item() for v in st["found_inf_per_device"].values())
        ret = None
        if not found:
            ret = optimizer.step(*args, **kwargs)
        st["stage"] = OptState.STEPPED
        return ret
