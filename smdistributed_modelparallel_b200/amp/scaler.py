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

    def _local_device(self):
        from ..state_mod import state

        if getattr(state, "initialized", False):
            return state.device
        return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")

    def _ensure_scale(self):
        # only the pipeline leader runs the step function, so only it has called scale(); the other stages
        # start from the same initial scale and stay in step through the shared verdict
        # (reference amp/scaler.py: lazy initialisation for pp_rank != 0 in step())
        if self._enabled and self._scale is None:
            self._lazy_init_scale_growth_tracker(self._local_device())

    def _sync_found_inf(self, optimizer_state):
        """One verdict for every rank that holds part of the model - exactly one collective per rank, also on
        ranks that own no gradient at all."""
        from ..state_mod import state

        found = optimizer_state["found_inf_per_device"]
        dev = self._local_device()
        total = torch.zeros(1, dtype=torch.float32, device=dev)
        for t in found.values():
            total += t.to(dev).reshape(-1)[:1]
        if getattr(state, "initialized", False) and state.core.mp_size() > 1:
            dist.all_reduce(total, op=dist.ReduceOp.MAX, group=state.mp_process_group)
        if not found:
            found[dev] = total
        for d, t in found.items():
            t.copy_(total.to(d).reshape(t.shape))

    def unscale_(self, optimizer):
        self._ensure_scale()
        inner = getattr(optimizer, "optimizer", optimizer)
        super().unscale_(inner)
        self._sync_found_inf(self._per_optimizer_states[id(inner)])

    def update(self, new_scale=None):
        self._ensure_scale()
        return super().update(new_scale)

    def step(self, optimizer, *args, **kwargs):
        if not self._enabled:
            return optimizer.step(*args, **kwargs)
        if "closure" in kwargs:
            raise SMPAMPError("Closure use is not supported with smp.amp.GradScaler.")
        self._ensure_scale()
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
