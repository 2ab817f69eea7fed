"""Optimizer-state sharding (ZeRO-1 style) across the data-parallel ranks.

Reference behaviour (reference smdistributed/modelparallel/torch/model.py:790-823 contiguous parameter
buffer ordered by reducer offsets, :1237-1340 "virtual parameters" = each rank's 1/N slice of that buffer
paired with the matching slice of the reducer's gradient buffer; torch/optimizers/optimizer.py:355-391
``_all_gather_base`` of the parameters after ``step``): every data-parallel rank keeps optimizer state
for, and updates, only its slice; the updated slices are all-gathered back into the full parameters.

Here:
* parameters live in one flat 16-bit (or fp32) buffer with exactly the layout of the fp32 gradient
  buffer (``GradBuffer`` segments = (reducer region, optimizer param group), each padded to a multiple
  of the group size);
* gradients are **reduce-scattered** in place (each rank receives only its slice, half the traffic of
  the reference's all-reduce);
* the slice update is one multi-tensor Adam launch that applies unscale x clip, updates the fp32 master
  slice and writes the 16-bit slice straight into the flat parameter buffer;
* the flat parameter buffer is all-gathered in place.
"""

from __future__ import annotations

import torch
import torch.distributed as dist

from ..backend.exceptions import MultipleDtypeOptShardingError, SMPUnsupportedError
from ..nn.utils import is_scaled_batch_parameter
from ..ops import multi_tensor as MT
from ..precision.loss_scaler import sync_overflow
from .grad_buffer import GradBuffer


class OptimizerStateSharding:
    def __init__(self, model, optimizer, bit16=None):
        from ..state_mod import state

        self.state = state
        self.model = model
        self.optimizer = optimizer
        self.bit16 = bit16
        core = state.core
        self.groups = {"scaled_batch": (state.rdp_process_group, core.rdp_size(), core.rdp_rank()),
                       "default": (state.dp_process_group, core.dp_size(), core.dp_rank())}
        self.dp_size = core.dp_size()
        if bit16 is not None:
            # undo the per-parameter master copies: this class owns flat master slices instead
            for group, halfs, fulls in zip(optimizer.param_groups, bit16.fp16_groups, bit16.fp32_from_fp32_groups):
                group["params"] = list(halfs) + list(fulls)
            optimizer.state.clear()
            # ... and drop them: a full fp32 copy per rank is exactly what state sharding is meant to avoid
            bit16.fp32_from_fp16_groups = [[] for _ in bit16.fp32_from_fp16_groups]
            bit16._master_of.clear()
            bit16._half_of.clear()
        params = []
        group_index = {}
        for gi, group in enumerate(optimizer.param_groups):
            for p in group["params"]:
                if p.requires_grad:
                    params.append(p)
                    group_index[p] = gi
        dtypes = {p.dtype for p in params}
        if len(dtypes) > 1:
            raise MultipleDtypeOptShardingError()
        self.param_dtype = dtypes.pop() if dtypes else torch.float32
        dev = state.device
        pad = max(core.dp_size(), core.rdp_size()) * 8
        # rebuild the gradient buffer group-aware and a parameter buffer with the same layout
        if model.grad_buffer is not None:
            model.grad_buffer.release()
        self.grad_buffer = GradBuffer(params, dev, pad_to=pad, group_of=lambda p: group_index[p])
        model.grad_buffer = self.grad_buffer
        if model.reducer is not None:
            model.reducer.grad_buffer = self.grad_buffer
            model.reducer.require_sync = False      # this class reduces (reduce-scatter) instead
        self.param_buffer = torch.zeros(self.grad_buffer.numel, dtype=self.param_dtype, device=dev)
        for p in self.grad_buffer.params:
            o = self.grad_buffer.offsets[p]
            view = self.param_buffer[o:o + p.numel()].view(p.shape)
            view.copy_(p.data)
            p.data = view
        # virtual parameters: my slice of every (region, param-group) segment
        self.slices = []          # dicts: region, gi, start, end (absolute offsets of MY slice), master, m, v
        new_groups = [[] for _ in optimizer.param_groups]
        for (region, gi), (a, b) in sorted(self.grad_buffer.segments.items(), key=lambda kv: kv[1][0]):
            if b <= a:
                continue
            _, n, r = self.groups[region]
            per = (b - a) // n
            s, e = a + r * per, a + (r + 1) * per
            master = torch.nn.Parameter(self.param_buffer[s:e].float().clone(), requires_grad=True)
            master.grad = self.grad_buffer.buffer[s:e]
            if region == "scaled_batch":
                master._smp_scaled_batch = True
            self.slices.append({"region": region, "gi": gi, "seg": (a, b), "start": s, "end": e, "master": master})
            new_groups[gi].append(master)
        for group, plist in zip(optimizer.param_groups, new_groups):
            group["params"] = plist
        self.overflow = False
        self._clip = None
        self._step = 0

    # ------------------------------------------------------------------ gradient reduction
    def reduce_gradients(self):
        """Reduce-scatter every segment in place over its group and average."""
        scale = 1.0 / self.dp_size
        buf = self.grad_buffer.buffer
        for sl in self.slices:
            group, n, r = self.groups[sl["region"]]
            a, b = sl["seg"]
            if n > 1:
                dist.reduce_scatter_tensor(buf[sl["start"]:sl["end"]], buf[a:b], group=group)
        if scale != 1.0:
            for sl in self.slices:
                buf[sl["start"]:sl["end"]].mul_(scale)

    def _local_grads(self):
        return [self.grad_buffer.buffer[sl["start"]:sl["end"]] for sl in self.slices]

    def check_overflow(self):
        dyn = self.bit16 is not None and self.bit16.dynamic_loss_scale
        if not dyn:
            self.overflow = False
            return False
        grads = self._local_grads()
        flag = False
        if grads:
            norm, _ = MT.multi_tensor_l2norm(MT.DEFAULT_CHUNK, None, [grads], False)
            flag = ~torch.isfinite(norm)
        self.overflow = sync_overflow(flag)
        if self.bit16 is not None:
            self.bit16.overflow = self.overflow
        return self.overflow

    def clip_grad_norm(self, max_norm, norm_type=2):
        if self.overflow:
            return -1
        scale = self.bit16.loss_scale if self.bit16 is not None else 1.0
        grads = self._local_grads()
        dev = self.state.device
        if grads:
            norm, _ = MT.multi_tensor_l2norm(MT.DEFAULT_CHUNK, None, [grads], False)
            sq = norm.float() ** 2
        else:
            sq = torch.zeros(1, device=dev)
        if self.state.core.size() > 1:
            dist.all_reduce(sq, group=self.state.world_process_group)
        total = sq.sqrt()
        self._clip = ((max_norm * scale) / (total + 1e-6)).clamp(max=1.0).float().reshape(1).contiguous()
        return total / scale

    # ------------------------------------------------------------------ update
    def step(self):
        b16 = self.bit16
        scale = b16.loss_scale if b16 is not None else 1.0
        if b16 is not None:
            b16.loss_scaler.update_scale(self.overflow)
        if self.overflow:
            self._clip = None
            return None
        inv = 1.0 / scale
        clip, self._clip = self._clip, None
        opt = self.optimizer
        name = type(opt).__name__
        adam_like = name in ("Adam", "AdamW", "FusedAdam")
        self._step += 1
        if adam_like:
            adamw = name == "AdamW" or getattr(opt, "adam_w_mode", 0) == 1
            pending = []
            for sl in self.slices:
                group = opt.param_groups[sl["gi"]]
                m = sl["master"]
                st = opt.state[m]
                if "exp_avg" not in st:
                    st["step"] = torch.tensor(0.0)
                    st["exp_avg"] = torch.zeros_like(m.data)
                    st["exp_avg_sq"] = torch.zeros_like(m.data)
                st["step"] = st["step"] + 1
                lists = [[m.grad], [m.data], [st["exp_avg"]], [st["exp_avg_sq"]]]
                if self.param_dtype != torch.float32:
                    lists.append([self.param_buffer[sl["start"]:sl["end"]]])
                b1, b2 = group["betas"]
                MT.multi_tensor_adam(MT.DEFAULT_CHUNK, None, lists, group["lr"], b1, b2, group["eps"], int(st["step"]),
                                     1 if adamw else 0, group.get("bias_correction", True), group["weight_decay"],
                                     inv, clip)
                if self.param_dtype == torch.float32:
                    self.param_buffer[sl["start"]:sl["end"]].copy_(m.data)
                # this slice is final: its all-gather starts now and runs under the update of the next slices
                self._gather_slice(sl, pending)
            for w in pending:
                w.wait()
            return None
        else:
            factor = inv * (float(clip.item()) if clip is not None else 1.0)
            if factor != 1.0:
                for sl in self.slices:
                    sl["master"].grad.mul_(factor)
            opt.step()
            for sl in self.slices:
                self.param_buffer[sl["start"]:sl["end"]].copy_(sl["master"].data)
        self.all_gather_params()
        return None

    def _gather_slice(self, sl, pending):
        """Asynchronous in-place all-gather of one updated slice's segment (reference torch/optimizers/optimizer.py:
        355-391 gathers after the whole update; here the gather of slice i overlaps the Adam kernel of slice i+1)."""
        group, n, r = self.groups[sl["region"]]
        if n > 1:
            a, b = sl["seg"]
            pending.append(dist.all_gather_into_tensor(self.param_buffer[a:b], self.param_buffer[sl["start"]:sl["end"]],
                                                       group=group, async_op=True))

    def all_gather_params(self):
        """Rebuild the full parameters from the updated slices (in-place all-gather per segment)."""
        for sl in self.slices:
            group, n, r = self.groups[sl["region"]]
            if n > 1:
                a, b = sl["seg"]
                dist.all_gather_into_tensor(self.param_buffer[a:b], self.param_buffer[sl["start"]:sl["end"]], group=group)

    def refresh_masters(self):
        """The model's parameters were overwritten from outside (``model.load_state_dict`` after the optimizer was
        built): take the fp32 master slices from them again, as ``copy_model_to_master`` does for per-parameter
        masters - otherwise the next step would update, and all-gather, the values from before the load."""
        for sl in self.slices:
            sl["master"].data.copy_(self.param_buffer[sl["start"]:sl["end"]])

    # ------------------------------------------------------------------ checkpointing
    def virtual_named_parameters(self):
        """(name, fp32 master slice) for every slice of the flat buffer this rank owns - the tensors the inner
        optimizer updates (the reference's "virtual parameters", model.py:1237-1340)."""
        for i, sl in enumerate(self.slices):
            yield f"{sl['region']}.group{sl['gi']}.slice{i}[{sl['start']}:{sl['end']}]", sl["master"]

    def shard_info(self):
        return {"slices": [{"region": s["region"], "group": s["gi"], "start": s["start"], "end": s["end"]}
                           for s in self.slices],
                "masters": [s["master"].detach().cpu() for s in self.slices]}

    def load_shard_info(self, info):
        for s, m in zip(self.slices, info["masters"]):
            s["master"].data.copy_(m.to(s["master"].device))
            self.param_buffer[s["start"]:s["end"]].copy_(s["master"].data)
        self.all_gather_params()
