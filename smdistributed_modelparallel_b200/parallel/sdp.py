"""Sharded data parallelism: a ZeRO-3 style engine (parameters, gradients and optimizer state all
sharded across ``sharded_data_parallel_degree`` ranks).

The reference exposes the feature but delegates the engine to an AWS-patched DeepSpeed that is not part
of its tree (reference smdistributed/modelparallel/torch/optimizers/optimizer.py:46-49,471-491;
backend/zero_config.py:13-131; knobs in backend/ds_config_defaults.json:6-27; call sites
torch/model.py:1136-1144,1201-1202; on-disk layout torch/checkpoint.py:300-319).  This is an independent
engine with the same observable contract:

* tp = pp = 1; the shard group is ``degree`` consecutive ranks, replica groups connect equal shard
  indices; every rank feeds its own batch;
* parameters smaller than ``sdp_param_persistence_threshold`` stay replicated; the rest are grouped
  into *units* (the largest modules below the prefetch bucket size, e.g. one transformer layer), each
  unit laid out flat, zero-padded to a multiple of the degree, and only the local 1/degree slice kept;
* forward / backward of a unit all-gathers its parameters just in time (the next unit is prefetched
  on NCCL's stream while the current one computes) and frees them afterwards;
* weight gradients are accumulated in fp32 directly by the wgrad GEMM epilogue into a per-unit full
  buffer, reduce-scattered to the owning slices as soon as the unit's backward finishes, accumulated
  over microbatches, then all-reduced across replica groups;
* one fused multi-tensor Adam launch per unit applies unscale x clip (``sdp_gradient_clipping``),
  updates the fp32 master slice and writes the 16-bit slice;
* checkpoints are per shard rank: ``model_{shard}.pt`` / ``optimizer_{shard}.pt``.
"""

from __future__ import annotations

import os
from collections import OrderedDict

import torch
import torch.distributed as dist
import torch.nn as nn

from ..backend.exceptions import SMPValidationError, UnsupportedShardedConfigError
from ..backend.logger import get_logger
from ..ops import multi_tensor as MT
from ..precision.loss_scaler import DynamicLossScaler, LossScaler
from .boundary import map_tensors, mark_inputs
from .grad_buffer import GradBuffer


class Unit:
    """A set of parameters gathered / released together; stored flat and sharded."""

    def __init__(self, name, module, params, gi, degree, rank, device, striped=False):
        self.name, self.module, self.gi = name, module, gi
        self.degree, self.rank = degree, rank
        # the weight the unit's module multiplies first goes to offset 0: the fused kernel gathers the unit
        # from inside that GEMM (parallel/sdp_fused.py)
        first = getattr(module, "sdp_first_weight", lambda: None)() if striped else None
        self.first_weight = first if any(first is p for p in params) and first.dim() == 2 else None
        if self.first_weight is not None:
            params = [self.first_weight] + [p for p in params if p is not self.first_weight]
        self.params = params
        self.offsets = {}
        off = 0
        for p in params:
            self.offsets[p] = off
            off = (off + p.numel() + 7) // 8 * 8
        if striped:
            # piece = 8 rows of the first weight (or 8 x 4096 elements); piece g lives on rank g % degree
            self.row = self.first_weight.shape[1] if self.first_weight is not None else 4096
            if self.row % 8:
                self.row, self.first_weight = 4096, None
            self.piece = 8 * self.row
            mult = degree * self.piece
            self.numel = (off + mult - 1) // mult * mult
            self.per = self.numel // degree
        else:
            mult = degree * 8
            self.numel = (off + mult - 1) // mult * mult
            self.per = self.numel // degree
            self.piece, self.row = self.per, self.per       # one piece per rank: plain contiguous sharding
        self.ppr = self.per // self.piece                   # pieces per rank
        self.dtype = params[0].dtype
        self.shapes = {p: tuple(p.shape) for p in params}
        self.numels = {p: p.numel() for p in params}
        full = torch.zeros(self.numel, dtype=self.dtype, device=device)
        for p in params:
            full[self.offsets[p]:self.offsets[p] + p.numel()].copy_(p.data.reshape(-1))
        self.shard = self.shard_of(full).clone()
        self.master = nn.Parameter(self.shard.float(), requires_grad=True)
        self.grad_shard = torch.zeros(self.per, dtype=torch.float32, device=device)
        self.master.grad = self.grad_shard
        self.full = None            # gathered parameters (param dtype), None when released
        self.full_grad = None       # fp32 [numel] while the unit's backward is in flight
        self.gather_work = None
        self.users = 0
        self.lazy_pending = False   # gather deferred to the first GEMM reading first_weight
        self.recompute_seen = False # forward re-run during backward (activation checkpointing)
        self.updated = False        # Adam already applied by the fused reduce kernel this step
        for p in params:
            p.data = torch.empty(0, dtype=p.dtype, device=device)
            p._smp_sdp_unit = self

    def shard_of(self, full, rank=None):
        """This rank's (or ``rank``'s) elements of a full-unit tensor, in shard order."""
        r = self.rank if rank is None else rank
        return full.view(self.ppr, self.degree, self.piece)[:, r].reshape(-1)


class ShardedDataParallelEngine:
    def __init__(self, model, params):
        from ..state_mod import state

        self.state = state
        self.model = model
        cfg = state.cfg
        core = state.core
        if cfg.tensor_parallel_degree != 1 or cfg.pipeline_parallel_degree != 1:
            raise UnsupportedShardedConfigError("sharded data parallelism requires tensor and pipeline degree 1")
        z = cfg.zero2d_config_dict().get("zero_optimization", {})
        self.degree = int(z.get("zero2d_shard_size", cfg.sharded_data_parallel_degree))
        self.shard_group = state.sdp_process_group
        self.replica_group = state.sdp_replica_process_group
        self.shard_rank = core.rank() % self.degree
        self.world = core.size()
        self.threshold = int(z.get("stage3_param_persistence_threshold", cfg.sdp_param_persistence_threshold))
        # a unit is both the prefetch granule and the reduce-scatter granule, so it honours the smaller of the
        # two bucket knobs (sdp_reduce_bucket_size / stage3_prefetch_bucket_size, in elements)
        self.unit_cap = int(os.environ.get("SMP_SDP_UNIT_ELEMENTS",
                                           min(float(z.get("stage3_prefetch_bucket_size", 5e8)),
                                               float(z.get("reduce_bucket_size", cfg.sdp_reduce_bucket_size)))))
        # sdp_hierarchical_allgather (two-level gather for shard groups spanning nodes) has nothing to do inside one
        # NVSwitch domain, where every peer is one hop away; shard groups across nodes fall back to NCCL, which
        # picks its own hierarchy
        self.hierarchical = bool(z.get("zero2d_hierarchical_allgather", cfg.sdp_hierarchical_allgather))
        self.max_live = int(z.get("stage3_max_live_parameters", cfg.sdp_max_live_parameters))
        self.clip = float(cfg.zero2d_config_dict().get("gradient_clipping", cfg.sdp_gradient_clipping))
        self.device = state.device
        self.all_params = [p for p in params if p.requires_grad]
        self.units: list[Unit] = []
        self.persistent: list = []
        self.persistent_buffer = None
        self.persistent_masters = {}
        self.optimizer = None
        self.loss_scaler = LossScaler(1.0)
        self.dynamic = False
        self.overflow = False
        self._built = False
        self.gather_stats = {"prefetched": 0, "on_demand": 0}      # units gathered ahead of use / when first needed
        self._order = []             # unit execution order observed in the first forward
        self._order_frozen = False
        self._pending_reduce = []    # (work, tmp_shard, unit, full_grad)
        self._in_backward = False
        self._step = 0
        self._hooks = []
        self._sync_step = True
        self.fused = False           # peer-memory kernels instead of NCCL (CUDA, shard group on one node)
        self.fabric = None
        self._want_fused = os.environ.get("SMP_SDP_FUSED", "1") != "0"

    # ------------------------------------------------------------------ construction
    def _build(self, group_of=None):
        if self._built:
            return
        self._built = True
        group_of = group_of or (lambda p: 0)
        from . import sdp_fused
        self.fused = self._want_fused and sdp_fused.supported(self)
        root = self.model.module
        names = {m: n for n, m in root.named_modules()}
        seen = set()

        def is_sharded(p):
            return p.requires_grad and p.numel() >= self.threshold

        # Sharded initialisation (the reference wraps model creation in deepspeed.zero.Init for this,
        # torch/parameter.py:23-32): with delayed parameter initialisation the model arrives here on the meta
        # device and every unit is materialised, initialised, sharded and released in turn, so a rank never holds
        # more than one unit in full.  Parameters are tracked by (owning module, name) slots because
        # materialisation replaces the Parameter objects.
        lazy = any(q.is_meta for q in root.parameters())
        owners, slot_group = {}, {}

        def index_owners(mods):
            for mod in mods:
                for n, q in mod.named_parameters(recurse=False):
                    slots = owners.setdefault(id(q), [])
                    if (mod, n) not in slots:
                        slots.append((mod, n))

        index_owners(root.modules())
        for pid, slots in owners.items():
            q = slots[0][0]._parameters[slots[0][1]]
            for slot in slots:
                slot_group[slot] = group_of(q)

        def gof(q):
            return slot_group.get(owners[id(q)][0], 0) if id(q) in owners else group_of(q)

        def realise(plist):
            if not lazy or not any(q.is_meta for q in plist):
                return plist
            from ..parameter import materialize_delayed
            where = [owners[id(q)][0] for q in plist]
            mods = {owners[id(q)][0][0] for q in plist if q.is_meta}
            materialize_delayed(root, local_only=lambda mod: mod in mods, device=self.device)
            index_owners(root.modules())
            return [mod._parameters[n] for mod, n in where]

        def subtree_count(m):
            return sum(p.numel() for p in m.parameters() if is_sharded(p))

        def add_unit(m, plist, tag=""):
            plist = list({id(p): p for p in plist if is_sharded(p) and id(p) not in seen}.values())
            for gi in sorted({gof(p) for p in plist}):
                sub = realise([p for p in plist if gof(p) == gi and id(p) not in seen])
                for p in sub:
                    seen.add(id(p))
                u = Unit(f"{names[m] or 'root'}{tag}#{gi}", m, sub, gi, self.degree, self.shard_rank, self.device,
                         striped=self.fused)
                self.units.append(u)

        def assign(m):
            n = subtree_count(m)
            if n == 0:
                return
            has_param_children = any(subtree_count(c) > 0 for c in m.children())
            if n <= self.unit_cap or not has_param_children:
                add_unit(m, list(m.parameters()))
                return
            own = list(m.parameters(recurse=False)) + list(getattr(m, "sdp_directly_used_parameters", lambda: [])())
            if own:
                add_unit(m, own, ".own")
            for c in m.children():
                assign(c)

        assign(root)
        if lazy:
            from ..parameter import materialize_delayed
            materialize_delayed(root, device=self.device)      # what is left: small (persistent) parameters, buffers
            index_owners(root.modules())
            self.all_params = [p for p in root.parameters() if p.requires_grad]
        self._group_of_current = gof
        in_units = {id(p) for u in self.units for p in u.params}
        self.persistent = [p for p in self.all_params if id(p) not in in_units]
        by_module = OrderedDict()
        for u in self.units:
            by_module.setdefault(u.module, []).append(u)
        # A parameter shared with a module outside its unit's subtree (tied input / output embeddings of a stock
        # HuggingFace model): that module gathers and releases the unit around its own forward / backward as well,
        # like a second use of the unit's module.
        hooked = {}
        for u in self.units:
            hooked.setdefault(u, set()).add(names[u.module])
        for m in root.modules():
            mn = names[m]
            for q in m.parameters(recurse=False):
                u = getattr(q, "_smp_sdp_unit", None)
                if u is None or any(h == "" or mn == h or mn.startswith(h + ".") for h in hooked[u]):
                    continue
                by_module.setdefault(m, []).append(u)
                hooked[u].add(mn)
        for m, us in by_module.items():
            self._hooks.append(m.register_forward_pre_hook(lambda mod, a, _us=us: self._pre_forward(_us, a)))
            self._hooks.append(m.register_forward_hook(lambda mod, a, out, _us=us: self._post_forward(_us, out)))
            self._hooks.append(m.register_full_backward_pre_hook(lambda mod, g, _us=us: self._pre_backward(_us)))
            self._hooks.append(m.register_full_backward_hook(lambda mod, gi, go, _us=us: self._post_backward(_us, gi)))
        if self.persistent:
            self.persistent_buffer = GradBuffer(self.persistent, self.device)
            for p in self.persistent:
                self.persistent_masters[p] = nn.Parameter(p.detach().float().clone(), requires_grad=True)
            if self.world > 1:                      # replicas must start identical
                for p in self.persistent:
                    dist.broadcast(p.data, src=0)
                    self.persistent_masters[p].data.copy_(p.data)
        if self.fused and self.units:
            self.fabric = sdp_fused.PeerFabric(self)
        total = sum(u.numel for u in self.units)
        get_logger().debug(f"sharded DP: {len(self.units)} units, {total / 1e6:.1f} M sharded + "
                           f"{sum(p.numel() for p in self.persistent) / 1e6:.2f} M persistent parameters, degree {self.degree}")

    # ------------------------------------------------------------------ gather / release
    def _issue_gather(self, u: Unit, lazy_ok=False):
        if u.full is not None or u.gather_work is not None or u.lazy_pending:
            return
        if self.fused:
            fab = self.fabric
            if lazy_ok and u.first_weight is not None:
                # the first GEMM reading first_weight gathers the whole unit (smpk_gemm_comm mode 3)
                full = torch.empty(u.numel, dtype=u.dtype, device=self.device)
                fab.defer_to_gemm(u, full)
            else:
                cur = torch.cuda.current_stream(self.device)
                fab.side.wait_stream(cur)
                with torch.cuda.stream(fab.side):
                    full = torch.empty(u.numel, dtype=u.dtype, device=self.device)
                    fab.gather(u, full)
                    ev = torch.cuda.Event()
                    ev.record(fab.side)
                u.gather_work = ev
            u._pending_full = full
            return
        full = torch.empty(u.numel, dtype=u.dtype, device=self.device)
        if self.degree > 1:
            u.gather_work = dist.all_gather_into_tensor(full, u.shard, group=self.shard_group, async_op=True)
        else:
            full.copy_(u.shard)
        u._pending_full = full

    def _ensure(self, u: Unit, lazy_ok=False):
        if u.full is None:
            if u.gather_work is None and not u.lazy_pending and getattr(u, "_pending_full", None) is None:
                self.gather_stats["on_demand"] += 1
            self._issue_gather(u, lazy_ok)
            if u.gather_work is not None:
                if self.fused:
                    cur = torch.cuda.current_stream(self.device)
                    cur.wait_event(u.gather_work)
                    u._pending_full.record_stream(cur)
                else:
                    u.gather_work.wait()
                u.gather_work = None
            u.full = u._pending_full
            u._pending_full = None
            for p in u.params:
                o = u.offsets[p]
                p.data = u.full[o:o + u.numels[p]].view(u.shapes[p])
        elif u.lazy_pending and not lazy_ok:
            self.fabric.cancel_deferred(u)

    def _release(self, u: Unit):
        if u.full is None:
            return
        if u.lazy_pending:
            from . import sdp_fused
            sdp_fused.pending_gathers.pop(u.lazy_key, None)
            u.lazy_pending = False
            raise SMPValidationError(
                f"sharded data parallelism: unit {u.name} was released but the GEMM expected to gather it "
                "(sdp_first_weight) never ran; the module read other parameters first")
        for p in u.params:
            p.data = torch.empty(0, dtype=p.dtype, device=self.device)
        u.full = None

    def _prefetch_after(self, u: Unit, direction: int):
        if not self._order_frozen or u not in self._order_index:
            return
        j = self._order_index[u] + direction
        if 0 <= j < len(self._order):
            nxt = self._order[j]
            if self.fused and nxt.first_weight is not None and (direction > 0 or nxt.recompute_seen):
                return          # gathered from inside its own first GEMM
            # sdp_max_live_parameters: a unit larger than the budget is never fetched ahead of its use
            if nxt.full is None and nxt.gather_work is None and not nxt.lazy_pending and \
                    getattr(nxt, "_pending_full", None) is None and nxt.numel <= self.max_live:
                self.gather_stats["prefetched"] += 1
                self._issue_gather(nxt)

    def _pre_forward(self, units, args=None):
        for u in units:
            if not self._order_frozen and u not in self._order:
                self._order.append(u)
            if self._in_backward:
                u.recompute_seen = True
            else:
                u.calls_in_mb = getattr(u, "calls_in_mb", 0) + 1
            self._ensure(u, lazy_ok=True)
            u.users += 1
        if not self._in_backward:            # a recomputed forward must not prefetch units that are already done
            self._prefetch_after(units[-1], +1)
        if args is None:
            return None
        # the marker's backward runs after every other node of the unit: all its weight gradients are final,
        # so that is where the unit's reduce-scatter is issued (None when there is nothing to hang it on: such
        # units are flushed at the end of backward)
        return mark_inputs(args, lambda _us=tuple(units): [self._reduce_unit(u) for u in _us])

    def _post_forward(self, units, out=None):
        if out is not None and torch.is_grad_enabled() and not self._in_backward:
            # module backward-pre hooks only see plain tensor / tuple outputs: for anything else (a dict, a
            # HuggingFace ModelOutput) the unit is gathered when the first gradient of an output tensor arrives
            if not (torch.is_tensor(out) or (isinstance(out, tuple) and all(torch.is_tensor(o) for o in out))):
                def arm(t, _us=tuple(units)):
                    if t.requires_grad:
                        t.register_hook(lambda g, _us=_us: self._pre_backward(_us))
                    return t
                map_tensors(out, arm)
        for u in units:
            u.users -= 1
            if u.users <= 0 and u.full_grad is None:
                u.users = 0
                if not self._in_backward:        # a recomputed forward is followed by its backward at once
                    self._release(u)

    def _pre_backward(self, units):
        for u in units:
            self._ensure(u)
            if u.full_grad is None:
                if self.fused:
                    u.full_grad = self.fabric.acquire_slot(u)
                else:
                    u.full_grad = torch.zeros(u.numel, dtype=torch.float32, device=self.device)
                for p in u.params:
                    o = u.offsets[p]
                    p.main_grad = u.full_grad[o:o + u.numels[p]].view(u.shapes[p])
        self._prefetch_after(units[0], -1)

    def _post_backward(self, units, grad_input=()):
        # PyTorch fires this hook *before* the module's backward when none of its inputs requires grad;
        # such units are flushed at the end of ShardedDataParallelEngine.backward instead
        if not any(g is not None for g in grad_input):
            return
        for u in units:
            self._reduce_unit(u)

    def _reduce_unit(self, u: Unit):
        if u.full_grad is None:
            return
        for p in u.params:                       # fold gradients autograd produced the ordinary way
            if p.grad is not None:
                p.main_grad.add_(p.grad.reshape(p.main_grad.shape).float())
                p.grad = None
            if hasattr(p, "main_grad"):
                del p.main_grad
        if self.fused:
            self.fabric.reduce(u, self._fused_adam_args(u))
        else:
            while len(self._pending_reduce) >= 2:    # bound the number of full gradient buffers alive
                self._finish_one_reduce()
            if self.degree > 1:
                tmp = torch.empty(u.per, dtype=torch.float32, device=self.device)
                work = dist.reduce_scatter_tensor(tmp, u.full_grad, group=self.shard_group, async_op=True)
                self._pending_reduce.append((work, tmp, u, u.full_grad))
            else:
                u.grad_shard.add_(u.full_grad[:u.per])
        u.full_grad = None
        if u.users <= 0:
            self._release(u)

    def _adam_in_backward(self) -> bool:
        """Opt-in (config ``sdp_adam_in_backward: true`` or ``SMP_SDP_FUSED_ADAM=1``): the reduce kernel of the last
        microbatch also applies Adam and rewrites the 16-bit shard, i.e. the parameter update happens inside
        ``model.backward`` and ``optimizer.step()`` only does the bookkeeping.  Consequences the caller accepts with
        the opt-in: skipping ``optimizer.step()`` does not skip the update, ``grad_norm()`` sees no gradient for the
        units already updated, and the first ``clip_master_grads`` call switches the engine back to the ordinary
        order (reduce in backward, clip + Adam in ``optimizer.step()``) from the next step on.  Off by default."""
        env = os.environ.get("SMP_SDP_FUSED_ADAM")
        if env is not None:
            return env == "1"
        return bool(getattr(self.state.cfg, "sdp_adam_in_backward", False))

    def _fused_adam_args(self, u: Unit):
        """Hyper-parameters for applying Adam inside the reduce kernel, or None when the update has to wait:
        the opt-in is off, not the last microbatch, a global quantity is needed first (clipping, dynamic loss
        scale), gradients still have to cross replica groups, or the optimizer is not Adam/AdamW."""
        opt = self.optimizer
        if not self._adam_in_backward():
            return None
        if opt is None or (self.clip and self.clip > 0) or self.dynamic or self.loss_scale != 1.0:
            return None
        if self.replica_group is not None and dist.get_world_size(self.replica_group) > 1:
            return None
        if type(opt).__name__ not in ("Adam", "AdamW", "FusedAdam"):
            return None
        if self.state.microbatch != self.state.num_microbatches() - 1 or not self._sync_step:
            return None
        if getattr(u, "calls_in_mb", 1) > 1:
            return None                      # a reused module is reduced once per use: the update must wait for all
        group = opt.param_groups[u.gi]
        if group.get("amsgrad", False) or group.get("maximize", False):
            return None
        st = opt.state[u.master]
        if "exp_avg" not in st:
            st["step"] = torch.tensor(0.0)
            st["exp_avg"] = torch.zeros_like(u.master.data)
            st["exp_avg_sq"] = torch.zeros_like(u.master.data)
        st["step"] = st["step"] + 1
        u.updated = True
        name = type(opt).__name__
        return {"state": st, "scale": 1.0 / self.world, "lr": group["lr"], "beta1": group["betas"][0],
                "beta2": group["betas"][1], "eps": group["eps"], "wd": group["weight_decay"], "step": self._step + 1,
                "adamw": name == "AdamW" or getattr(opt, "adam_w_mode", 0) == 1,
                "bias_correction": group.get("bias_correction", True)}

    def _finish_one_reduce(self):
        work, tmp, u, _keep = self._pending_reduce.pop(0)
        work.wait()
        u.grad_shard.add_(tmp)

    # ------------------------------------------------------------------ DistributedModel entry points
    def forward(self, module, args, kwargs):
        self._build()
        if self.state.microbatch == 0 and any(u.updated for u in self.units):
            # in-backward Adam ran last step but optimizer.step() was never called: the update is already in the
            # shards, so close that step here (bias-correction count, flags, peers' view of the parameters)
            self._step += 1
            for u in self.units:
                u.updated = False
            if self.fabric is not None:
                self.fabric.publish_params()
        for u in self.units:
            u.calls_in_mb = 0
        out = module(*args, **kwargs)
        if not self._order_frozen and self._order:
            self._order_frozen = True
            self._order_index = {u: i for i, u in enumerate(self._order)}
        return out

    def backward(self, tensors, grads):
        self._in_backward = True
        try:
            torch.autograd.backward(tensors, grads)
        finally:
            self._in_backward = False
        # units whose module-level backward hook cannot fire (no input requires grad) are flushed here
        for u in self.units:
            if u.full_grad is not None:
                self._reduce_unit(u)
            elif u.full is not None and u.users <= 0:
                self._release(u)

    def reduce_epilogue(self, sync=True):
        """End of the minibatch: finish outstanding reduce-scatters; average over all data-parallel ranks;
        all-reduce across replica groups and the replicated (persistent) parameters' gradients."""
        while self._pending_reduce:
            self._finish_one_reduce()
        if self.fabric is not None:
            self.fabric.wait_reduces()
        for u in self.units:                 # a prefetch nobody consumed would be stale after the optimizer step
            if u.full is None and u.gather_work is not None:
                if not self.fused:
                    u.gather_work.wait()
                else:
                    torch.cuda.current_stream(self.device).wait_event(u.gather_work)
                u.gather_work, u._pending_full = None, None
        if not sync:
            return
        scale = 1.0 / self.world
        for u in self.units:
            if u.updated:
                continue
            if self.replica_group is not None and dist.get_world_size(self.replica_group) > 1:
                dist.all_reduce(u.grad_shard, group=self.replica_group)
            u.grad_shard.mul_(scale)
        if self.persistent_buffer is not None:
            buf = self.persistent_buffer.buffer
            if self.world > 1:
                dist.all_reduce(buf)
            buf.mul_(scale)
        self._check_overflow()

    def zero_grad(self):
        for u in self.units:
            u.grad_shard.zero_()
        if self.persistent_buffer is not None:
            self.persistent_buffer.zero()

    # ------------------------------------------------------------------ optimizer
    def attach_optimizer(self, optimizer, clip=None, static_loss_scale=1.0, dynamic_loss_scale=False,
                         dynamic_loss_args=None):
        group_index = {}
        for gi, g in enumerate(optimizer.param_groups):
            for p in g["params"]:
                group_index[id(p)] = gi
        self._build(lambda p: group_index.get(id(p), 0))
        self.optimizer = optimizer
        if clip is not None:
            self.clip = float(clip)
        cfg = self.state.cfg
        if cfg.fp16:
            self.dynamic = bool(dynamic_loss_scale)
            self.loss_scaler = DynamicLossScaler(**(dynamic_loss_args or {})) if self.dynamic else LossScaler(static_loss_scale)
        new_groups = [[] for _ in optimizer.param_groups]
        for u in self.units:
            new_groups[u.gi].append(u.master)
        for p in self.persistent:
            m = self.persistent_masters[p]
            m.grad = p.main_grad
            new_groups[self._group_of_current(p)].append(m)
        for g, plist in zip(optimizer.param_groups, new_groups):
            g["params"] = plist
        optimizer.state.clear()

    @property
    def loss_scale(self):
        return self.loss_scaler.loss_scale

    def _local_grads(self):
        gs = [u.grad_shard for u in self.units]
        return gs

    def _check_overflow(self):
        if not self.dynamic:
            self.overflow = False
            return
        grads = self._local_grads() + ([self.persistent_buffer.buffer] if self.persistent_buffer is not None else [])
        norm, _ = MT.multi_tensor_l2norm(MT.DEFAULT_CHUNK, None, [grads], False)
        flag = (~torch.isfinite(norm)).float()
        if self.world > 1:
            dist.all_reduce(flag, op=dist.ReduceOp.MAX)
        self.overflow = bool(flag.item() > 0)

    def grad_norm(self):
        """Global L2 norm of the (still loss-scaled) averaged gradient: shard norms summed over the shard
        group; replicated parameters counted once."""
        dev = self.device
        sq = torch.zeros(1, device=dev)
        if self.units:
            n, _ = MT.multi_tensor_l2norm(MT.DEFAULT_CHUNK, None, [self._local_grads()], False)
            sq = n.float() ** 2
        if self.degree > 1:
            dist.all_reduce(sq, group=self.shard_group)
        if self.persistent_buffer is not None:
            n, _ = MT.multi_tensor_l2norm(MT.DEFAULT_CHUNK, None, [[self.persistent_buffer.buffer]], False)
            sq = sq + n.float() ** 2
        return sq.sqrt()

    def clip_grad_norm(self, max_norm):
        self.clip = float(max_norm)
        return self.grad_norm() / self.loss_scale

    def optimizer_step(self):
        opt = self.optimizer
        if opt is None:
            raise SMPValidationError("smp.DistributedOptimizer must wrap the optimizer before stepping")
        scale = self.loss_scale
        self.loss_scaler.update_scale(self.overflow)
        if self.overflow:
            get_logger().info(f"OVERFLOW! Skipping step. Attempted loss scale: {scale}")
            return None
        coef = None
        if self.clip and self.clip > 0:
            total = self.grad_norm()
            coef = ((self.clip * scale) / (total + 1e-6)).clamp(max=1.0).float().reshape(1).contiguous()
        inv = 1.0 / scale
        name = type(opt).__name__
        self._step += 1
        if name in ("Adam", "AdamW", "FusedAdam"):
            adamw = name == "AdamW" or getattr(opt, "adam_w_mode", 0) == 1
            for gi, group in enumerate(opt.param_groups):
                b1, b2 = group["betas"]
                args = (group["lr"], b1, b2, group["eps"], self._step, 1 if adamw else 0,
                        group.get("bias_correction", True), group["weight_decay"], inv, coef)
                l5 = [[], [], [], [], []]
                for u in self.units:
                    if u.gi != gi or u.updated:
                        continue
                    st = opt.state[u.master]
                    if "exp_avg" not in st:
                        st["step"] = torch.tensor(0.0)
                        st["exp_avg"] = torch.zeros_like(u.master.data)
                        st["exp_avg_sq"] = torch.zeros_like(u.master.data)
                    st["step"] = st["step"] + 1
                    for lst, t in zip(l5, (u.grad_shard, u.master.data, st["exp_avg"], st["exp_avg_sq"], u.shard)):
                        lst.append(t)
                masters = [m for m in group["params"] if not any(m is u.master for u in self.units)]
                low = {id(self.persistent_masters[p]): p for p in self.persistent}
                for m in masters:
                    st = opt.state[m]
                    if "exp_avg" not in st:
                        st["step"] = torch.tensor(0.0)
                        st["exp_avg"] = torch.zeros_like(m.data)
                        st["exp_avg_sq"] = torch.zeros_like(m.data)
                    st["step"] = st["step"] + 1
                    for lst, t in zip(l5, (low[id(m)].main_grad, m.data, st["exp_avg"], st["exp_avg_sq"], low[id(m)].data)):
                        lst.append(t)
                if l5[0]:
                    if l5[4][0].dtype == torch.float32:      # fp32 training: no separate low-precision copy
                        MT.multi_tensor_adam(MT.DEFAULT_CHUNK, None, l5[:4], *args)
                        for mdat, lo in zip(l5[1], l5[4]):
                            lo.copy_(mdat)
                    else:
                        MT.multi_tensor_adam(MT.DEFAULT_CHUNK, None, l5, *args)
        else:
            factor = inv * (float(coef.item()) if coef is not None else 1.0)
            if factor != 1.0:
                for u in self.units:
                    u.grad_shard.mul_(factor)
                if self.persistent_buffer is not None:
                    self.persistent_buffer.buffer.mul_(factor)
            opt.step()
            for u in self.units:
                u.shard.copy_(u.master.data)
            for p in self.persistent:
                p.data.copy_(self.persistent_masters[p].data)
        for u in self.units:
            u.updated = False
        if self.fabric is not None:
            self.fabric.publish_params()
        return None

    def _before_shard_write(self):
        """Host-driven rewrite of the shards (checkpoint load): no peer may still be reading them."""
        if self.fabric is not None:
            torch.cuda.synchronize()
            dist.barrier(group=self.shard_group)

    def _after_shard_write(self):
        if self.fabric is not None:
            self.fabric.publish_params()
            torch.cuda.synchronize()
            dist.barrier(group=self.shard_group)

    # ------------------------------------------------------------------ state dicts / checkpoints
    def full_state_dict(self):
        """Gather every unit and return an ordinary (un-sharded) state dict on every rank."""
        self._build()
        for u in self.units:
            self._ensure(u)
        sd = OrderedDict((k, v.detach().cpu().clone()) for k, v in self.model.module.state_dict().items())
        for u in self.units:
            self._release(u)
        return sd

    def load_full_state_dict(self, sd, strict=True):
        self._build()
        for u in self.units:
            self._ensure(u)
        res = self.model.module.load_state_dict(sd, strict=strict)
        self._before_shard_write()
        for u in self.units:                         # re-shard what was loaded
            u.shard.copy_(u.shard_of(u.full))
            u.master.data.copy_(u.shard)
            self._release(u)
        self._after_shard_write()
        for p in self.persistent:
            self.persistent_masters[p].data.copy_(p.data)
        return res

    def _model_shard_dict(self):
        self._build()
        return {"module": {"units": {u.name: u.shard.detach().cpu() for u in self.units},
                           "persistent": {self.model.get_param_name(p): p.detach().cpu() for p in self.persistent},
                           "buffers": {n: b.detach().cpu() for n, b in self.model.module.named_buffers()}},
                "shard_size": self.degree, "_smp_zero2d": True, "striped": self.fused}

    def optimizer_state_dict(self):
        opt = self.optimizer
        return {"optimizer_state_dict": opt.state_dict() if opt is not None else None,
                "fp32_flat_groups": [u.master.detach().cpu() for u in self.units],
                "persistent_masters": {self.model.get_param_name(p): m.detach().cpu()
                                       for p, m in self.persistent_masters.items()},
                "partition_count": self.degree, "zero_stage": 3, "dynamic_loss_scale": self.dynamic,
                "loss_scaler": self.loss_scaler.state_dict(), "overflow": self.overflow, "_smp_zero2d": True,
                "step": self._step, "striped": self.fused}

    def load_optimizer_state_dict(self, sd, load_sharded_optimizer_state=True):
        self.loss_scaler.load_state_dict(sd["loss_scaler"])
        self.overflow = sd.get("overflow", False)
        self._step = sd.get("step", 0)
        if not load_sharded_optimizer_state:
            return
        if sd.get("partition_count") != self.degree:
            raise SMPValidationError("sharded_data_parallel_degree differs from the checkpoint's")
        if sd.get("striped", False) != self.fused:
            raise SMPValidationError("checkpoint shard layout (striped vs contiguous) differs from this run's; "
                                     "convert through a full checkpoint (partial=False)")
        self._before_shard_write()
        for u, m in zip(self.units, sd["fp32_flat_groups"]):
            u.master.data.copy_(m.to(self.device))
            u.shard.copy_(u.master.data)
        self._after_shard_write()
        names = {self.model.get_param_name(p): p for p in self.persistent}
        for n, m in sd.get("persistent_masters", {}).items():
            if n in names:
                self.persistent_masters[names[n]].data.copy_(m.to(self.device))
                names[n].data.copy_(self.persistent_masters[names[n]].data)
        if sd.get("optimizer_state_dict") is not None and self.optimizer is not None:
            self.optimizer.load_state_dict(sd["optimizer_state_dict"])

    def _writes_checkpoint(self):
        # only the first replica of every shard index writes (reference checkpoint.py:300-319)
        return self.state.core.rank() == self.shard_rank

    def save_model_checkpoint(self, path, tag):
        if self._writes_checkpoint():
            torch.save(self._model_shard_dict(), os.path.join(path, tag, f"model_{self.shard_rank}.pt"))
        self.state.comm.barrier()

    def save_optimizer_checkpoint(self, path, tag):
        if self._writes_checkpoint():
            torch.save(self.optimizer_state_dict(), os.path.join(path, tag, f"optimizer_{self.shard_rank}.pt"))
        self.state.comm.barrier()

    def load_model_checkpoint(self, path, tag, strict=True):
        self._build()
        sd = torch.load(os.path.join(path, tag, f"model_{self.shard_rank}.pt"), map_location="cpu", weights_only=False)
        if not sd.get("_smp_zero2d") or sd.get("shard_size") != self.degree:
            raise SMPValidationError("checkpoint was not written with the same sharded_data_parallel_degree")
        if sd.get("striped", False) != self.fused:
            raise SMPValidationError("checkpoint shard layout (striped vs contiguous) differs from this run's; "
                                     "convert through a full checkpoint (partial=False)")
        mod = sd["module"]
        self._before_shard_write()
        for u in self.units:
            if u.name in mod["units"]:
                u.shard.copy_(mod["units"][u.name].to(self.device))
                u.master.data.copy_(u.shard)
            elif strict:
                raise SMPValidationError(f"unit {u.name} missing from checkpoint")
        names = {self.model.get_param_name(p): p for p in self.persistent}
        for n, t in mod["persistent"].items():
            if n in names:
                names[n].data.copy_(t.to(self.device))
                self.persistent_masters[names[n]].data.copy_(names[n].data)
        bufs = dict(self.model.module.named_buffers())
        for n, t in mod.get("buffers", {}).items():
            if n in bufs:
                bufs[n].copy_(t.to(bufs[n].device))
        self._after_shard_write()

    def load_optimizer_checkpoint(self, path, tag, load_sharded_optimizer_state=True):
        sd = torch.load(os.path.join(path, tag, f"optimizer_{self.shard_rank}.pt"), map_location="cpu",
                        weights_only=False)
        self.load_optimizer_state_dict(sd, load_sharded_optimizer_state)
