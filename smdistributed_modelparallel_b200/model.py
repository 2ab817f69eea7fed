"""``smp.DistributedModel`` and ``smp.model_creation``.

Public surface follows the reference (reference smdistributed/modelparallel/torch/model.py:78-107
model_creation, :111-251 constructor, :703-753 post_partition, :863-1096 state-dict API, :1113-1146
backward, :1154-1235 _step, :1359-1441 local_* accessors, :1548-1608 hooks).  Construction order is
the same - cast to 16 bit, replace modules marked for tensor parallelism, name the tree, partition
(deferred to the first step when pp > 1), move local modules to the GPU, build the gradient buffer and
reducers, broadcast parameters over the data-parallel group - with the B200-first pieces underneath:
one contiguous fp32 gradient buffer written by the wgrad GEMM epilogues, fused multi-tensor updates,
and a stage executor instead of a per-request module server.
"""

from __future__ import annotations

from collections import OrderedDict
from contextlib import contextmanager

import torch
import torch.distributed as dist
import torch.nn as nn
from torch.utils.hooks import RemovableHandle

from .backend.exceptions import (BwdExecutionNotInStepFnError, FwdExecutionNotInStepFnError,
                                 ModelNotPartitionedError, MultipleDistributedModelError, NotInitializedError,
                                 SMPUnsupportedError, SMPValidationError)
from .backend.logger import get_logger
from .nn.dist_module import DistributedModule
from .nn.utils import is_distributed_parameter, is_one_rank_parameter, is_scaled_batch_parameter
from .parallel.grad_buffer import GradBuffer, GradReducer
from .state_mod import state

_HALF = (torch.float16, torch.bfloat16)


@contextmanager
def model_creation(tensor_parallelism=False, dtype=None, **tensor_parallel_config):
    """Context in which the model is constructed: sets the default dtype, marks every module created
    inside for tensor parallelism (when requested) and applies delayed parameter initialisation when the
    config asks for it (reference model.py:78-107)."""
    from .parameter import delay_param_initialization

    if not state.initialized:
        raise NotInitializedError()
    if dtype is None:
        dtype = torch.float16 if state.cfg.fp16 else (torch.bfloat16 if state.cfg.bf16 else None)
    prev = torch.get_default_dtype()
    if dtype is not None:
        torch.set_default_dtype(dtype)
    try:
        with state.module_manager.tensor_parallelism(enabled=tensor_parallelism, **tensor_parallel_config):
            with delay_param_initialization(enabled=state.cfg.delayed_parameter_initialization):
                yield
    finally:
        torch.set_default_dtype(prev)


class DistributedModel(nn.Module):
    def __init__(self, module, trace_device="gpu", trace_execution_times=False, trace_memory_usage=False,
                 overlapping_allreduce=True, backward_passes_per_step=1, average_grads_across_microbatches=True,
                 bucket_cap_mb=100, find_unused_parameters=False, broadcast_buffers=True,
                 gradient_as_bucket_view=True):
        super().__init__()
        if not state.initialized:
            raise NotInitializedError()
        if state.model is not None:
            raise MultipleDistributedModelError()
        if trace_device not in ("cpu", "gpu"):
            raise SMPValidationError("trace_device must be 'cpu' or 'gpu'")
        cfg = state.cfg
        self.trace_device = trace_device
        self.trace_execution_times = trace_execution_times
        self.trace_memory_usage = trace_memory_usage
        self.overlapping_allreduce = overlapping_allreduce
        self.backward_passes_per_step = backward_passes_per_step
        self.average_grads_across_microbatches = average_grads_across_microbatches
        self.bucket_cap_mb = bucket_cap_mb
        self.find_unused_parameters = find_unused_parameters
        self.broadcast_buffers = broadcast_buffers
        self.gradient_as_bucket_view = gradient_as_bucket_view
        self._post_partition_hooks = OrderedDict()
        self._post_step_hooks = OrderedDict()
        self._backward_passes = 0
        self._in_backward = False
        self.grad_buffer = None
        self.reducer = None
        self.sdp_engine = None
        self._partitioned = False
        self._num_steps = 0
        self._join = None

        # 1. precision
        self.dtype = torch.float16 if cfg.fp16 else (torch.bfloat16 if cfg.bf16 else None)
        if self.dtype is not None:
            module = _cast_floating(module, self.dtype)

        # 2. tensor-parallel replacements
        state.tp_registry.register_builtin()
        mm = state.module_manager
        if cfg.tensor_parallel_degree > 1 or any(mm.is_marked_for_tp(m) for m in module.modules()):
            mm.simplify_tensor_parallelism_modules(module, lambda m: state.tp_registry.is_supported(type(m)))
            module = self._replace_tp_counterparts(module)
        self.module = module
        state.model = self

        # 3. name the tree
        mm.name_modules_and_create_parent_map(self.module)

        # 4. partition now when there is nothing to decide (pp == 1); otherwise at the first step
        if cfg.pipeline_parallel_degree == 1:
            mm.assign_partitions({n: 0 for n in mm.module_names()})
            self.post_partition()
        else:
            from .runtime.server import ExecutionServer
            state.exec_server = ExecutionServer(self)
        if state.loaded_model_state is not None:
            sd, kwargs = state.loaded_model_state
            state.loaded_model_state = None
            self.load_state_dict(sd, **kwargs)

    # ------------------------------------------------------------------ construction helpers
    def _replace_tp_counterparts(self, module):
        reg, mm = state.tp_registry, state.module_manager
        memo = {}

        def convert(m):
            if id(m) in memo:
                return memo[id(m)]
            out = m
            if not isinstance(m, DistributedModule) and mm.is_marked_for_tp(m) and reg.is_supported(type(m)):
                out = reg.distribute(m)
                if out is not m:
                    part = mm.manual_partition_of(m)
                    if part is not None:
                        mm.set_partition(out, part)
                    ck = mm.get_checkpoint_config(m)
                    if ck is not None:
                        mm._ckpt_configs[out] = ck
                    if self.dtype is not None:
                        out = _cast_floating(out, self.dtype)
            memo[id(m)] = out
            if out is m and not isinstance(m, DistributedModule):     # never look inside distributed modules
                for name, child in list(m.named_children()):
                    new = convert(child)
                    if new is not child:
                        setattr(m, name, new)
            return out

        return convert(module)

    def post_partition(self):
        """Everything that needs the partition: materialise + move local modules, gradient buffer,
        reducers, parameter broadcast, hooks (reference model.py:703-753)."""
        cfg, core = state.cfg, state.core
        from .parameter import materialize_delayed

        if not cfg.zero2d_enabled():
            materialize_delayed(self.module, local_only=lambda m: self._is_local_module(m))
        # (sharded data parallelism: the engine materialises, shards and releases one unit at a time -
        # parallel/sdp.py::_build - so no rank ever holds the whole model)
        dev = state.device
        if dev.type == "cuda":
            for m in self.local_modules():
                _move_own_tensors(m, dev)
        self._release_nonlocal()
        self._partitioned = True
        params = list(self.local_parameters())
        if cfg.zero2d_enabled():
            from .parallel.sdp import ShardedDataParallelEngine
            self.sdp_engine = ShardedDataParallelEngine(self, params)
        else:
            if self.dtype is not None or cfg._fp32_grad_accumulation:
                half = [p for p in params if p.dtype in _HALF]
                if half:
                    pad = core.dp_size() * 8 if cfg.shard_optimizer_state else 1
                    self.grad_buffer = GradBuffer(half, dev, pad_to=pad)
            if core.dp_size() > 1:
                self._broadcast_parameters_and_buffers()
                self.reducer = GradReducer(params, state.rdp_process_group, state.dp_process_group, core.dp_size(),
                                           self.bucket_cap_mb, self.grad_buffer)
                if self.overlapping_allreduce and not cfg.shard_optimizer_state:
                    self.reducer.install_overlap(
                        self.module, lambda: state.microbatch == state.num_microbatches() - 1)
                    self.reducer.install_hook_overlap(state.num_microbatches(), lambda: state.microbatch)
        from .runtime.checkpointing import apply_configured_checkpointing
        apply_configured_checkpointing(self.module)
        for hook in list(self._post_partition_hooks.values()):
            hook(self, state.optimizer)
        if core.rank() == 0:
            get_logger().debug(f"model ready: {sum(p.numel() for p in params) / 1e6:.1f} M local parameters")

    def _is_local_module(self, m) -> bool:
        mm = state.module_manager
        return (not mm.partitioned) or mm.get_partition(m) == state.core.pp_rank()

    def _release_nonlocal(self):
        """Parameters of modules that live on other pipeline ranks are dropped to empty tensors."""
        if state.core.pp_size() == 1:
            return
        for m in self.module.modules():
            if self._is_local_module(m):
                continue
            for name, p in list(m.named_parameters(recurse=False)):
                p.data = torch.empty(0, dtype=p.dtype, device=p.device)
                p.requires_grad_(False)
                p._smp_nonlocal = True
            for name, b in list(m.named_buffers(recurse=False)):
                b.data = torch.empty(0, dtype=b.dtype, device=b.device)
                b._smp_nonlocal = True

    def _broadcast_parameters_and_buffers(self):
        """Make data-parallel replicas start identical: distributed / scaled-batch tensors over the RDP group,
        replicated ones over the whole DP group (source: group rank 0), in 250 MB coalesced buckets."""
        core = state.core
        tensors_rdp, tensors_dp = [], []
        for t in list(self.local_parameters()) + (list(self.local_buffers()) if self.broadcast_buffers else []):
            if t.numel() == 0:
                continue
            (tensors_rdp if is_distributed_parameter(t) or is_one_rank_parameter(t) else tensors_dp).append(t.data)
        for tensors, group, ranks in ((tensors_rdp, state.rdp_process_group, core.get_rdp_group()),
                                      (tensors_dp, state.dp_process_group, core.get_dp_group())):
            if dist.get_world_size(group) == 1:
                continue
            _broadcast_coalesced(tensors, group, ranks[0], 250 * 1024 * 1024)

    # ------------------------------------------------------------------ execution
    def forward(self, *args, **kwargs):
        if not state.in_step_func:
            raise FwdExecutionNotInStepFnError()
        if self.dtype is not None:
            args, kwargs = _cast_inputs(args, self.dtype), _cast_inputs(kwargs, self.dtype)
        if state.core.pp_size() > 1:
            return state.exec_server.run_model_forward(self, args, kwargs)
        if self.sdp_engine is not None:
            return self.sdp_engine.forward(self.module, args, kwargs)
        return self.module(*args, **kwargs)

    def backward(self, tensors, grad_tensors=None, retain_graph=None):
        """Backward of one microbatch; must be called inside an ``@smp.step`` function.

        Gradients are seeded with ``loss_scale / num_microbatches`` (averaging over microbatches, as the
        reference's reducers do: ddp_model.py:336-374) so no second scaling pass is needed."""
        if not state.in_step_func:
            raise BwdExecutionNotInStepFnError()
        tensors = (tensors,) if torch.is_tensor(tensors) else tuple(tensors)
        factor = 1.0
        if self.average_grads_across_microbatches:
            factor /= state.num_microbatches()
        opt = state.optimizer
        if opt is not None and getattr(opt, "loss_scale", None) is not None and state.cfg.fp16:
            factor *= float(opt.loss_scale)
        if grad_tensors is None:
            grad_tensors = tuple(torch.full_like(t, factor) for t in tensors)
        else:
            grad_tensors = (grad_tensors,) if torch.is_tensor(grad_tensors) else tuple(grad_tensors)
            grad_tensors = tuple(g * factor for g in grad_tensors)
        if state.core.pp_size() > 1:
            return state.exec_server.run_model_backward(self, tensors, grad_tensors)
        if self.sdp_engine is not None:
            return self.sdp_engine.backward(tensors, grad_tensors)
        torch.autograd.backward(tensors, grad_tensors, retain_graph=retain_graph)

    @contextmanager
    def _step(self):
        """Wraps one ``@smp.step`` call: decides whether gradients are synchronised this call, then after
        the microbatches reduces gradients, refreshes master gradients and runs post-step hooks
        (reference model.py:1154-1235)."""
        self._backward_passes += 1
        sync = (self._backward_passes % self.backward_passes_per_step) == 0
        if getattr(self, "_join", None) is not None and sync and torch.is_grad_enabled():
            active = self._join_count(1)                 # joined ranks answer with 0 from _join_drain
            self._join["steps"] += 1
            if self.reducer is not None:
                self.reducer.scale_override = None if self._join["by_world"] else 1.0 / max(active, 1)
        if self.reducer is not None:
            self.reducer.require_sync = sync
            if sync:
                self.reducer.prepare_for_backward()
        if self.sdp_engine is not None:
            self.sdp_engine._sync_step = sync
        ran_backward = {"v": False}
        self._ran_backward = ran_backward
        # An exception inside the step propagates untouched: the reductions and the barrier below are
        # collectives, and entering them on the ranks that failed would turn the error into a hang.
        yield
        if torch.is_grad_enabled() and self._partitioned:
            opt = state.optimizer
            shard = getattr(opt, "_opt_shard", None) if opt is not None else None
            if self.sdp_engine is not None:
                self.sdp_engine.reduce_epilogue(sync)
            elif shard is not None:
                if sync:
                    shard.reduce_gradients()
                    shard.check_overflow()
            else:
                if self.reducer is not None and sync:
                    self.reducer.synchronize()
                if opt is not None and sync and hasattr(opt, "update_master_grads"):
                    opt.update_master_grads()
            for hook in list(self._post_step_hooks.values()):
                hook(self, opt)
        self._num_steps += 1
        if state.core.pp_size() > 1:
            state.comm.pp_barrier()

    def zero_grad_buffer(self):
        if self.grad_buffer is not None:
            self.grad_buffer.zero()
        if self.sdp_engine is not None:
            self.sdp_engine.zero_grad()

    def zero_grad(self, set_to_none: bool = True):
        self.zero_grad_buffer()
        for p in self.local_parameters():
            p.grad = None

    # ------------------------------------------------------------------ accessors
    @property
    def partitioned(self):
        return self._partitioned

    def _validate_partitioned(self):
        if not self._partitioned:
            raise ModelNotPartitionedError()

    def get_module(self):
        return self.module

    def local_modules(self):
        for m in self.module.modules():
            if self._is_local_module(m):
                yield m

    def local_named_modules(self, memo=None, prefix=""):
        for n, m in self.module.named_modules(memo, prefix):
            if self._is_local_module(m):
                yield n, m

    def local_named_parameters(self, recurse=True):
        seen = set()
        for mn, m in self.module.named_modules():
            if not self._is_local_module(m):
                continue
            for pn, p in m.named_parameters(recurse=False):
                if id(p) in seen or getattr(p, "_smp_nonlocal", False):
                    continue
                seen.add(id(p))
                yield (f"{mn}.{pn}" if mn else pn), p

    def local_parameters(self, recurse=True):
        for _, p in self.local_named_parameters(recurse):
            yield p

    def local_named_buffers(self, prefix="", recurse=True):
        seen = set()
        for mn, m in self.module.named_modules():
            if not self._is_local_module(m):
                continue
            for bn, b in m.named_buffers(recurse=False):
                if id(b) in seen or getattr(b, "_smp_nonlocal", False):
                    continue
                seen.add(id(b))
                yield (f"{mn}.{bn}" if mn else bn), b

    def local_buffers(self, recurse=True):
        for _, b in self.local_named_buffers(recurse=recurse):
            yield b

    # names as the wrapped module sees them, without a "module." prefix (reference model.py:1098-1108)
    def named_parameters(self, *args, **kwargs):
        return self.module.named_parameters(*args, **kwargs)

    def parameters(self, *args, **kwargs):
        return self.module.parameters(*args, **kwargs)

    def named_buffers(self, *args, **kwargs):
        return self.module.named_buffers(*args, **kwargs)

    def buffers(self, *args, **kwargs):
        return self.module.buffers(*args, **kwargs)

    def is_local_buffer(self, buffer):
        """By name or by tensor (reference model.py:1418-1428)."""
        self._validate_partitioned()
        if isinstance(buffer, str):
            return any(n == buffer for n, _ in self.local_named_buffers())
        if torch.is_tensor(buffer):
            return any(b is buffer for _, b in self.local_named_buffers())
        raise SMPUnsupportedError(f"is_local_buffer takes a name or a tensor, not {type(buffer).__name__}")

    def get_module_for_param(self, param):
        """The module that directly owns ``param`` (reference model.py:348-349)."""
        for m in self.module.modules():
            for p in m.parameters(recurse=False):
                if p is param:
                    return m
        raise KeyError("parameter does not belong to this model")

    def scaled_batch_reducer_named_parameters(self):
        """Parameters whose gradients are averaged over the reduced-data-parallel group only (they already saw
        the whole tensor-parallel group's batch; reference model.py:586-613, 1430-1432)."""
        for n, p in self.local_named_parameters():
            if p.requires_grad and is_scaled_batch_parameter(p):
                yield n, p

    def default_reducer_named_parameters(self):
        """Parameters averaged over the full data-parallel group (reference model.py:1434-1436)."""
        for n, p in self.local_named_parameters():
            if p.requires_grad and not is_scaled_batch_parameter(p):
                yield n, p

    def virtual_named_parameters(self):
        """Under optimizer-state sharding: this rank's slices of the flat parameter buffer, the tensors the
        inner optimizer really updates (reference model.py:1237-1340, 1371-1373)."""
        shard = getattr(state.optimizer, "_opt_shard", None) if state.optimizer is not None else None
        if shard is None:
            return
        yield from shard.virtual_named_parameters()

    def display_partition(self):
        """Log the module tree down to the first level at which a subtree lives on one pipeline rank
        (reference model.py:668-701).  Returns the lines."""
        self._validate_partitioned()
        mm = state.module_manager
        lines = []

        def ranks_below(m):
            return {mm.get_partition(x) for x in m.modules()}

        def walk(m, name, depth):
            below = ranks_below(m)
            if len(below) <= 1:
                lines.append(f"{'  ' * depth}{name or 'main'}: {next(iter(below)) if below else mm.get_partition(m)}")
                return
            lines.append(f"{'  ' * depth}{name or 'main'}: {mm.get_partition(m)} (children on {sorted(below)})")
            for cn, c in m.named_children():
                walk(c, f"{name}/{cn}" if name else cn, depth + 1)

        walk(self.module, "", 0)
        if state.core.rank() == 0:
            log = get_logger()
            log.info("Partition assignments:")
            for ln in lines:
                log.info(ln)
        return lines

    def load_partition(self, partitioning_and_trace_results=None):
        """Adopt a saved module -> pipeline-rank assignment (reference model.py:846-849)."""
        if partitioning_and_trace_results is not None:
            info = partitioning_and_trace_results
            if isinstance(info, dict) and "partition_info" in info:
                info = info["partition_info"]
            self.load_saved_partition(info)

    def get_ddp_logging_data(self):
        """What the reducer did in the steps so far (bucket count, overlapped / trailing elements)."""
        r = self.reducer
        if r is None:
            return {}
        out = dict(getattr(r, "stats", {}))
        out["num_buckets"] = len(getattr(r, "_hook_buckets", None) or getattr(r, "buckets", None) or [])
        out["dp_size"] = getattr(r, "dp_size", 1)
        return out

    def is_distributed_parameter(self, param):
        return is_distributed_parameter(param)

    def is_scaled_batch_parameter(self, param):
        return is_scaled_batch_parameter(param)

    def is_distributed_buffer(self, buf):
        return is_distributed_parameter(buf)

    def is_scaled_batch_buffer(self, buf):
        return is_scaled_batch_parameter(buf)

    def is_local_parameter(self, param):
        return not getattr(param, "_smp_nonlocal", False)

    def get_param_name(self, param):
        if getattr(self, "_param_names", None) is None or param not in self._param_names:
            self._param_names = {p: n for n, p in self.module.named_parameters()}
        return self._param_names.get(param)

    def distributed_modules(self):
        for m in self.module.modules():
            if isinstance(m, DistributedModule):
                yield m

    def size(self):
        return sum(p.numel() * p.element_size() for p in self.module.parameters())

    def cpu(self):
        raise SMPValidationError("DistributedModel placement is managed by smp; .cpu() is not supported")

    def cuda(self, *ignore):
        return self

    # ------------------------------------------------------------------ hooks
    def register_post_partition_hook(self, hook, prepend=False) -> RemovableHandle:
        """hook(model, optimizer) runs right after partitioning (immediately if already partitioned)."""
        handle = RemovableHandle(self._post_partition_hooks)
        self._post_partition_hooks[handle.id] = hook
        if prepend:
            self._post_partition_hooks.move_to_end(handle.id, last=False)
        if self._partitioned:
            hook(self, state.optimizer)
        return handle

    def register_post_step_hook(self, hook) -> RemovableHandle:
        handle = RemovableHandle(self._post_step_hooks)
        self._post_step_hooks[handle.id] = hook
        return handle

    def _call_after_partitioning(self, fn):
        if self._partitioned:
            fn(self, state.optimizer)
        else:
            self.register_post_partition_hook(fn)

    def register_comm_hook(self, hook_state, hook):
        if self.reducer is None:
            self._call_after_partitioning(lambda m, o: m.reducer and m.reducer.register_comm_hook(hook_state, hook))
        else:
            self.reducer.register_comm_hook(hook_state, hook)

    @contextmanager
    def join(self, divide_by_initial_world_size=True, enable=True, throw_on_early_termination=False):
        """Training with a different number of steps per data-parallel rank (reference torch/ddp_model.py:652-775, the
        adaptation of ``DistributedDataParallel.join`` to its reducers; on PyTorch >= 1.10 the reference refuses).

        Inside the context every synchronising step first counts the ranks still training.  A rank that leaves the
        context keeps answering that count and shadows the others' gradient reductions with zeros until nobody is
        left, then every rank takes the parameters and buffers of the rank that trained longest.  Gradients are
        averaged over all ranks (``divide_by_initial_world_size``) or over the ranks still training.  While the
        context is active the reduction is not overlapped with backward.  Plain data parallelism only."""
        if not enable or state.core.dp_size() == 1:
            yield
            return
        cfg = state.cfg
        if (state.core.pp_size() > 1 or state.core.tp_size() > 1 or self.sdp_engine is not None
                or cfg.shard_optimizer_state or cfg._fp32_grad_accumulation):
            raise SMPUnsupportedError("join supports plain data parallelism only (no pipeline / tensor / sharded "
                                      "data parallelism, optimizer-state sharding or fp32 gradient accumulation)")
        if self.reducer is not None and self.reducer.comm_hook is not None:
            raise SMPUnsupportedError("join cannot be combined with a communication hook")
        group = state.dp_process_group
        self._join = {"group": group, "by_world": bool(divide_by_initial_world_size), "steps": 0,
                      "throw": bool(throw_on_early_termination)}
        if self.reducer is not None:
            self.reducer.joining = True
        failed = False
        try:
            yield
        except BaseException:
            failed = True                # the other ranks cannot be answered consistently any more
            raise
        finally:
            ctx, self._join = self._join, None
            if self.reducer is not None:
                self.reducer.joining = False
                self.reducer.scale_override = None
            if not failed:
                self._join_drain(ctx)

    def _join_count(self, active: int) -> int:
        t = torch.tensor([float(active)], device=state.device)
        dist.all_reduce(t, group=self._join["group"] if self._join else None)
        return int(t.item())

    def _join_drain(self, ctx):
        """After leaving ``join``: shadow the remaining ranks' steps, then adopt the final model."""
        group = ctx["group"]
        self._join = ctx
        try:
            while True:
                if self._join_count(0) == 0:
                    break
                if ctx["throw"]:
                    raise RuntimeError("join: another rank is still training after this rank ran out of inputs")
                if self._partitioned and self.reducer is not None:
                    self.reducer.shadow_synchronize()
        finally:
            self._join = None
        # the rank that trained longest is authoritative (ties: lowest rank)
        me = dist.get_rank(group)
        key = torch.tensor([float(ctx["steps"]) * 4096.0 + (4095.0 - me)], device=state.device)
        dist.all_reduce(key, op=dist.ReduceOp.MAX, group=group)
        src_in_group = 4095 - int(key.item()) % 4096
        src = dist.get_global_rank(group, src_in_group)
        with torch.no_grad():
            for t in list(self.module.parameters()) + list(self.module.buffers()):
                if not t.is_meta:
                    dist.broadcast(t.data, src=src, group=group)
            # 16-bit training: the fp32 master weights must follow, or the next step would restore the old values
            b16 = getattr(state.optimizer, "fp16_optimizer", None) if state.optimizer is not None else None
            for grp in (getattr(b16, "fp32_from_fp16_groups", None) or []):
                for m in grp:
                    dist.broadcast(m.data, src=src, group=group)

    # ------------------------------------------------------------------ state dict API
    def local_state_dict(self, destination=None, prefix="", keep_vars=False):
        """Parameters / buffers this rank holds, plus the metadata needed to reload under the same or a
        different tensor-parallel degree (reference model.py:1482-1528)."""
        self._validate_partitioned()
        if self.sdp_engine is not None:
            return self.sdp_engine._model_shard_dict()
        full = self.module.state_dict(prefix=prefix + "module.", keep_vars=keep_vars)
        if state.core.pp_size() > 1:
            # tensors of modules living on other pipeline ranks were released to empty storage at partition time
            full = OrderedDict((k, v) for k, v in full.items() if v.numel() > 0)
        sd = destination if destination is not None else OrderedDict()
        for k, v in full.items():
            sd[k] = v
        sd["_smp_is_partial"] = True
        sd["_smp_load_info"] = {
            "tensor_parallel_degree": state.core.tp_size(),
            "pipeline_parallel_degree": state.core.pp_size(),
            "partition_info": state.module_manager.partition_info(),
            "tp_rank": state.core.tp_rank(),
            "pp_rank": state.core.pp_rank(),
        }
        return sd

    def state_dict(self, destination=None, prefix="", keep_vars=False, cast_to_cpu=True, gather_to_rank0=False):
        """Full (un-sharded, un-partitioned) state dict: tensor-parallel shards are concatenated on their
        distribution axis, pipeline partitions are unioned (reference model.py:863-931, utils.py:481-658)."""
        from .checkpoint_utils import gather_full_state_dict

        self._validate_partitioned()
        if self.sdp_engine is not None:
            return self.sdp_engine.full_state_dict()
        return gather_full_state_dict(self, prefix=prefix, cast_to_cpu=cast_to_cpu, gather_to_rank0=gather_to_rank0)

    def load_state_dict(self, state_dict, strict=True, same_partition_load=False, translate_function=None):
        from .checkpoint_utils import load_into_model

        def _do(model, _opt=None):
            if model.sdp_engine is not None and not (isinstance(state_dict, dict) and state_dict.get("_smp_zero2d")):
                sd = {k: v for k, v in state_dict.items() if not k.startswith("_smp_")}
                if translate_function is not None:
                    sd = translate_function(sd)
                return model.sdp_engine.load_full_state_dict(sd, strict=strict)
            return load_into_model(model, state_dict, strict=strict, translate_function=translate_function)

        if not self._partitioned:
            info = state_dict.get("_smp_load_info") if isinstance(state_dict, dict) else None
            info = info if isinstance(info, dict) else None
            if same_partition_load and info and info.get("partition_info"):
                self.load_saved_partition(info["partition_info"])
            # a deferred model load runs before the optimizer wrapper's own post-partition work: the wrapper first
            # copies model -> fp32 masters, then a pending optimizer checkpoint restores the exact masters (the same
            # order as loading into an already partitioned model)
            self.register_post_partition_hook(_do, prepend=True)
            return None
        return _do(self)

    def load_saved_partition(self, partition_info):
        """Use the module -> pipeline-rank assignment of an earlier run (``state.module_manager.
        get_model_partition_info()``) instead of partitioning again (reference model.py:936-942)."""
        state.module_manager.load_partition(partition_info)


# ---------------------------------------------------------------------------------------- helpers
def _cast_floating(module: nn.Module, dtype):
    """Cast floating-point parameters and buffers (keeps integer buffers and fp32-pinned tensors)."""
    def fn(t):
        if t.is_floating_point() and not getattr(t, "_smp_keep_fp32", False):
            return t.to(dtype)
        return t
    return module._apply(fn)


def _cast_inputs(obj, dtype):
    if torch.is_tensor(obj):
        return obj.to(dtype) if obj.is_floating_point() and obj.dtype == torch.float32 else obj
    if isinstance(obj, (list, tuple)):
        out = [_cast_inputs(o, dtype) for o in obj]
        return type(obj)(*out) if hasattr(obj, "_fields") else type(obj)(out)
    if isinstance(obj, dict):
        return {k: _cast_inputs(v, dtype) for k, v in obj.items()}
    return obj


def _move_own_tensors(module: nn.Module, device):
    """Move the tensors a module owns directly (not its children's), keeping Parameter identity and tags."""
    for name, p in module.named_parameters(recurse=False):
        if p.device != device and p.device.type != "meta":
            p.data = p.data.to(device)
            if p.grad is not None:
                p.grad.data = p.grad.data.to(device)
    for name, b in module.named_buffers(recurse=False):
        if b.device != device and b.device.type != "meta":
            module._buffers[name] = b.to(device)


def _broadcast_coalesced(tensors, group, src_global_rank, bucket_bytes):
    bucket, size = [], 0

    def flush():
        nonlocal bucket, size
        if not bucket:
            return
        by_dtype = {}
        for t in bucket:
            by_dtype.setdefault(t.dtype, []).append(t)
        for dt, ts in by_dtype.items():
            flat = torch.cat([t.reshape(-1) for t in ts])
            dist.broadcast(flat, src=src_global_rank, group=group)
            off = 0
            for t in ts:
                t.copy_(flat[off:off + t.numel()].view_as(t))
                off += t.numel()
        bucket, size = [], 0

    for t in tensors:
        bucket.append(t)
        size += t.numel() * t.element_size()
        if size >= bucket_bytes:
            flush()
    flush()
