"""Book-keeping about the user's module tree: names, parents, pipeline partition of every module,
tensor-parallel marks, activation-checkpoint settings, per-microbatch execution records.

Role of the reference's ``ModuleManager`` (reference smdistributed/modelparallel/torch/
module_manager.py:60-1392; API at :969-1010 set_activation_checkpointing, :1012-1042
set_tensor_parallelism, :1061-1092 set_partition, :1094-1114 tensor_parallelism, :1160-1178 partition,
:1189-1231 naming, :1337-1360 metrics).  Modules are recorded as they are constructed (the constructor
hook in ``patches``) and named ``main/child/grandchild`` once the DistributedModel exists.
"""

from __future__ import annotations

import weakref
from collections import defaultdict
from contextlib import contextmanager
from dataclasses import dataclass, field

import torch.nn as nn

from ..backend.exceptions import (CheckpointingConfigError, InvalidPartitionIDError, MissingModuleError,
                                  SMPInvalidArgumentError)


@dataclass
class CheckpointConfig:
    enabled: bool = True
    preserve_rng_state: bool = True
    module_name: str = ""
    strategy: str = "each"
    pack_args_as_tuple: bool = False


@dataclass
class TraceResult:
    order: list = field(default_factory=list)        # module names in first-execution order
    times: dict = field(default_factory=dict)        # name -> seconds
    memory: dict = field(default_factory=dict)       # name -> bytes of activations produced
    out_bytes: dict = field(default_factory=dict)    # name -> bytes of outputs
    in_bytes: dict = field(default_factory=dict)     # name -> bytes of inputs


class ModuleManager:
    def __init__(self):
        self.cfg = None
        self._partition_ctx = None           # partition id set by `with smp.partition(i)`
        self._tp_ctx = None                  # (enabled, config) set by `with smp.tensor_parallelism(...)`
        self._module_partition = weakref.WeakKeyDictionary()     # module -> manual partition id
        self._tp_marks = weakref.WeakKeyDictionary()             # module -> (enabled, config dict)
        self._ckpt_configs = weakref.WeakKeyDictionary()         # module -> CheckpointConfig
        self._names = weakref.WeakKeyDictionary()                # module -> "main/..."
        self._by_name = {}
        self._parents = weakref.WeakKeyDictionary()              # module -> set(parent modules)
        self._partition_by_name = {}                             # final assignment name -> pp_rank
        self.partitioned = False
        self.trace = TraceResult()
        self.comm_volume_bytes = 0
        self.hops = 0
        self._exec_counts = defaultdict(int)                     # (mb, name) -> calls this step

    def set_config(self, cfg):
        self.cfg = cfg

    # ------------------------------------------------------------------ construction-time hooks
    def record_module_construction(self, module: nn.Module):
        """Called from the patched ``nn.Module.__init__``."""
        if self._partition_ctx is not None:
            self._module_partition[module] = self._partition_ctx
        if self._tp_ctx is not None:
            self._tp_marks[module] = self._tp_ctx

    @contextmanager
    def partition(self, index: int):
        self._check_partition_id(index)
        prev, self._partition_ctx = self._partition_ctx, index
        try:
            yield
        finally:
            self._partition_ctx = prev

    @contextmanager
    def tensor_parallelism(self, enabled: bool = True, **tp_config):
        prev, self._tp_ctx = self._tp_ctx, (bool(enabled), dict(tp_config))
        try:
            yield
        finally:
            self._tp_ctx = prev

    def _check_partition_id(self, index):
        pp = self.cfg.pipeline_parallel_degree if self.cfg is not None else None
        if not isinstance(index, int) or index < 0 or (pp is not None and index >= pp):
            raise InvalidPartitionIDError(f"partition id must be an int in [0, {pp}), got {index!r}")

    # ------------------------------------------------------------------ explicit setters (public API)
    def set_partition(self, module: nn.Module, partition: int):
        self._check_partition_id(partition)
        if self.partitioned:
            from ..backend.exceptions import StepFunctionCalledError
            raise StepFunctionCalledError("smp.set_partition must be called before the first step")
        self._module_partition[module] = partition

    def set_tensor_parallelism(self, module: nn.Module, enabled: bool = True, **tp_config):
        for m in module.modules():
            self._tp_marks[m] = (bool(enabled), dict(tp_config))

    def set_activation_checkpointing(self, module: nn.Module, preserve_rng_state=True, pack_args_as_tuple=False,
                                     strategy="each"):
        if not isinstance(module, nn.Module):
            raise CheckpointingConfigError("set_activation_checkpointing expects an nn.Module")
        if isinstance(module, nn.Sequential):
            ok = strategy in ("each", "contiguous") or (isinstance(strategy, str) and strategy.startswith("group_")
                                                         and strategy[6:].isdigit() and int(strategy[6:]) > 0)
            if not ok:
                raise CheckpointingConfigError(
                    f"strategy must be 'each', 'contiguous' or 'group_N' for nn.Sequential, got {strategy!r}")
        if self.cfg is not None and self.cfg.zero2d_enabled():
            own = {id(p) for p in module.parameters()}
            # a checkpointed module may not share parameters with modules outside it under sharded DP
            root = getattr(self, "_root", None)
            if root is not None:
                for other in root.modules():
                    if other is module or any(other is m for m in module.modules()):
                        continue
                    if any(id(p) in own for p in other.parameters(recurse=False)):
                        raise CheckpointingConfigError(
                            "cannot checkpoint a module that shares parameters with another module when sharded "
                            "data parallelism is enabled")
        self._ckpt_configs[module] = CheckpointConfig(True, preserve_rng_state, self._names.get(module, ""), strategy,
                                                      pack_args_as_tuple)

    def get_checkpoint_config(self, module):
        return self._ckpt_configs.get(module)

    def should_checkpoint_activations(self, module) -> bool:
        c = self._ckpt_configs.get(module)
        return bool(c and c.enabled)

    # ------------------------------------------------------------------ tensor-parallel marks
    def get_tp_mark(self, module):
        return self._tp_marks.get(module)

    def is_marked_for_tp(self, module) -> bool:
        m = self._tp_marks.get(module)
        return bool(m and m[0])

    def get_tp_config(self, module) -> dict:
        m = self._tp_marks.get(module)
        return dict(m[1]) if m else {}

    def simplify_tensor_parallelism_modules(self, root: nn.Module, supported):
        """Keep only the outermost marked modules that have a distributed counterpart, and drop marks on
        modules sharing parameters with modules outside themselves (reference module_manager.py:1116-1158)."""
        def visit(m, inside_marked):
            marked = self.is_marked_for_tp(m) and supported(m)
            if inside_marked and m in self._tp_marks:
                del self._tp_marks[m]
                marked = False
            elif self.is_marked_for_tp(m) and not supported(m):
                marked = False
            for c in m.children():
                visit(c, inside_marked or marked)

        visit(root, False)

    # ------------------------------------------------------------------ naming / tree
    def name_modules_and_create_parent_map(self, root: nn.Module):
        self._root = root
        self._names = weakref.WeakKeyDictionary()
        self._by_name = {}
        self._parents = weakref.WeakKeyDictionary()

        def walk(m, name):
            if m not in self._names:            # first path wins for reused modules
                self._names[m] = name
                self._by_name[name] = m
            for cname, c in m.named_children():
                self._parents.setdefault(c, set()).add(m)
                walk(c, f"{name}/{cname}")

        walk(root, "main")

    def get_module_name(self, module) -> str:
        try:
            return self._names[module]
        except KeyError:
            raise MissingModuleError(f"module {type(module).__name__} is not part of the DistributedModel") from None

    def has_module(self, module) -> bool:
        return module in self._names

    def get_module(self, name: str) -> nn.Module:
        try:
            return self._by_name[name]
        except KeyError:
            raise MissingModuleError(f"no module named {name}") from None

    def get_parent_modules(self, module):
        return list(self._parents.get(module, ()))

    def module_names(self):
        return list(self._by_name)

    # ---- read-only accessors under the reference's names (torch/module_manager.py:330-520)
    def get_module_names(self):
        return self.module_names()

    def get_parent_module(self, module):
        parents = self.get_parent_modules(module)
        return parents[0] if parents else None

    def is_main_module(self, module) -> bool:
        return module is getattr(self, "_root", None)

    def get_parameters(self, module, recurse=True):
        return list(module.parameters(recurse=recurse))

    @staticmethod
    def is_distributed(tensor) -> bool:
        from ..nn.utils import is_distributed_parameter
        return is_distributed_parameter(tensor)

    is_distributed_parameter = is_distributed_buffer = is_distributed

    @staticmethod
    def is_scaled_batch_parameter(tensor) -> bool:
        from ..nn.utils import is_scaled_batch_parameter
        return is_scaled_batch_parameter(tensor)

    is_scaled_batch_buffer = is_scaled_batch_parameter

    @staticmethod
    def is_one_rank_parameter(tensor) -> bool:
        from ..nn.utils import is_one_rank_parameter
        return is_one_rank_parameter(tensor)

    is_one_rank_buffer = is_one_rank_parameter

    @staticmethod
    def get_parameter_distribution_axis(tensor):
        return getattr(tensor, "_smp_dist_axis", None)

    get_buffer_distribution_axis = get_parameter_distribution_axis

    # ------------------------------------------------------------------ partition map
    def manual_partition_of(self, module):
        return self._module_partition.get(module)

    def assign_partitions(self, mapping: dict):
        """mapping: module name -> pp_rank for every named module."""
        pp = self.cfg.pipeline_parallel_degree
        for n, r in mapping.items():
            if not (0 <= r < pp):
                raise InvalidPartitionIDError(f"module {n} assigned to invalid partition {r}")
        self._partition_by_name = dict(mapping)
        self.partitioned = True

    def get_partition(self, module) -> int:
        if not self._partition_by_name:
            return 0
        return self._partition_by_name[self.get_module_name(module)]

    def get_partition_by_name(self, name) -> int:
        return self._partition_by_name.get(name, 0)

    def partition_info(self) -> dict:
        return dict(self._partition_by_name)

    def get_model_partition_info(self) -> dict:
        """Module name -> pipeline rank of the partitioned model: what ``load_saved_partition`` takes (reference
        torch/module_manager.py:1054-1059, test/torch/mpi_4ps/test_load_partition.py)."""
        return self.partition_info()

    def load_partition(self, partition_info):
        """Adopt a saved assignment instead of partitioning: it wins over ``auto_partition`` and over
        ``smp.set_partition`` calls made afterwards; modules the saved assignment does not name follow their
        closest named ancestor.  Only before the first step."""
        if self.partitioned:
            from ..backend.exceptions import StepFunctionCalledError
            raise StepFunctionCalledError("a saved partition must be loaded before the first step")
        loaded = {}
        for name, part in dict(partition_info).items():
            self._check_partition_id(int(part))
            loaded[str(name)] = int(part)
        self._loaded_partition = loaded

    def loaded_partition(self):
        return getattr(self, "_loaded_partition", None)

    def assignment_from_saved(self, loaded) -> dict:
        default = self.cfg.default_partition if (self.cfg is not None and self.cfg.default_partition is not None) else 0
        out = {}
        for name in self._by_name:
            probe = name
            while probe not in loaded and "/" in probe:
                probe = probe.rsplit("/", 1)[0]
            out[name] = loaded.get(probe, default)
        return out

    def is_executor(self, module, pp_rank) -> bool:
        return self.get_partition(module) == pp_rank

    def local_module_names(self, pp_rank):
        return [n for n, r in self._partition_by_name.items() if r == pp_rank] if self._partition_by_name \
            else list(self._by_name)

    # ------------------------------------------------------------------ per-step records / metrics
    def record_execution(self, mb, name):
        self._exec_counts[(mb, name)] += 1
        return self._exec_counts[(mb, name)] - 1

    def clear_minibatch_state(self):
        self._exec_counts.clear()

    def record_hop(self, nbytes):
        self.hops += 1
        self.comm_volume_bytes += int(nbytes)

    def get_metrics(self, pp_rank=0):
        total = sum(p.numel() for p in self._root.parameters()) if getattr(self, "_root", None) is not None else 0
        local_names = set(self.local_module_names(pp_rank))
        local_params = 0
        seen = set()
        for n in local_names:
            for p in self._by_name[n].parameters(recurse=False):
                if id(p) not in seen:
                    seen.add(id(p))
                    local_params += p.numel()
        return {
            "total_communication_volume_mb": self.comm_volume_bytes / 2 ** 20,
            "num_hops_between_devices": self.hops,
            "parameter_count": local_params,
            "total_parameter_count": total,
            "module_fraction": (len(local_names) / max(1, len(self._by_name))),
        }

    def validate_manual_partitions(self):
        if self.cfg is None or self.cfg.auto_partition:
            return
        if self.cfg.default_partition is None:
            raise SMPInvalidArgumentError("auto_partition=False requires default_partition")
