"""smdistributed_modelparallel_b200 - a Blackwell-native model-parallel training library with the
public API of SageMaker's ``smdistributed.modelparallel.torch`` 1.12 (``import
smdistributed_modelparallel_b200 as smp``).

Export list mirrors the reference (reference smdistributed/modelparallel/torch/__init__.py:31-79;
``init`` :88-138; context-manager partials :153-163).
"""

from __future__ import annotations

import os

import torch

from . import amp, nn, optimizers  # noqa: F401  (sub-namespaces of the public API)
from .backend import exceptions  # noqa: F401
from .backend.collectives import CommGroup, RankType
from .backend.config import ModelParallelConfig
from .backend.core import ModelParallelCore
from .backend.exceptions import *  # noqa: F401,F403
from .backend.logger import get_logger
from .backend.split import StepOutput
from .checkpoint import load, resume_from_checkpoint, save, save_checkpoint
from .comm import (allgather, barrier, broadcast, dp_barrier, gather, get_dp_process_group, get_mp_process_group,
                   get_pp_process_group, get_rdp_process_group, get_tp_process_group, get_world_process_group,
                   mp_barrier, pp_barrier, rdp_barrier, recv_from, send, tp_barrier)
from .model import DistributedModel, model_creation
from .optimizers.optimizer import DistributedOptimizer
from .parameter import delay_param_initialization
from .runtime.checkpointing import checkpoint, checkpoint_sequential
from .state_mod import ModelParallelState, state
from .step import step

__version__ = "1.12.1+b200.0"

WORLD, PP_GROUP, TP_GROUP, RDP_GROUP, DP_GROUP, MP_GROUP = (CommGroup.WORLD, CommGroup.PP_GROUP, CommGroup.TP_GROUP,
                                                           CommGroup.RDP_GROUP, CommGroup.DP_GROUP, CommGroup.MP_GROUP)

core = ModelParallelCore()
ModelParallelState.__init__(state)


def init(config=None):
    """Initialise the library: validate ``config`` (merged with ``SM_HP_MP_PARAMETERS``), bootstrap ranks
    from the launcher environment, open the control-plane bus, create every process group."""
    cfg = ModelParallelConfig(config)
    for key in ("horovod", "herring"):
        if getattr(cfg, key, False):
            # silently training without gradient averaging would be worse than refusing (the reference itself
            # refuses Herring: torch/allreduce/herring.py:9-73)
            from .backend.exceptions import SMPUnsupportedError

            raise SMPUnsupportedError(f"'{key}: true': data parallelism runs on the built-in NCCL reducers here; "
                                      "set 'ddp: true' instead")
    if state.initialized:
        reset()
    core.initialize(cfg)
    state.initialize(cfg, core)
    state.init_model_specific_state()
    from .patches import patch_manager

    patch_manager.patch_constructor()
    state.tp_registry.register_builtin()
    if torch.cuda.is_available():
        torch.cuda.set_device(core.local_rank())
    # tensor-parallel collectives: the fused peer-memory kernels are the default path on GPUs (they still check
    # per call that the group sits on one NVSwitch domain and that the shapes tile); `fused_tp_collectives: false`
    # or SMP_TP_COMM=nccl selects the NCCL + GEMM composition (the measured baseline)
    from .parallel import tp_ops

    tp_ops.set_comm_mode(tp_ops.select_comm_mode(cfg, torch.cuda.is_available()))
    # the native tensor-communication runtime (gradient reducer, uneven collectives, p2p handles): a missing or
    # stale library is reported here, at start-up, not at the first collective that needs it
    from .backend import native_torch

    if native_torch.enabled():
        native_torch.load()
    if core.rank() == 0:
        cfg.display_config()


def is_initialized() -> bool:
    return bool(state.initialized)


def is_tracing() -> bool:
    return bool(state.is_tracing_flag)


def reset():
    """Forget the model / optimizer / partition (the process groups and the bus stay up)."""
    from .patches import patch_manager

    patch_manager.reset()
    if state.model is not None and getattr(state.model, "grad_buffer", None) is not None:
        state.model.grad_buffer.release()
    state.tp_registry.reset()
    state.init_model_specific_state()
    state.step_func = {}
    # device-side caches of the op wrappers (scratch buffers, rotary tables, deferred gathers)
    from .ops import attention as _attention
    from .ops import workspace as _workspace
    from .ops._sdp_hook import pending_gathers as _pending

    _workspace.release_all()
    _attention._table_cache.clear()
    _pending.clear()


def shutdown():
    core.shutdown(True)
    state.initialized = False


# ---- ranks and sizes
def rank(): return core.rank()
def size(): return core.size()
def local_rank(): return core.local_rank()
def local_size(): return core.local_size()
def dp_rank(): return core.dp_rank()
def dp_size(): return core.dp_size()
def pp_rank(): return core.pp_rank()
def pp_size(): return core.pp_size()
def tp_rank(): return core.tp_rank()
def tp_size(): return core.tp_size()
def rdp_rank(): return core.rdp_rank()
def rdp_size(): return core.rdp_size()
def mp_rank(): return core.mp_rank()
def mp_size(): return core.mp_size()
def num_microbatches(): return state.num_microbatches()


# ---- global ranks of the groups this rank belongs to (reference torch/core.py:30-34)
def get_dp_group(pp_rank=None): return core.get_dp_group(pp_rank)
def get_pp_group(dp_rank=None): return core.get_pp_group(dp_rank)
def get_mp_group(rdp_rank=None): return core.get_mp_group(rdp_rank)
def get_tp_group(pp_rank=None, rdp_rank=None): return core.get_tp_group(pp_rank, rdp_rank)
def get_rdp_group(pp_rank=None, tp_rank=None): return core.get_rdp_group(pp_rank, tp_rank)


def param_shard_rank():
    """Rank inside the sharded-data-parallel shard group (reference torch/core.py:37-60)."""
    deg = state.cfg.sharded_data_parallel_degree
    return core.rank() % deg if state.cfg.zero2d_enabled() else 0


def param_shard_size():
    return state.cfg.sharded_data_parallel_degree if state.cfg.zero2d_enabled() else 1


# ---- module-tree annotations
def partition(index):
    return state.module_manager.partition(index)


def set_partition(module, partition):
    return state.module_manager.set_partition(module, partition)


def tensor_parallelism(enabled=True, **tp_config):
    return state.module_manager.tensor_parallelism(enabled, **tp_config)


def set_tensor_parallelism(module, enabled=True, **tp_config):
    return state.module_manager.set_tensor_parallelism(module, enabled, **tp_config)


def set_activation_checkpointing(module, preserve_rng_state=True, pack_args_as_tuple=False, strategy="each"):
    return state.module_manager.set_activation_checkpointing(module, preserve_rng_state, pack_args_as_tuple, strategy)


def tp_register(dist_module, init_hook=None, forward_hook=None, return_hook=None):
    return state.tp_registry.register(dist_module, init_hook, forward_hook, return_hook)


def tp_register_with_module(module_cls, dist_module, init_hook=None, forward_hook=None, return_hook=None,
                            translate_functions=None):
    return state.tp_registry.register_with_module(module_cls, dist_module, init_hook, forward_hook, return_hook,
                                                  translate_functions)


__all__ = [
    "init", "is_initialized", "is_tracing", "reset", "shutdown", "DistributedModel", "DistributedOptimizer", "step",
    "StepOutput", "partition", "set_partition", "tensor_parallelism", "set_tensor_parallelism",
    "set_activation_checkpointing", "checkpoint", "checkpoint_sequential", "model_creation",
    "delay_param_initialization", "tp_register", "tp_register_with_module", "rank", "size", "local_rank",
    "local_size", "dp_rank", "dp_size", "pp_rank", "pp_size", "tp_rank", "tp_size", "rdp_rank", "rdp_size",
    "mp_rank", "mp_size", "num_microbatches", "param_shard_rank", "param_shard_size", "get_dp_group", "get_pp_group",
    "get_mp_group", "get_tp_group", "get_rdp_group", "get_dp_process_group",
    "get_mp_process_group", "get_pp_process_group", "get_tp_process_group", "get_rdp_process_group",
    "get_world_process_group", "CommGroup", "RankType", "WORLD", "PP_GROUP", "TP_GROUP", "DP_GROUP", "RDP_GROUP",
    "MP_GROUP", "broadcast", "send", "recv_from", "allgather", "gather", "barrier", "pp_barrier", "tp_barrier",
    "dp_barrier", "rdp_barrier", "mp_barrier", "save", "load", "save_checkpoint", "resume_from_checkpoint", "nn",
    "amp", "optimizers", "state", "core",
]

if os.environ.get("SMP_MANUAL_INIT", "1") == "0" and os.environ.get("SM_HP_MP_PARAMETERS"):
    init({})
