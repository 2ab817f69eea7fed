"""User-facing communication API: object collectives (tensors inside the objects travel as tensors) and
process-group getters (reference smdistributed/modelparallel/torch/comm.py:21-115 and
torch/collectives.py:20-212).

Objects are pickled over the control-plane bus; tensors found inside are moved to host memory for the
trip and restored on the receiver's current device - these calls are for state dicts, metrics and
hand-shakes, not for the training hot path (which uses NCCL / peer-memory kernels directly).
"""

from __future__ import annotations

import torch

from .backend.collectives import CommGroup, RankType
from .backend.exceptions import InvalidCommGroupError, NotInitializedError
from .backend.split import map_structure
from .state_mod import state


def _check():
    if not state.initialized:
        raise NotInitializedError()


class _TensorOnWire:
    __slots__ = ("cpu", "was_cuda")

    def __init__(self, t):
        self.was_cuda = t.is_cuda
        self.cpu = t.detach().cpu()


def _pack(obj):
    return map_structure(lambda x: _TensorOnWire(x) if torch.is_tensor(x) else x, obj)


def _unpack(obj):
    dev = state.device
    def f(x):
        if isinstance(x, _TensorOnWire):
            return x.cpu.to(dev) if x.was_cuda and dev.type == "cuda" else x.cpu
        return x
    return map_structure(f, obj)


def _group(g):
    if not isinstance(g, CommGroup):
        raise InvalidCommGroupError(f"expected a smp.CommGroup, got {g!r}")
    return g


def broadcast(obj, group=CommGroup.WORLD):
    _check()
    state.comm.broadcast(_pack(obj), _group(group))


def send(obj, dest_rank, rank_type=RankType.WORLD_RANK):
    _check()
    state.comm.send(_pack(obj), dest_rank, rank_type)


def recv_from(src_rank, rank_type=RankType.WORLD_RANK):
    _check()
    return _unpack(state.comm.recv_from(src_rank, rank_type))


def allgather(obj, group=CommGroup.WORLD):
    _check()
    return [_unpack(o) for o in state.comm.allgather(_pack(obj), _group(group))]


def gather(obj, group=CommGroup.WORLD, rank=0):
    _check()
    out = state.comm.gather(_pack(obj), _group(group), rank)
    return None if out is None else [_unpack(o) for o in out]


def barrier(group=CommGroup.WORLD):
    _check()
    state.comm.barrier(_group(group))


def pp_barrier(): barrier(CommGroup.PP_GROUP)
def tp_barrier(): barrier(CommGroup.TP_GROUP)
def dp_barrier(): barrier(CommGroup.DP_GROUP)
def rdp_barrier(): barrier(CommGroup.RDP_GROUP)
def mp_barrier(): barrier(CommGroup.MP_GROUP)


def get_world_process_group(): _check(); return state.world_process_group
def get_dp_process_group(): _check(); return state.dp_process_group
def get_mp_process_group(): _check(); return state.mp_process_group
def get_pp_process_group(): _check(); return state.pp_process_group
def get_tp_process_group(): _check(); return state.tp_process_group
def get_rdp_process_group(): _check(); return state.rdp_process_group
