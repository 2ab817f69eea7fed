"""Process-wide state: config, core, communicator, process groups, managers, microbatch bookkeeping.

Plays the role of the reference's ``PTModelParallelState`` singleton (reference
smdistributed/modelparallel/torch/state_mod.py:31-418 on top of backend/state_mod.py:14-93).  Process
groups for WORLD / DP / MP / PP / TP / RDP are created on every rank in the same order (a
``torch.distributed`` requirement, reference state_mod.py:116-160); in addition one 2-rank group per
ordered pair of pipeline ranks carries activation / gradient p2p so opposite directions can never
block each other on one NCCL communicator.
"""

from __future__ import annotations

import itertools
from contextlib import contextmanager

import torch
import torch.distributed as dist

from .backend.collectives import CollectiveCommunicator, TransactionIdentifier
from .backend.exceptions import InvalidLinkIDError, SMPUnsupportedError


class ModelParallelState:
    def __init__(self):
        self.initialized = False
        self.cfg = None
        self.core = None
        self.comm = None
        self.step_func = {}
        self._reset_groups()
        self.init_model_specific_state()

    # ------------------------------------------------------------------ lifecycle
    def _reset_groups(self):
        self.world_process_group = None
        self.dp_process_group = None
        self.mp_process_group = None
        self.pp_process_group = None
        self.tp_process_group = None
        self.rdp_process_group = None
        self.sdp_process_group = None
        self.sdp_replica_process_group = None
        self.pp_p2p_groups = {}
        self._group_cache = {}

    def init_model_specific_state(self):
        from .pipeline.module_manager import ModuleManager
        from .nn.registry import TensorParallelismRegistry
        from .utils.random import RngManager

        self.microbatch = 0
        self.model = None
        self.optimizer = None
        self.current_step_func_id = None
        self.no_grad_context = False
        self.in_step_func = False
        self.is_tracing_flag = False
        self.loaded_model_state = None
        self.loaded_optimizer_state = None
        self.exec_server = None
        self.module_manager = ModuleManager()
        if not hasattr(self, "tp_registry") or self.tp_registry is None:
            self.tp_registry = TensorParallelismRegistry()
        self.rng_manager = RngManager()
        self.offloaders = {}
        self.checkpoint_activations_depth = 0
        self._tx_counter = itertools.count(64)
        self._link_counter = itertools.count(0)
        self.num_hops = 0
        self.metrics_uploaded = False
        if self.cfg is not None:
            self.module_manager.set_config(self.cfg)
            self.rng_manager.set_state(self.cfg.tensor_parallel_seed)

    def initialize(self, cfg, core):
        self.cfg = cfg
        self.core = core
        self.comm = CollectiveCommunicator(core)
        self._create_process_groups()
        self.create_pipeline()
        self.module_manager.set_config(cfg)
        self.rng_manager.set_state(cfg.tensor_parallel_seed)
        self.initialized = True

    def _new_group(self, ranks):
        key = tuple(ranks)
        if key not in self._group_cache:
            if len(ranks) == dist.get_world_size() and list(ranks) == list(range(len(ranks))):
                self._group_cache[key] = dist.group.WORLD
            else:
                self._group_cache[key] = dist.new_group(ranks=list(ranks))
        return self._group_cache[key]

    def _create_process_groups(self):
        core = self.core
        self._reset_groups()
        self.world_process_group = dist.group.WORLD
        me = core.rank()
        r = core.ranker
        # every rank walks every group of every family in the same order
        for kind, attr in (("dp", "dp_process_group"), ("mp", "mp_process_group"), ("pp", "pp_process_group"),
                           ("tp", "tp_process_group"), ("rdp", "rdp_process_group")):
            for ranks in r.all_groups(kind):
                g = self._new_group(ranks)
                if me in ranks:
                    setattr(self, attr, g)
        # directed p2p groups between pipeline ranks
        if core.pp_size() > 1:
            for ranks in r.all_groups("pp"):
                for i, a in enumerate(ranks):
                    for j, b in enumerate(ranks):
                        if i == j:
                            continue
                        g = dist.new_group(ranks=sorted((a, b)))
                        if me in (a, b) and me in ranks:
                            self.pp_p2p_groups[(a, b)] = g
        # sharded data parallelism: shard group (consecutive ranks) and replica group (same shard index)
        deg = self.cfg.sharded_data_parallel_degree
        if self.cfg.zero2d_enabled() and deg > 1:
            world = core.size()
            for start in range(0, world, deg):
                ranks = list(range(start, start + deg))
                g = self._new_group(ranks)
                if me in ranks:
                    self.sdp_process_group = g
            for off in range(deg):
                ranks = list(range(off, world, deg))
                g = self._new_group(ranks)
                if me in ranks:
                    self.sdp_replica_process_group = g

    def create_pipeline(self):
        from .runtime.pipeline import InterleavedPipeline, OnlyForwardPipeline, SimplePipeline

        kind = self.cfg.pipeline
        if kind == "simple":
            self.pipeline = SimplePipeline(self.cfg.microbatches, self.cfg.active_microbatches)
        elif kind == "interleaved":
            self.pipeline = InterleavedPipeline(self.cfg.microbatches, self.cfg.active_microbatches)
        elif kind == "_only_forward":
            self.pipeline = OnlyForwardPipeline(self.cfg.microbatches, self.cfg.active_microbatches)
        else:
            raise SMPUnsupportedError("Pipeline not supported!")

    # ------------------------------------------------------------------ Megatron-style getters
    # (reference state_mod.py:83-166: code written against Megatron's mpu can ask the state object)
    def get_data_parallel_group(self):
        return self.dp_process_group

    def get_data_parallel_rank(self) -> int:
        return self.core.dp_rank()

    def get_data_parallel_world_size(self) -> int:
        return self.core.dp_size()

    def get_model_parallel_group(self):
        return self.mp_process_group

    def get_model_parallel_rank(self) -> int:
        return self.core.mp_rank()

    def get_model_parallel_world_size(self) -> int:
        return self.core.mp_size()

    def is_tracing(self) -> bool:
        return bool(self.is_tracing_flag)

    # ------------------------------------------------------------------ ids
    def get_transaction_id(self) -> TransactionIdentifier:
        return TransactionIdentifier(next(self._tx_counter), False)

    def get_link_id(self) -> int:
        v = next(self._link_counter)
        if v >= (1 << 31):
            raise InvalidLinkIDError("link id space exhausted")
        return v

    # ------------------------------------------------------------------ microbatch bookkeeping
    def current_step_func(self):
        """Id of the step function being executed (reference torch/state_mod.py: keys of per-step-function records)."""
        return self.current_step_func_id

    def current_minibatch(self):
        srv = self.exec_server
        return getattr(srv, "_step", 0) if srv is not None else 0

    def num_microbatches(self) -> int:
        return self.cfg.microbatches

    def clear_minibatch_state(self):
        self.microbatch = 0
        if self.module_manager is not None:
            self.module_manager.clear_minibatch_state()
        for off in self.offloaders.values():
            off.reset()

    # ------------------------------------------------------------------ rng
    @contextmanager
    def fork_smp_rng_state(self, seed=None, enabled=True):
        """Run a block under the tensor-parallel RNG stream (same across tp ranks for replicated
        tensors), restoring the caller's generator afterwards (reference torch/state_mod.py:354-373)."""
        if not enabled:
            yield
            return
        with self.rng_manager.consistent_rng_state(seed):
            yield

    @property
    def device(self):
        if torch.cuda.is_available():
            # the process's GPU, not the calling thread's current device: pipeline tasks run in their own threads,
            # and a new thread starts on device 0
            core = getattr(self, "core", None)
            if core is not None and core.initialized:
                return torch.device("cuda", core.local_rank())
            return torch.device("cuda", torch.cuda.current_device())
        return torch.device("cpu")


state = ModelParallelState.__new__(ModelParallelState)
