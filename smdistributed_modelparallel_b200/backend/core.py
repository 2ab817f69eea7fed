"""Process bootstrap, rank/group queries, timeline and metrics accessors.

Same public surface as the reference's ``ModelParallelCore`` (reference
smdistributed/modelparallel/backend/core.py:191-562) with a different substrate: ranks come from the
launcher environment (``torchrun``: RANK / WORLD_SIZE / LOCAL_RANK; OpenMPI variables are honoured
too) instead of MPI_Init inside a native library, the control plane is the TCP message bus of
``libsmpb200_core.so``, and collectives run on ``torch.distributed``.
"""

from __future__ import annotations

import atexit
import ctypes
import os
import socket
import sys

import torch
import torch.distributed as dist

from . import native_core
from .exceptions import InvalidEnvironmentError, NotInitializedError, WorkerSizeError
from .logger import get_logger
from .topology import Ranker


def _env_int(*names, default=None):
    for n in names:
        v = os.environ.get(n)
        if v is not None and v != "":
            return int(v)
    return default


def _my_address(master_addr: str, master_port: int) -> str:
    """IP of the interface that routes to the rendezvous host (127.0.0.1 for single-node runs)."""
    if master_addr in ("127.0.0.1", "localhost", "::1"):
        return "127.0.0.1"
    try:
        s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
        s.connect((master_addr, master_port or 29500))
        ip = s.getsockname()[0]
        s.close()
        return ip
    except OSError:
        return "127.0.0.1"


class ExitHook:
    """Records a non-zero ``sys.exit`` or an uncaught exception so shutdown can tell the peers."""

    def __init__(self):
        self.exit_code = None
        self.exception = None
        self._installed = False

    def hook(self):
        if self._installed:
            return
        self._installed = True
        self._orig_exit = sys.exit
        self._orig_excepthook = sys.excepthook

        def _exit(code=0):
            self.exit_code = code
            self._orig_exit(code)

        def _exc(exc_type, exc, *args):
            self.exception = exc
            self._orig_excepthook(exc_type, exc, *args)

        sys.exit = _exit
        sys.excepthook = _exc

    @property
    def success(self) -> bool:
        return self.exception is None and self.exit_code in (None, 0)


class ModelParallelCore:
    def __init__(self):
        self._initialized = False
        self.cfg = None
        self.ranker: Ranker | None = None
        self.exit_hook = ExitHook()
        self._lib = None
        self._rank = 0
        self._size = 1
        self._local_rank = 0
        self._local_size = 1
        self._node_of: list[int] = [0]
        self._host_group = None
        self._atexit_registered = False
        self.bus_enabled = False

    # ------------------------------------------------------------------ lifecycle
    @property
    def initialized(self) -> bool:
        return self._initialized

    def _validate_initialized(self):
        if not self._initialized:
            raise NotInitializedError()

    def initialize(self, cfg, timeline_path: str | None = None):
        log = get_logger()
        self.cfg = cfg
        self._lib = native_core.load()
        rank = _env_int("RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK")
        size = _env_int("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE")
        if (rank is None) != (size is None):
            raise InvalidEnvironmentError("RANK and WORLD_SIZE must both be set by the launcher (torchrun) or both unset")
        if rank is None:
            rank, size = 0, 1
        self._rank, self._size = rank, size
        self._local_rank = _env_int("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", default=0)
        self._local_size = _env_int("LOCAL_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_SIZE", default=size if size <= 8 else 8)

        use_cuda = torch.cuda.is_available()
        if use_cuda:
            ndev = torch.cuda.device_count()
            if self._local_rank >= ndev:
                raise WorkerSizeError(f"local rank {self._local_rank} has no GPU ({ndev} visible): fewer devices than "
                                      "processes per node")
            torch.cuda.set_device(self._local_rank)

        # torch.distributed world (NCCL on GPUs, gloo on CPU) - created once, reused across smp.reset()
        if not dist.is_initialized():
            backend = "nccl" if (use_cuda and cfg.ddp_dist_backend == "nccl") else "gloo"
            if size == 1 and "MASTER_ADDR" not in os.environ:
                store = dist.HashStore()
                dist.init_process_group(backend, store=store, rank=0, world_size=1)
            else:
                os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
                if cfg.ddp_port is not None:
                    os.environ["MASTER_PORT"] = str(cfg.ddp_port)
                os.environ.setdefault("MASTER_PORT", "29500")
                kw = {}
                if backend == "nccl":
                    kw["device_id"] = torch.device("cuda", self._local_rank)
                dist.init_process_group(backend, rank=rank, world_size=size, **kw)
        if dist.get_world_size() != size:
            raise InvalidEnvironmentError("torch.distributed world size disagrees with the launcher environment")
        # host-side (gloo) group for object exchange during bootstrap
        if self._host_group is None:
            self._host_group = dist.new_group(backend="gloo") if dist.get_backend() != "gloo" else dist.group.WORLD

        self._validate_worker_size()
        tp, pp = cfg.tensor_parallel_degree, cfg.pipeline_parallel_degree
        self.ranker = Ranker(cfg.placement_strategy, size // (tp * pp), pp, tp)

        # control-plane bus: full TCP mesh keyed by (peer, transaction id)
        if not self._lib.smpc_bus_initialized():
            # listen on the interface that reaches the rendezvous host only; peers must present the job's token
            # (drawn by rank 0, handed out through the process group) before their frames are read
            import secrets

            master = os.environ.get("MASTER_ADDR", "127.0.0.1")
            my_ip = _my_address(master, _env_int('MASTER_PORT', default=29500))
            local_only = master in ("127.0.0.1", "localhost", "::1")
            # (a non-loopback rendezvous whose route could not be determined falls back to every interface)
            bind_ip = my_ip if (local_only or my_ip != "127.0.0.1") else ""
            port = self._lib.smpc_bus_listen_on(bind_ip.encode())
            if port <= 0:
                raise InvalidEnvironmentError("message bus could not open a listening socket")
            me = f"{my_ip}:{port}"
            node_name = socket.gethostname()
            gathered = [None] * size
            mine = (me, node_name, secrets.randbits(63))
            if size > 1:
                dist.all_gather_object(gathered, mine, group=self._host_group)
            else:
                gathered = [mine]
            self._lib.smpc_bus_set_token(int(gathered[0][2]))
            peers = ",".join(g[0] for g in gathered)
            names = [g[1] for g in gathered]
            uniq = {n: i for i, n in enumerate(dict.fromkeys(names))}
            self._node_of = [uniq[n] for n in names]
            rc = self._lib.smpc_bus_connect(rank, size, peers.encode(), 120)
            if rc != 0:
                raise InvalidEnvironmentError(f"message bus mesh setup failed (code {rc})")
        self.bus_enabled = True

        self.exit_hook.hook()
        if not self._atexit_registered:
            atexit.register(self.shutdown)
            self._atexit_registered = True

        cfg.construct_zero2d_config_dict(self)
        tl = timeline_path or os.environ.get("SMP_TIMELINE_PATH")
        if tl:
            self._lib.smpc_timeline_create(f"{tl}.rank{rank}.json".encode(), rank)
        os.environ["RANK"] = str(rank)
        os.environ["LOCAL_RANK"] = str(self._local_rank)
        os.environ["WORLD_SIZE"] = str(size)
        self._initialized = True
        if rank == 0:
            log.debug(f"initialized: world {size}, pp {pp}, tp {tp}, rdp {self.rdp_size()}, placement {self.ranker.order}")

    def _validate_worker_size(self):
        cfg, size = self.cfg, self._size
        tp, pp = cfg.tensor_parallel_degree, cfg.pipeline_parallel_degree
        if size % (tp * pp) != 0:
            raise WorkerSizeError(
                f"The number of processes ({size}) must be divisible by pipeline_parallel_degree x "
                f"tensor_parallel_degree ({pp} x {tp}).")
        if size // (tp * pp) > 1 and not (cfg.ddp or cfg.horovod or cfg.zero2d_enabled()):
            raise WorkerSizeError(
                f"{size} processes with pp={pp}, tp={tp} implies data parallelism; enable it with ddp=True.")
        if cfg.zero2d_enabled() and size % cfg.sharded_data_parallel_degree != 0:
            raise WorkerSizeError(
                f"sharded_data_parallel_degree ({cfg.sharded_data_parallel_degree}) must divide the number of "
                f"processes ({size}).")

    def shutdown(self, success: bool | None = None):
        if self._lib is None:
            return
        ok = self.exit_hook.success if success is None else success
        try:
            if self._lib.smpc_timeline_enabled():
                self._lib.smpc_timeline_flush()
            self._lib.smpc_bus_shutdown(1 if ok else 0)
        except Exception:  # pragma: no cover - interpreter teardown
            pass
        self.bus_enabled = False

    # ------------------------------------------------------------------ ranks and sizes
    def rank(self): self._validate_initialized(); return self._rank
    def size(self): self._validate_initialized(); return self._size
    def local_rank(self): self._validate_initialized(); return self._local_rank
    def local_size(self): self._validate_initialized(); return self._local_size
    def pp_rank(self): return self.ranker.get_pp_rank(self.rank())
    def tp_rank(self): return self.ranker.get_tp_rank(self.rank())
    def rdp_rank(self): return self.ranker.get_rdp_rank(self.rank())
    def dp_rank(self): return self.ranker.get_dp_rank(self.rank())
    def mp_rank(self): return self.ranker.get_mp_rank(self.rank())
    def pp_size(self): self._validate_initialized(); return self.cfg.pipeline_parallel_degree
    def tp_size(self): self._validate_initialized(); return self.cfg.tensor_parallel_degree
    def rdp_size(self): return self.size() // (self.pp_size() * self.tp_size())
    def dp_size(self): return self.rdp_size() * self.tp_size()
    def mp_size(self): return self.pp_size() * self.tp_size()

    # ------------------------------------------------------------------ groups (lists of global ranks)
    def get_pp_group(self, dp_rank=None):
        r = self.rank() if dp_rank is None else self.ranker.translate(
            0, self.ranker.get_tp_rank_from_dp_rank(dp_rank), self.ranker.get_rdp_rank_from_dp_rank(dp_rank))
        return self.ranker.get_pp_group(r)

    def get_dp_group(self, pp_rank=None):
        r = self.rank() if pp_rank is None else self.ranker.translate(pp_rank, 0, 0)
        return self.ranker.get_dp_group(r)

    def get_mp_group(self, rdp_rank=None):
        r = self.rank() if rdp_rank is None else self.ranker.translate(0, 0, rdp_rank)
        return self.ranker.get_mp_group(r)

    def get_tp_group(self, pp_rank=None, rdp_rank=None):
        if pp_rank is None and rdp_rank is None:
            return self.ranker.get_tp_group(self.rank())
        pp = self.pp_rank() if pp_rank is None else pp_rank
        rdp = self.rdp_rank() if rdp_rank is None else rdp_rank
        return self.ranker.get_tp_group(self.ranker.translate(pp, 0, rdp))

    def get_rdp_group(self, pp_rank=None, tp_rank=None):
        if pp_rank is None and tp_rank is None:
            return self.ranker.get_rdp_group(self.rank())
        pp = self.pp_rank() if pp_rank is None else pp_rank
        tp = self.tp_rank() if tp_rank is None else tp_rank
        return self.ranker.get_rdp_group(self.ranker.translate(pp, tp, 0))

    def pp_rank_to_rank(self, pp_rank): return self.ranker.translate(pp_rank, self.tp_rank(), self.rdp_rank())
    def tp_rank_to_rank(self, tp_rank): return self.ranker.translate(self.pp_rank(), tp_rank, self.rdp_rank())
    def rdp_rank_to_rank(self, rdp_rank): return self.ranker.translate(self.pp_rank(), self.tp_rank(), rdp_rank)

    def dp_rank_to_rank(self, dp_rank):
        return self.ranker.translate(self.pp_rank(), self.ranker.get_tp_rank_from_dp_rank(dp_rank),
                                     self.ranker.get_rdp_rank_from_dp_rank(dp_rank))

    def mp_rank_to_rank(self, mp_rank):
        return self.ranker.translate(self.ranker.get_pp_rank_from_mp_rank(mp_rank),
                                     self.ranker.get_tp_rank_from_mp_rank(mp_rank), self.rdp_rank())

    def is_in_same_instance(self, rank): return self._node_of[rank] == self._node_of[self.rank()]
    def is_multi_node(self): return len(set(self._node_of)) > 1
    def instance_id(self): return self._node_of[self.rank()]

    def barrier(self):
        self._validate_initialized()
        if self._size > 1:
            dist.barrier(group=self._host_group)

    @property
    def host_group(self):
        return self._host_group

    # ------------------------------------------------------------------ timeline / metrics
    def timeline_start_step(self):
        if self._lib is not None:
            self._lib.smpc_timeline_start_step()

    def timeline_end_step(self):
        if self._lib is not None:
            self._lib.smpc_timeline_end_step()

    def timeline_record_pipeline_event(self, microbatch, label):
        if self._lib is not None and self._lib.smpc_timeline_enabled():
            self._lib.smpc_timeline_record_pipeline_event(int(microbatch), str(label).encode())

    def timeline_flush(self):
        if self._lib is not None and self._lib.smpc_timeline_enabled():
            return self._lib.smpc_timeline_flush()
        return -1

    def pool_record(self, delta_alloc, delta_reserved, ok=True):
        """Account bytes entering (+) / leaving (-) the receive-side device buffers of the pipeline transport."""
        if self._lib is not None:
            self._lib.smpc_pool_record(int(delta_alloc), int(delta_reserved), 1 if ok else 0)

    def get_and_reset_memory_stats(self):
        out = (ctypes.c_double * 4)()
        self._lib.smpc_get_and_reset_memory_metrics(out)
        return {"d2d_peak_allocated_mb": out[0], "d2d_peak_reserved_mb": out[1], "gpu_free_mb": out[2],
                "gpu_total_mb": out[3]}

    def get_and_reset_alloc_stats(self):
        out = (ctypes.c_longlong * 2)()
        self._lib.smpc_get_and_reset_alloc_metrics(out)
        return {"success": out[0], "failure": out[1]}
