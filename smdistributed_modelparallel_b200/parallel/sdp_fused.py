"""Peer-memory fabric of the sharded-data-parallel engine (CUDA, shard group inside one NVSwitch domain).

Holds, in symmetric (CUDA-IPC mapped) memory shared by the shard group:

* every unit's 16-bit parameter shard - peers read them directly, there is no send side;
* ``NSLOT`` rotating fp32 unit-gradient buffers - the owner's wgrad GEMMs accumulate into one, the peers'
  reduce kernels read their stripes out of it;
* epoch flags: ``params_ready[n]`` (shard j final for optimizer step e), and per gradient slot
  ``grad_ready[n]`` / ``consumed[n]``.

and launches the three kernels that replace NCCL all-gather / reduce-scatter + optimizer here:

* ``smpk_gemm_comm`` mode 3 - parameter all-gather fused into the first GEMM consuming the unit;
* ``smpk_sdp_gather`` - the same gather stand-alone, on a side stream (backward prefetch);
* ``smpk_sdp_reduce_adam`` - gradient reduce-scatter by peer loads fused with scale / accumulate and, when
  nothing has to be known globally first (no clipping, no dynamic loss scale), the Adam update.

(reference: engine delegated to DeepSpeed - torch/optimizers/optimizer.py:46-49,471-491; knobs
backend/ds_config_defaults.json:6-27)
"""

from __future__ import annotations

import ctypes
import os
import sys

import torch
import torch.distributed as dist

from ..ops import _native as N
from .fused_tp import MAX_RANKS, SymmetricBuffer, _CommDesc, _ptr_array

NSLOT = 3                     # gradient slots allocated up front
MAX_SLOTS = 8
_DEBUG = os.environ.get("SMP_SDP_DEBUG", "0") == "1"


def _dbg(fab, *what):
    if _DEBUG:
        print(f"[sdp r{fab.rank}]", *what, file=sys.stderr, flush=True)

COPY_ROWS = 8
DEFAULT_ROW = 4096            # row width (elements) of units without a designated first weight

from ..ops._sdp_hook import pending_gathers  # data_ptr -> callable(gemm_args) run by the consuming GEMM


def supported(engine) -> bool:
    if not torch.cuda.is_available() or engine.device.type != "cuda" or not N.available():
        return False
    if engine.degree < 2 or engine.degree > MAX_RANKS:
        return False
    core = engine.state.core
    base = core.rank() - engine.shard_rank
    return all(core.is_in_same_instance(base + j) for j in range(engine.degree))


class PeerFabric:
    def __init__(self, engine):
        self.engine = engine
        self.n, self.rank, self.pg = engine.degree, engine.shard_rank, engine.shard_group
        self.dev = engine.device
        units = engine.units
        es = torch.empty((), dtype=units[0].dtype).element_size() if units else 2
        self.shard_off = {}
        total = 0
        for u in units:
            self.shard_off[u] = total
            total += (u.per * es + 255) // 256 * 256
        self.shards = SymmetricBuffer(max(total, 256), None, self.rank, None, pg=self.pg)
        self.max_numel = max((u.numel for u in units), default=8)
        self.slot_bytes = (self.max_numel * 4 + 255) // 256 * 256
        # gradient slots: one symmetric allocation each; the pool grows (collectively, same order on every rank)
        # when more units are in backward at once than there are slots - e.g. a tied embedding whose gradient
        # stays open from the LM head to the very end of backward
        self.slots = []
        # flag words: params_ready[8] | slot s: grad_ready[8] consumed[8]
        self.flags = SymmetricBuffer((1 + 2 * MAX_SLOTS) * MAX_RANKS * 4, None, self.rank, None, pg=self.pg)
        max_rows = max((u.numel // u.row for u in units), default=64)
        self.blk_flags = torch.zeros(max_rows // 64 + 8, dtype=torch.int32, device=self.dev)
        self.counter = torch.zeros(MAX_SLOTS, dtype=torch.int32, device=self.dev)
        self.param_epoch = 0
        for _ in range(NSLOT):
            self._add_slot()
        self.side = torch.cuda.Stream(device=self.dev)
        self.reduce_stream = torch.cuda.Stream(device=self.dev)
        self.reduce_events = []
        for u in units:                              # move the shards into peer-visible memory
            view = self.shards.view(u.dtype, u.per, self.shard_off[u])
            view.copy_(u.shard)
            u.shard = view
        self.publish_params()
        torch.cuda.synchronize()
        dist.barrier(group=self.pg)

    # ------------------------------------------------------------------ flags
    def _flag_ptr(self, peer, block):
        return self.flags.peer_ptrs[peer] + 4 * MAX_RANKS * block

    def publish_params(self):
        """Everything queued on the current stream so far (optimizer writes of the 16-bit shard) becomes
        visible to the peers as parameter epoch ``param_epoch + 1``."""
        self.param_epoch += 1
        _dbg(self, "publish params epoch", self.param_epoch)
        peers = _ptr_array([self._flag_ptr(j, 0) for j in range(self.n)])
        N.call("smpk_flags_signal", N.stream_ptr(self.dev), ctypes.cast(peers, ctypes.c_void_p), self.n, self.rank,
               self.param_epoch)
        N.count_launch()

    # ------------------------------------------------------------------ parameter gather
    def _peer_shards(self, u):
        return [self.shards.peer_ptrs[j] + self.shard_off[u] for j in range(self.n)]

    def gather(self, u, full, stream=None):
        """Stand-alone striped all-gather of ``u`` into ``full`` (current or given stream)."""
        es = full.element_size()
        _dbg(self, "gather", u.name, "epoch", self.param_epoch)
        ps = _ptr_array(self._peer_shards(u))
        N.call("smpk_sdp_gather", N.stream_ptr(self.dev) if stream is None else stream.cuda_stream,
               ctypes.cast(ps, ctypes.c_void_p), full.data_ptr(), u.piece * es, u.ppr, self.flags.ptr, self.param_epoch,
               self.n, self.rank, 0)
        N.count_launch()

    def defer_to_gemm(self, u, full):
        """Register ``full`` so that the next GEMM whose B operand is the unit's first weight gathers the unit."""
        key = full.data_ptr()

        def run(gemm_args):
            pending_gathers.pop(key, None)
            u.lazy_pending = False
            _dbg(self, "gemm-gather", u.name, gemm_args, "epoch", self.param_epoch)
            if gemm_args is None:                     # consumer cannot run the fused kernel: gather now
                self.gather(u, full)
                return None
            d = _CommDesc()
            d.mode, d.nranks, d.rank, d.rows_per_rank, d.comm_ctas, d.epoch = 3, self.n, self.rank, u.per // u.row, 0, self.param_epoch
            for j, p in enumerate(self._peer_shards(u)):
                d.peer_src[j] = p
            d.local_buf = key
            d.blk_flags = self.blk_flags.data_ptr()
            d.ready_flags = self.flags.ptr
            return d

        pending_gathers[key] = run
        u.lazy_pending = True
        u.lazy_key = key

    def cancel_deferred(self, u):
        if getattr(u, "lazy_pending", False):
            run = pending_gathers.get(u.lazy_key)
            if run is not None:
                run(None)

    # ------------------------------------------------------------------ gradient slots
    def _add_slot(self):
        if len(self.slots) >= MAX_SLOTS:
            raise RuntimeError("sharded data parallelism: more than 8 units have gradients in flight at once")
        buf = SymmetricBuffer(self.slot_bytes, None, self.rank, None, pg=self.pg)
        self.slots.append({"buf": buf, "uses": 0, "busy": False, "event": None})
        return len(self.slots) - 1

    def acquire_slot(self, u):
        """fp32 [numel] gradient buffer for ``u`` in peer-visible memory (zeroed), once every peer has read the
        slot's previous contents.  Every rank runs the same program, so slot choices agree without talking."""
        s = next((i for i, sl in enumerate(self.slots) if not sl["busy"]), None)
        if s is None:
            s = self._add_slot()
        sl = self.slots[s]
        sl["uses"] += 1
        sl["busy"] = True
        e = sl["uses"]
        _dbg(self, "acquire", u.name, "slot", s, "use", e)
        stream = N.stream_ptr(self.dev)
        if e > 1:
            # peers: epoch flags in peer memory; this rank's own reduce kernel (side stream): a CUDA event, so
            # no kernel on this GPU ever spins on another kernel of this GPU
            N.call("smpk_flags_wait", stream, self.flags.ptr + 4 * MAX_RANKS * (2 + 2 * s), self.n, self.rank, e - 1)
            N.count_launch()
            if sl["event"] is not None:
                torch.cuda.current_stream(self.dev).wait_event(sl["event"])
        buf = sl["buf"].view(torch.float32, u.numel)
        buf.zero_()
        u.slot = (s, e)
        return buf

    def reduce(self, u, adam=None):
        """Signal this rank's gradient slot complete, then (side stream) sum every rank's slot over this rank's
        stripes into ``u.grad_shard`` - or straight through Adam when ``adam`` carries the hyper-parameters."""
        s, e = u.slot
        _dbg(self, "reduce", u.name, "slot", s, "use", e, "adam" if adam else "")
        cur = torch.cuda.current_stream(self.dev)
        peers = _ptr_array([self._flag_ptr(j, 1 + 2 * s) for j in range(self.n)])
        N.call("smpk_flags_signal", cur.cuda_stream, ctypes.cast(peers, ctypes.c_void_p), self.n, self.rank, e)
        self.reduce_stream.wait_stream(cur)
        pg = _ptr_array([self.slots[s]["buf"].peer_ptrs[j] for j in range(self.n)])
        pc = _ptr_array([self._flag_ptr(j, 2 + 2 * s) for j in range(self.n)])
        ready = self.flags.ptr + 4 * MAX_RANKS * (1 + 2 * s)
        rs = self.reduce_stream.cuda_stream
        common = (rs, ctypes.cast(pg, ctypes.c_void_p), u.grad_shard.data_ptr())
        tail = (u.piece, u.ppr, ready, ctypes.cast(pc, ctypes.c_void_p), self.counter[s:s + 1].data_ptr(), e, self.n,
                self.rank)
        if adam is None:
            N.call("smpk_sdp_reduce_adam", *common, None, None, None, None, 0, *tail, 1.0, 1.0, 0,
                   0.0, 0.0, 0.0, 0.0, 0.0, 1, 0, 0, 1.0, 0)
        else:
            st = adam["state"]
            N.call("smpk_sdp_reduce_adam", *common, u.master.data.data_ptr(), st["exp_avg"].data_ptr(),
                   st["exp_avg_sq"].data_ptr(), u.shard.data_ptr(), N.dtype_code(u.shard.dtype), *tail,
                   float(adam["scale"]), 1.0, 1, float(adam["lr"]), float(adam["beta1"]), float(adam["beta2"]),
                   float(adam["eps"]), float(adam["wd"]), int(adam["step"]), int(adam["adamw"]),
                   int(adam["bias_correction"]), 1.0, 0)
        N.count_launch(2)
        ev = torch.cuda.Event()
        ev.record(self.reduce_stream)
        self.reduce_events.append(ev)
        self.slots[s]["event"] = ev
        self.slots[s]["busy"] = False

    def wait_reduces(self):
        cur = torch.cuda.current_stream(self.dev)
        for ev in self.reduce_events:
            cur.wait_event(ev)
        self.reduce_events = []
