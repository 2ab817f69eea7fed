"""Activation offloading to pinned host memory on dedicated copy streams.

Reference: smdistributed/modelparallel/torch/offload.py:76-327 (``TensorOffloader``: high-priority D2H
and H2D streams :85-87, pinned allocation :229-235, ``offload`` :211-249, ``load`` :261-281, bounded
look-ahead ``activation_loading_horizon``).  The saved inputs of every activation checkpoint are copied
out with ``cudaMemcpyAsync`` on a side stream while the forward continues; during backward the
checkpoints are needed in reverse order, so each ``load`` also prefetches the next few (at most
``activation_loading_horizon`` resident at once).

The reference's tensor-parallel "only tp_rank 0 offloads, then broadcast" optimisation
(``_shard_offloaded_activations``, offload.py:152-185) exists because its activations are replicated
across tp ranks; in the token-parallel layout used here every rank's checkpointed activations are
distinct shards already, so each rank offloads exactly 1/t of the reference's volume with no broadcast.
"""

from __future__ import annotations

import torch


class _Handle:
    __slots__ = ("idx", "cpu", "shape", "dtype", "device", "d2h_done", "gpu", "h2d_done", "requires_grad")

    def __init__(self, idx):
        self.idx = idx
        self.cpu = None
        self.gpu = None
        self.d2h_done = None
        self.h2d_done = None


class TensorOffloader:
    def __init__(self, horizon: int = 4):
        self.horizon = max(0, int(horizon))
        self._handles: dict[int, _Handle] = {}
        self._next = 0
        self._pinned_pool: dict = {}
        self.d2h_stream = None
        self.h2d_stream = None
        self.bytes_offloaded = 0
        self.peak_host_bytes = 0
        self._host_bytes = 0

    def _streams(self, device):
        if self.d2h_stream is None:
            lo, hi = torch.cuda.Stream.priority_range()
            self.d2h_stream = torch.cuda.Stream(device=device, priority=hi)
            self.h2d_stream = torch.cuda.Stream(device=device, priority=hi)
        return self.d2h_stream, self.h2d_stream

    def _pinned(self, shape, dtype):
        key = (tuple(shape), dtype)
        pool = self._pinned_pool.setdefault(key, [])
        if pool:
            return pool.pop()
        t = torch.empty(shape, dtype=dtype, pin_memory=True)
        self._host_bytes += t.numel() * t.element_size()
        self.peak_host_bytes = max(self.peak_host_bytes, self._host_bytes)
        return t

    # ------------------------------------------------------------------ forward side
    def offload(self, t: torch.Tensor) -> _Handle:
        d2h, _ = self._streams(t.device)
        h = _Handle(self._next)
        self._next += 1
        h.shape, h.dtype, h.device = t.shape, t.dtype, t.device
        src = t.detach()
        if not src.is_contiguous():
            src = src.contiguous()
        h.cpu = self._pinned(src.shape, src.dtype)
        d2h.wait_stream(torch.cuda.current_stream(t.device))
        d2h.wait_stream(self.h2d_stream)   # a recycled pinned buffer may still be feeding an earlier H2D copy
        with torch.cuda.stream(d2h):
            h.cpu.copy_(src, non_blocking=True)
            h.d2h_done = torch.cuda.Event()
            h.d2h_done.record(d2h)
        src.record_stream(d2h)          # keep the source alive until the copy has run
        self._handles[h.idx] = h
        self.bytes_offloaded += src.numel() * src.element_size()
        return h

    # ------------------------------------------------------------------ backward side
    def _start_load(self, h: _Handle):
        if h.gpu is not None:
            return
        _, h2d = self._streams(h.device)
        with torch.cuda.stream(h2d):
            h2d.wait_event(h.d2h_done)
            h.gpu = torch.empty(h.shape, dtype=h.dtype, device=h.device)
            h.gpu.copy_(h.cpu, non_blocking=True)
            h.h2d_done = torch.cuda.Event()
            h.h2d_done.record(h2d)

    def load(self, h: _Handle) -> torch.Tensor:
        self._start_load(h)
        # look ahead: checkpoints are consumed in reverse creation order
        resident = sum(1 for x in self._handles.values() if x.gpu is not None)
        k = h.idx - 1
        while resident < self.horizon and k >= 0:
            nxt = self._handles.get(k)
            if nxt is not None and nxt.gpu is None:
                self._start_load(nxt)
                resident += 1
            k -= 1
        cur = torch.cuda.current_stream(h.device)
        cur.wait_event(h.h2d_done)
        out = h.gpu
        out.record_stream(cur)
        self._release(h)
        return out

    def _release(self, h: _Handle):
        self._handles.pop(h.idx, None)
        if h.cpu is not None:
            self._pinned_pool.setdefault((tuple(h.cpu.shape), h.cpu.dtype), []).append(h.cpu)
        h.cpu = None
        h.gpu = None

    def reset(self):
        """End of a minibatch: drop handles whose backward never ran (e.g. evaluation under grad mode)."""
        for h in list(self._handles.values()):
            self._release(h)
        self._next = 0

    def saved_tensors(self):
        return len(self._handles)


def get_offloader() -> TensorOffloader:
    from .state_mod import state

    key = state.microbatch if state.core.pp_size() > 1 else 0
    off = state.offloaders.get(key)
    if off is None:
        off = TensorOffloader(state.cfg.activation_loading_horizon)
        state.offloaders[key] = off
    return off
