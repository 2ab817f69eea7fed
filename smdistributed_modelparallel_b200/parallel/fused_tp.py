"""Fused tensor-parallel GEMM + collective over NVLink peer memory.

``ag_gemm``  : token all-gather fused into a column-parallel GEMM (copy CTAs pull the peers' shards while
               the tcgen05 CTAs already multiply the local shard; remote row blocks are consumed as they land).
``gemm_rs``  : row-parallel GEMM whose epilogue stores each tile straight into the owner rank's staging
               slot (remote tiles first), followed by a small reduce kernel on the owner.
``all_gather``: the copy kernel alone (wgrad needs the gathered activations again in backward).

Buffers are *symmetric*: every rank of the tensor-parallel group allocates the same layout with
``smpc_symm_alloc`` (cudaMalloc + CUDA IPC handle, csrc/symm/symm.cu), exchanges handles once through the
control-plane bus and maps its peers.  Ordering uses uint32 epoch flags in that memory (written with
``st.release.sys``, polled with ``ld.acquire.sys``); each logical buffer is double-buffered and a producer
waits for its consumers' "done" flags of two epochs ago before overwriting.

These replace the NCCL collectives the reference issues around its cuBLAS GEMMs (reference
smdistributed/modelparallel/torch/nn/transformer.py:1436,1474,1522,1141; torch/nn/utils.py:540-571,701-821;
SURVEY 2.5 C1-C6), which run on the compute stream and are never overlapped with the GEMMs.
"""

from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from ..backend import native_core
from ..backend.collectives import CommGroup
from ..ops import _native as N
from ..ops.gelu import gelu_backward
from ..ops.gemm import _EPI_BY_NAME, colsum, gemm

MAX_RANKS = 8
BLOCK_M = 128
_NBUF = 2                      # double buffering of every logical buffer
_MAX_BLOCKS = 4096             # row blocks tracked per gathered buffer (512k tokens)


class _CudaMem:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class _CommDesc(ctypes.Structure):
    _fields_ = [("mode", ctypes.c_int), ("nranks", ctypes.c_int), ("rank", ctypes.c_int),
                ("rows_per_rank", ctypes.c_int), ("comm_ctas", ctypes.c_int), ("epoch", ctypes.c_uint),
                ("peer_src", ctypes.c_void_p * MAX_RANKS), ("local_buf", ctypes.c_void_p),
                ("blk_flags", ctypes.c_void_p), ("ready_flags", ctypes.c_void_p),
                ("peer_done", ctypes.c_void_p * MAX_RANKS), ("counters", ctypes.c_void_p),
                ("peer_stage", ctypes.c_void_p * MAX_RANKS), ("peer_arrived", ctypes.c_void_p * MAX_RANKS)]


def _ptr_array(ptrs):
    arr = (ctypes.c_void_p * MAX_RANKS)()
    for i, p in enumerate(ptrs):
        arr[i] = p
    return arr


class SymmetricBuffer:
    """One allocation of ``nbytes`` on every rank of the group, with every peer's copy mapped locally."""

    def __init__(self, nbytes, ranks, me_index, comm, pg=None):
        lib = native_core.load()
        self.nbytes = int(nbytes)
        ptr = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        if lib.smpc_symm_alloc(self.nbytes, ctypes.byref(ptr), handle) != 0:
            raise RuntimeError(f"symmetric allocation failed: {lib.smpc_symm_last_error().decode()}")
        self.ptr = ptr.value
        if pg is not None:                  # exchange over a torch.distributed group (sharded data parallelism)
            handles = [None] * dist.get_world_size(pg)
            dist.all_gather_object(handles, bytes(handle.raw), group=pg)
        else:
            handles = comm.allgather(bytes(handle.raw), CommGroup.TP_GROUP)
        self.peer_ptrs = []
        for j, h in enumerate(handles):
            if j == me_index:
                self.peer_ptrs.append(self.ptr)
                continue
            pp = ctypes.c_void_p()
            if lib.smpc_symm_open(h, ctypes.byref(pp)) != 0:
                raise RuntimeError(f"mapping peer memory failed: {lib.smpc_symm_last_error().decode()}")
            self.peer_ptrs.append(pp.value)
        self.tensor = torch.as_tensor(_CudaMem(self.ptr, self.nbytes), device="cuda")

    def view(self, dtype, numel, byte_offset=0):
        es = torch.empty((), dtype=dtype).element_size()
        return self.tensor[byte_offset:byte_offset + numel * es].view(dtype)


class FusedTP:
    """Per-process manager of the symmetric buffers and epochs of one tensor-parallel group."""

    def __init__(self, state):
        self.state = state
        core = state.core
        self.group = state.tp_process_group
        self.n = core.tp_size()
        self.rank = core.tp_rank()
        self.ranks = core.get_tp_group()
        self.dev = state.device
        self.comm_ctas = 20
        # flag words (uint32), one symmetric block: per buffer parity b: ready[n] done[n] arrived[n] reduced[n]
        self._flag_words = _NBUF * 4 * MAX_RANKS
        self.flags = SymmetricBuffer(self._flag_words * 4, self.ranks, self.rank, state.comm)
        self.gather = None          # SymmetricBuffer holding _NBUF gathered-activation buffers
        self.gather_bytes = 0
        self.stage = None           # SymmetricBuffer holding _NBUF staging buffers
        self.stage_bytes = 0
        self.blk_flags = torch.zeros(_NBUF, _MAX_BLOCKS, dtype=torch.int32, device=self.dev)
        self.counters = torch.zeros(_NBUF, 2, MAX_RANKS, dtype=torch.int32, device=self.dev)
        self.red_counter = torch.zeros(_NBUF, dtype=torch.int32, device=self.dev)
        self.g_epoch = [0] * _NBUF
        self.s_epoch = [0] * _NBUF
        self.g_next = 0
        self.s_next = 0
        torch.cuda.synchronize()
        state.comm.barrier(CommGroup.TP_GROUP)

    # ---- flag addressing (uint32 index inside the flags block)
    def _fidx(self, b, kind):
        return (b * 4 + kind) * MAX_RANKS

    def _flag_ptr(self, peer, b, kind):
        return self.flags.peer_ptrs[peer] + 4 * self._fidx(b, kind)

    def _ensure(self, which, nbytes):
        """(Re)allocate the gathered / staging buffers collectively when a larger size is needed."""
        cur = self.gather_bytes if which == "gather" else self.stage_bytes
        if nbytes <= cur:
            return
        nbytes = (int(nbytes) + (1 << 21) - 1) >> 21 << 21
        torch.cuda.synchronize()
        self.state.comm.barrier(CommGroup.TP_GROUP)
        buf = SymmetricBuffer(_NBUF * nbytes, self.ranks, self.rank, self.state.comm)
        if which == "gather":
            self.gather, self.gather_bytes = buf, nbytes
        else:
            self.stage, self.stage_bytes = buf, nbytes
        torch.cuda.synchronize()
        self.state.comm.barrier(CommGroup.TP_GROUP)

    # ------------------------------------------------------------------ all-gather side
    def _publish_shard(self, x_local):
        """Put this rank's shard into its slot of the next gathered buffer and tell the peers."""
        T_l, K = x_local.shape
        slot = T_l * K * 2
        self._ensure("gather", slot * self.n)
        b = self.g_next
        self.g_next = (b + 1) % _NBUF
        prev, self.g_epoch[b] = self.g_epoch[b], self.g_epoch[b] + 1
        epoch = self.g_epoch[b]
        stream = N.stream_ptr(self.dev)
        if prev > 0:        # peers must have finished reading my slot of this buffer's previous use
            N.call("smpk_flags_wait", stream, self.flags.ptr + 4 * self._fidx(b, 1), self.n, self.rank, prev)
        base = b * self.gather_bytes
        full = self.gather.view(x_local.dtype, self.n * T_l * K, base).view(self.n * T_l, K)
        full[self.rank * T_l:(self.rank + 1) * T_l].copy_(x_local)
        peers = _ptr_array([self._flag_ptr(j, b, 0) for j in range(self.n)])
        N.call("smpk_flags_signal", stream, ctypes.cast(peers, ctypes.c_void_p), self.n, self.rank, epoch)
        N.count_launch(2)
        return b, epoch, full, base, slot

    def _desc_ag(self, b, epoch, base, T_l):
        d = _CommDesc()
        d.mode, d.nranks, d.rank, d.rows_per_rank, d.comm_ctas, d.epoch = 1, self.n, self.rank, T_l, self.comm_ctas, epoch
        for j in range(self.n):
            d.peer_src[j] = self.gather.peer_ptrs[j] + base
            d.peer_done[j] = self._flag_ptr(j, b, 1)
        d.local_buf = self.gather.ptr + base
        d.blk_flags = self.blk_flags[b].data_ptr()
        d.ready_flags = self.flags.ptr + 4 * self._fidx(b, 0)
        d.counters = self.counters[b, 0].data_ptr()
        return d

    def all_gather(self, x_local):
        """[T/t, K] -> view of the gathered [T, K] inside the symmetric buffer (valid until that buffer is
        reused two gathers later - consume or clone it immediately)."""
        b, epoch, full, base, slot = self._publish_shard(x_local)
        peer_src = _ptr_array([self.gather.peer_ptrs[j] + base for j in range(self.n)])
        peer_done = _ptr_array([self._flag_ptr(j, b, 1) for j in range(self.n)])
        N.count_launch()
        N.call("smpk_p2p_allgather", N.stream_ptr(self.dev), ctypes.cast(peer_src, ctypes.c_void_p),
               self.gather.ptr + base, slot, self.flags.ptr + 4 * self._fidx(b, 0),
               ctypes.cast(peer_done, ctypes.c_void_p), self.counters[b, 0].data_ptr(), self.n, self.rank, epoch, 32)
        return full

    def ag_gemm(self, x_local, weight, *, b_mn_major=False, bias=None, epilogue="none", aux_out=None, out=None):
        """y = epilogue(allgather(x_local) @ W^T); returns (y, gathered view)."""
        T_l, K = x_local.shape
        Nn = weight.shape[1] if b_mn_major else weight.shape[0]
        b, epoch, full, base, _ = self._publish_shard(x_local)
        desc = self._desc_ag(b, epoch, base, T_l)
        if out is None:
            out = torch.empty(self.n * T_l, Nn, dtype=x_local.dtype, device=self.dev)
        epi = _EPI_BY_NAME[epilogue]
        N.count_launch()
        N.call("smpk_gemm_comm", N.stream_ptr(self.dev), N.dtype_code(x_local.dtype), N.dtype_code(out.dtype),
               self.n * T_l, Nn, K, full.data_ptr(), full.stride(0), 0, weight.data_ptr(), weight.stride(0),
               int(b_mn_major), out.data_ptr(), out.stride(0), N.ptr(bias),
               N.dtype_code(bias.dtype) if bias is not None else 0, None, 0, 0, 0.0, N.ptr(aux_out), 1.0, epi, 0,
               ctypes.byref(desc))
        return out, full

    # ------------------------------------------------------------------ reduce-scatter side
    def gemm_rs(self, x, weight, *, b_mn_major=False, bias=None, residual=None, out_residual=None, out_bias=None):
        """y_local[T/t, N] = reduce_scatter(x @ W^T (+ bias) (+ residual)) (+ out_bias) (+ out_residual)."""
        T, K = x.shape
        T_l = T // self.n
        Nn = weight.shape[1] if b_mn_major else weight.shape[0]
        stage_elems = self.n * T_l * Nn
        self._ensure("stage", stage_elems * 2)
        b = self.s_next
        self.s_next = (b + 1) % _NBUF
        prev, self.s_epoch[b] = self.s_epoch[b], self.s_epoch[b] + 1
        epoch = self.s_epoch[b]
        stream = N.stream_ptr(self.dev)
        if prev > 0:        # every peer must have reduced what I stored into this staging buffer last time
            N.call("smpk_flags_wait", stream, self.flags.ptr + 4 * self._fidx(b, 3), self.n, -1, prev)
            N.count_launch()
        base = b * self.stage_bytes
        d = _CommDesc()
        d.mode, d.nranks, d.rank, d.rows_per_rank, d.comm_ctas, d.epoch = 2, self.n, self.rank, T_l, 0, epoch
        for j in range(self.n):
            d.peer_stage[j] = self.stage.peer_ptrs[j] + base
            d.peer_arrived[j] = self._flag_ptr(j, b, 2)
        d.counters = self.counters[b, 1].data_ptr()
        N.count_launch()
        N.call("smpk_gemm_comm", stream, N.dtype_code(x.dtype), N.dtype_code(x.dtype), T, Nn, K, x.data_ptr(),
               x.stride(0), 0, weight.data_ptr(), weight.stride(0), int(b_mn_major), None, Nn, N.ptr(bias),
               N.dtype_code(bias.dtype) if bias is not None else 0, N.ptr(residual),
               residual.stride(0) if residual is not None else 0,
               N.dtype_code(residual.dtype) if residual is not None else 0, 1.0, None, 1.0, 0, 0, ctypes.byref(d))
        out = torch.empty(T_l, Nn, dtype=x.dtype, device=self.dev)
        peer_reduced = _ptr_array([self._flag_ptr(j, b, 3) for j in range(self.n)])
        N.count_launch()
        N.call("smpk_rs_reduce", stream, self.stage.ptr + base, self.flags.ptr + 4 * self._fidx(b, 2),
               ctypes.cast(peer_reduced, ctypes.c_void_p), self.red_counter[b:b + 1].data_ptr(), N.ptr(out_residual),
               N.dtype_code(out_residual.dtype) if out_residual is not None else 0, N.ptr(out_bias),
               N.dtype_code(out_bias.dtype) if out_bias is not None else 0, out.data_ptr(), N.dtype_code(out.dtype),
               T_l, Nn, self.n, self.rank, N.dtype_code(x.dtype), epoch)
        return out


_manager = None


def available(group=None) -> bool:
    """Fused kernels need CUDA, a tensor-parallel group of 2..8 ranks inside one node and mutual peer access."""
    from ..state_mod import state

    if not torch.cuda.is_available() or not getattr(state, "initialized", False):
        return False
    n = state.core.tp_size()
    if n < 2 or n > MAX_RANKS:
        return False
    return all(state.core.is_in_same_instance(r) for r in state.core.get_tp_group())


def manager() -> FusedTP:
    global _manager
    from ..state_mod import state

    if _manager is None or _manager.state is not state or _manager.group is not state.tp_process_group:
        _manager = FusedTP(state)
    return _manager


def usable(x_local_rows: int, *tensors) -> bool:
    """Shape / dtype gate for the fused kernels (128-row blocks per rank, 16-bit operands, aligned strides)."""
    if x_local_rows % BLOCK_M != 0:
        return False
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda or t.dtype not in (torch.bfloat16, torch.float16) or t.stride(-1) != 1:
            return False
        if t.dim() == 2 and (t.stride(0) % 8 or t.data_ptr() % 16):
            return False
    return True


def all_gather_tokens(x, group=None):
    return manager().all_gather(x.contiguous()).clone()


def reduce_scatter_tokens(x, group=None):
    from .tp_ops import scatter_sum_tokens
    return scatter_sum_tokens(x, group or manager().group)


def _wgrad(dy, x, weight):
    mg = getattr(weight, "main_grad", None)
    if mg is not None:
        gemm(dy, x, a_mn_major=True, b_mn_major=True, out=mg, residual=mg, beta=1.0)
        return None
    return gemm(dy, x, a_mn_major=True, b_mn_major=True)


def _bgrad(dy, bias):
    mg = getattr(bias, "main_grad", None)
    if mg is not None:
        colsum(dy, out=mg, accumulate=True)
        return None
    return colsum(dy, out_dtype=bias.dtype)


class FusedAGLinear(torch.autograd.Function):
    """all-gather(x_local) @ W^T (+bias, +GELU) in one kernel.  Optionally also returns the gathered input
    (a private copy) so a second column-parallel GEMM can share it; gradients arriving for that copy are
    folded into the reduce-scatter of this function's input gradient."""

    @staticmethod
    def forward(ctx, x_local, weight, bias, activation, want_gathered):
        m = manager()
        ctx.activation, ctx.want = activation, want_gathered
        pre = None
        if activation is None:
            y, full = m.ag_gemm(x_local, weight, bias=bias)
        else:
            need = any(ctx.needs_input_grad[:3])
            T = x_local.shape[0] * m.n
            pre = torch.empty(T, weight.shape[0], dtype=x_local.dtype, device=x_local.device) if need else None
            y, full = m.ag_gemm(x_local, weight, bias=bias, epilogue=f"gelu_{activation}", aux_out=pre)
        ctx.save_for_backward(x_local, weight, bias, pre)
        if want_gathered:
            return y, full.clone()
        return y

    @staticmethod
    def backward(ctx, dy, dxg=None):
        x_local, weight, bias, pre = ctx.saved_tensors
        m = manager()
        dy = dy if dy.is_contiguous() else dy.contiguous()
        if ctx.activation is not None:
            dy = gelu_backward(dy, pre, ctx.activation)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            res = dxg.contiguous() if (dxg is not None) else None
            dx = m.gemm_rs(dy, weight, b_mn_major=True, residual=res)
        if ctx.needs_input_grad[1]:
            X = m.all_gather(x_local)
            dw = _wgrad(dy, X, weight)
        if bias is not None and ctx.needs_input_grad[2]:
            db = _bgrad(dy, bias)
        return dx, dw, db, None, None


class FusedLinearRS(torch.autograd.Function):
    """reduce_scatter(x @ W^T (+bias) (+residual_full)) -> this rank's tokens, one GEMM kernel with peer stores
    plus the reduce kernel."""

    @staticmethod
    def forward(ctx, x, weight, bias, residual_full):
        m = manager()
        ctx.has_res = residual_full is not None
        y = m.gemm_rs(x, weight, bias=bias, residual=residual_full)
        ctx.save_for_backward(x, weight, bias)
        return y

    @staticmethod
    def backward(ctx, dy_local):
        x, weight, bias = ctx.saved_tensors
        m = manager()
        dy_local = dy_local if dy_local.is_contiguous() else dy_local.contiguous()
        dx = dw = db = dres = None
        dx_full, dY = m.ag_gemm(dy_local, weight, b_mn_major=True)      # dgrad on the gathered output grads
        if not ctx.needs_input_grad[0]:
            dx_full = None
        if ctx.needs_input_grad[1]:
            dw = _wgrad(dY, x, weight)
        if bias is not None and ctx.needs_input_grad[2]:
            db = _bgrad(dY, bias)
        if ctx.has_res and ctx.needs_input_grad[3]:
            dres = dY.clone()
        return dx_full, dw, db, dres
