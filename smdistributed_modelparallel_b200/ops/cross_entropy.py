"""Fused (vocab-parallel) softmax cross-entropy on the sm_100a kernels.

Forward computes per-token statistics over the local vocabulary shard, reduces them across the
tensor-parallel group when there is one, then overwrites the logits with the gradient in place - the
[tokens, vocab/tp] tensor is held once.  Replaces the reference's eager-op ``_VocabParallelCrossEntropy``
(reference smdistributed/modelparallel/torch/nn/cross_entropy.py:28-98; its three all-reduces MAX /
SUM / SUM at :34,:60,:66 are the same three here).
"""

from __future__ import annotations

import torch
import torch.distributed as dist

from . import _native as N


def _stats_ref(logits, targets, vocab_start):
    lf = logits.float()
    m = lf.max(dim=-1).values
    s = torch.exp(lf - m[:, None]).sum(-1)
    t = targets - vocab_start
    inside = (t >= 0) & (t < logits.shape[-1])
    tl = lf.gather(1, t.clamp(0, logits.shape[-1] - 1)[:, None])[:, 0]
    return m, s, torch.where(inside, tl, torch.zeros_like(tl))


class _VocabParallelCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets, vocab_start, group, ignore_index, inplace_grad):
        V = logits.shape[-1]
        l2 = logits.reshape(-1, V)
        if not l2.is_contiguous():
            l2 = l2.contiguous()
        t1 = targets.reshape(-1).to(torch.int64).contiguous()
        rows = l2.shape[0]
        native = l2.is_cuda and l2.dtype in N._DTYPE_CODE
        if native:
            st = torch.empty(3, rows, dtype=torch.float32, device=l2.device)
            N.count_launch()
            N.call("smpk_ce_stats", N.stream_ptr(l2.device), l2.data_ptr(), N.dtype_code(l2.dtype), t1.data_ptr(),
                   int(vocab_start), rows, V, l2.stride(0), st[0].data_ptr(), st[1].data_ptr(), st[2].data_ptr())
            m, s, tl = st[0], st[1], st[2]
        else:
            m, s, tl = _stats_ref(l2, t1, vocab_start)
        if group is not None and dist.get_world_size(group) > 1:
            gm = m.clone()
            dist.all_reduce(gm, op=dist.ReduceOp.MAX, group=group)
            pack = torch.stack([s * torch.exp(m - gm), tl])
            dist.all_reduce(pack, op=dist.ReduceOp.SUM, group=group)
            m, s, tl = gm, pack[0].contiguous(), pack[1].contiguous()
        need_grad = ctx.needs_input_grad[0]
        if native:
            loss = torch.empty(rows, dtype=torch.float32, device=l2.device)
            grad_buf = l2 if inplace_grad else (l2.clone() if need_grad else l2)
            N.count_launch()
            N.call("smpk_ce_finish", N.stream_ptr(l2.device), grad_buf.data_ptr(), N.dtype_code(l2.dtype),
                   t1.data_ptr(), int(vocab_start), rows, V, grad_buf.stride(0), m.data_ptr(), s.data_ptr(),
                   tl.data_ptr(), 1.0, int(ignore_index), loss.data_ptr(), int(need_grad))
            grad = grad_buf if need_grad else None
        else:
            loss = torch.log(s) + m - tl
            ign = t1 == ignore_index
            loss = loss.masked_fill(ign, 0.0)
            grad = None
            if need_grad:
                p = torch.exp(l2.float() - m[:, None]) / s[:, None]
                t = t1 - vocab_start
                inside = (t >= 0) & (t < V)
                idx = t.clamp(0, V - 1)
                p[torch.arange(rows, device=p.device)[inside], idx[inside]] -= 1.0
                p = p.masked_fill(ign[:, None], 0.0)
                grad = p.to(l2.dtype)
        ctx.save_for_backward(grad)
        ctx.shape = logits.shape
        return loss.reshape(targets.shape)

    @staticmethod
    def backward(ctx, dloss):
        (grad,) = ctx.saved_tensors
        g = grad * dloss.reshape(-1, 1).to(grad.dtype)
        return g.reshape(ctx.shape), None, None, None, None, None


def vocab_parallel_cross_entropy(logits: torch.Tensor, targets: torch.Tensor, vocab_start: int = 0, group=None,
                                 ignore_index: int = -100, inplace_grad: bool = False) -> torch.Tensor:
    """Per-token loss (fp32, shape of ``targets``) for logits sharded on the vocabulary dimension.

    ``inplace_grad=True`` lets the kernel overwrite ``logits`` with the gradient (the caller must not
    use ``logits`` afterwards) - what the LM head does.
    """
    return _VocabParallelCE.apply(logits, targets, vocab_start, group, ignore_index, inplace_grad)


def cross_entropy(logits, targets, ignore_index: int = -100, reduction: str = "mean", inplace_grad: bool = False):
    loss = vocab_parallel_cross_entropy(logits, targets, 0, None, ignore_index, inplace_grad)
    if reduction == "none":
        return loss
    if reduction == "sum":
        return loss.sum()
    valid = (targets != ignore_index).sum().clamp(min=1)
    return loss.sum() / valid
