"""Multi-tensor apply front-end (``amp_C`` work-alike) over csrc/kernels/sm100/multi_tensor.cu.

The call shape follows Apex so optimizer code ported from the reference reads the same:
``multi_tensor_applier(op, noop_flag, tensor_lists, *args)`` with ``op`` one of the functions below
(reference smdistributed/modelparallel/torch/apex/multi_tensor_apply/multi_tensor_apply.py and the call
sites reference torch/fp16/fp16.py:132,159,555, torch/fp16/fp16util.py:205,
torch/apex/optimizers/fused_adam.py:148-170).  CUDA tensors run the native kernels; CPU tensors take
the per-tensor PyTorch restatement (used by the gloo test tier).
"""

from __future__ import annotations

import ctypes
import math

import torch

from . import _native as N
from .workspace import get_workspace

DEFAULT_CHUNK = 2048 * 32


def _pack(tensor_lists):
    """ctypes views for smpk_mt_*: (n, lists**, sizes*, dtypes*) plus keep-alive references."""
    depth = len(tensor_lists)
    n = len(tensor_lists[0])
    for lst in tensor_lists:
        assert len(lst) == n, "multi_tensor: lists must have equal length"
    arrs = []
    for lst in tensor_lists:
        arrs.append((ctypes.c_void_p * max(n, 1))(*[t.data_ptr() for t in lst]))
    lists = (ctypes.POINTER(ctypes.c_void_p) * depth)(*[ctypes.cast(a, ctypes.POINTER(ctypes.c_void_p)) for a in arrs])
    sizes = (ctypes.c_longlong * max(n, 1))(*[t.numel() for t in tensor_lists[0]])
    dtypes = (ctypes.c_int * depth)(*[N.dtype_code(lst[0].dtype) if n else 0 for lst in tensor_lists])
    for lst in tensor_lists:
        for t, ref in zip(lst, tensor_lists[0]):
            assert t.numel() == ref.numel() and t.is_contiguous() and t.dtype == lst[0].dtype
    keep = (arrs, lists, sizes, dtypes)
    as_ptr = lambda a: ctypes.cast(a, ctypes.c_void_p)
    return n, as_ptr(lists), as_ptr(sizes), as_ptr(dtypes), keep


def _is_cuda(tensor_lists):
    return len(tensor_lists[0]) > 0 and tensor_lists[0][0].is_cuda


def _noop_ptr(noop_flag):
    return None if noop_flag is None else noop_flag.data_ptr()


def _skip(noop_flag):
    return noop_flag is not None and not noop_flag.is_cuda and int(noop_flag.item()) != 0


def multi_tensor_applier(op, noop_flag, tensor_lists, *args, chunk_size: int = DEFAULT_CHUNK):
    return op(chunk_size, noop_flag, tensor_lists, *args)


def _adaptive_chunk(chunk, tensor_lists):
    # huge flat buffers: grow the chunk so one launch covers them (kernel arg block holds 640 chunks)
    biggest = max((t.numel() for t in tensor_lists[0]), default=0)
    while biggest // chunk > 4096 and chunk < (1 << 26):
        chunk *= 2
    return chunk


def multi_tensor_scale(chunk, noop_flag, tensor_lists, scale, scale_dev=None):
    """out = in * scale for every pair; noop_flag is set to 1 when an input is inf/nan."""
    ins, outs = tensor_lists
    if not ins:
        return
    if _is_cuda(tensor_lists):
        n, lists, sizes, dtypes, keep = _pack(tensor_lists)
        N.count_launch()
        N.call("smpk_mt_scale", N.stream_ptr(ins[0].device), n, lists, sizes, dtypes, _adaptive_chunk(chunk, tensor_lists),
               float(scale), N.ptr(scale_dev), _noop_ptr(noop_flag), None)
        return
    s = float(scale) * (float(scale_dev.item()) if scale_dev is not None else 1.0)
    for i, o in zip(ins, outs):
        if noop_flag is not None and not torch.isfinite(i).all():
            noop_flag.fill_(1)
        o.copy_((i.float() * s).to(o.dtype))


def multi_tensor_axpby(chunk, noop_flag, tensor_lists, a, b, arg_to_check=-1):
    xs, ys, outs = tensor_lists
    if not xs:
        return
    if _is_cuda(tensor_lists):
        n, lists, sizes, dtypes, keep = _pack(tensor_lists)
        N.count_launch()
        N.call("smpk_mt_axpby", N.stream_ptr(xs[0].device), n, lists, sizes, dtypes, _adaptive_chunk(chunk, tensor_lists),
               float(a), float(b), int(arg_to_check), _noop_ptr(noop_flag), None)
        return
    for x, y, o in zip(xs, ys, outs):
        bad = (arg_to_check != 1 and not torch.isfinite(x).all()) or (arg_to_check != 0 and not torch.isfinite(y).all())
        if bad and noop_flag is not None:
            noop_flag.fill_(1)
        o.copy_((a * x.float() + b * y.float()).to(o.dtype))


def multi_tensor_l2norm(chunk, noop_flag, tensor_lists, per_tensor=False):
    """Returns (global L2 norm [1], per-tensor norms [n] or empty)."""
    (xs,) = tensor_lists
    dev = xs[0].device if xs else torch.device("cpu")
    if xs and xs[0].is_cuda:
        n, lists, sizes, dtypes, keep = _pack(tensor_lists)
        out = torch.empty(1, dtype=torch.float32, device=dev)
        pt = torch.empty(n, dtype=torch.float32, device=dev) if per_tensor else None
        scratch = get_workspace(dev, "l2norm", N.query("smpk_mt_l2norm_scratch_floats") * 4)
        N.count_launch(2)
        N.call("smpk_mt_l2norm", N.stream_ptr(dev), n, lists, sizes, dtypes, _adaptive_chunk(chunk, tensor_lists),
               out.data_ptr(), N.ptr(pt), scratch.data_ptr(), _noop_ptr(noop_flag), None)
        return out, (pt if per_tensor else torch.empty(0, device=dev))
    norms = [x.float().norm() for x in xs]
    tot = torch.sqrt(sum((n_ * n_ for n_ in norms), torch.zeros((), device=dev))).reshape(1)
    if noop_flag is not None and not torch.isfinite(tot).all():
        noop_flag.fill_(1)
    return tot, (torch.stack(norms) if per_tensor and norms else torch.empty(0, device=dev))


def multi_tensor_adam(chunk, noop_flag, tensor_lists, lr, beta1, beta2, eps, step, mode, bias_correction,
                      weight_decay, grad_scale=1.0, grad_scale_dev=None):
    """lists: [g, p, m, v] or [g, p, m, v, p_lowprec]; mode 0 = L2 (Adam), 1 = decoupled (AdamW)."""
    if not tensor_lists[0]:
        return
    if _is_cuda(tensor_lists):
        n, lists, sizes, dtypes, keep = _pack(tensor_lists)
        N.count_launch()
        N.call("smpk_mt_adam", N.stream_ptr(tensor_lists[0][0].device), n, lists, sizes, dtypes,
               _adaptive_chunk(chunk, tensor_lists), len(tensor_lists), float(lr), float(beta1), float(beta2),
               float(eps), int(step), int(mode), int(bool(bias_correction)), float(weight_decay), float(grad_scale),
               N.ptr(grad_scale_dev), _noop_ptr(noop_flag))
        return
    if _skip(noop_flag):
        return
    gs = float(grad_scale) * (float(grad_scale_dev.item()) if grad_scale_dev is not None else 1.0)
    bc1 = 1 - beta1 ** step if bias_correction else 1.0
    bc2 = 1 - beta2 ** step if bias_correction else 1.0
    for i in range(len(tensor_lists[0])):
        g = tensor_lists[0][i].float() * gs
        p, m, v = tensor_lists[1][i], tensor_lists[2][i], tensor_lists[3][i]
        pf = p.float()
        if mode == 0:
            g = g + weight_decay * pf
        m.copy_(beta1 * m + (1 - beta1) * g)
        v.copy_(beta2 * v + (1 - beta2) * g * g)
        upd = (m / bc1) / ((v / bc2).sqrt() + eps)
        if mode == 1:
            upd = upd + weight_decay * pf
        p.copy_((pf - lr * upd).to(p.dtype))
        if len(tensor_lists) == 5:
            tensor_lists[4][i].copy_(p.to(tensor_lists[4][i].dtype))


def multi_tensor_sgd(chunk, noop_flag, tensor_lists, weight_decay, momentum, dampening, lr, nesterov, first_run,
                     wd_after_momentum, scale=1.0):
    """lists: [g, p, momentum] or [g, p, momentum, p_lowprec]."""
    if not tensor_lists[0]:
        return
    if _is_cuda(tensor_lists):
        n, lists, sizes, dtypes, keep = _pack(tensor_lists)
        N.count_launch()
        N.call("smpk_mt_sgd", N.stream_ptr(tensor_lists[0][0].device), n, lists, sizes, dtypes,
               _adaptive_chunk(chunk, tensor_lists), len(tensor_lists), float(lr), float(momentum), float(dampening),
               float(weight_decay), int(bool(nesterov)), int(bool(first_run)), int(bool(wd_after_momentum)),
               float(scale), _noop_ptr(noop_flag))
        return
    if _skip(noop_flag):
        return
    for i in range(len(tensor_lists[0])):
        g = tensor_lists[0][i].float() * scale
        p, mom = tensor_lists[1][i], tensor_lists[2][i]
        pf = p.float()
        if weight_decay and not wd_after_momentum:
            g = g + weight_decay * pf
        if momentum:
            mom.copy_(g if first_run else mom * momentum + (1 - dampening) * g)
            g = g + momentum * mom if nesterov else mom.float()
        if weight_decay and wd_after_momentum:
            g = g + weight_decay * pf
        p.copy_((pf - lr * g).to(p.dtype))
        if len(tensor_lists) == 4:
            tensor_lists[3][i].copy_(p.to(tensor_lists[3][i].dtype))


def multi_tensor_adagrad(chunk, noop_flag, tensor_lists, lr, eps, mode, weight_decay, grad_scale=1.0):
    """lists: [g, p, h] (+ p_lowprec); mode 0 = L2, 1 = decoupled weight decay."""
    if not tensor_lists[0]:
        return
    if _is_cuda(tensor_lists):
        n, lists, sizes, dtypes, keep = _pack(tensor_lists)
        N.count_launch()
        N.call("smpk_mt_adagrad", N.stream_ptr(tensor_lists[0][0].device), n, lists, sizes, dtypes,
               _adaptive_chunk(chunk, tensor_lists), len(tensor_lists), float(lr), float(eps), float(weight_decay),
               int(mode), float(grad_scale), _noop_ptr(noop_flag))
        return
    if _skip(noop_flag):
        return
    for i in range(len(tensor_lists[0])):
        g = tensor_lists[0][i].float() * grad_scale
        p, h = tensor_lists[1][i], tensor_lists[2][i]
        pf = p.float()
        if mode == 0:
            g = g + weight_decay * pf
        h.add_(g * g)
        upd = g / (h.sqrt() + eps)
        if mode == 1:
            upd = upd + weight_decay * pf
        p.copy_((pf - lr * upd).to(p.dtype))
        if len(tensor_lists) == 4:
            tensor_lists[3][i].copy_(p.to(tensor_lists[3][i].dtype))


def multi_tensor_lamb(chunk, noop_flag, tensor_lists, lr, beta1, beta2, eps, step, bias_correction, weight_decay,
                      grad_averaging, mode, global_grad_norm, max_grad_norm, use_nvlamb=False, grad_scale=1.0):
    """LAMB over lists [g, p, m, v] (+ p_lowprec): stage 1 builds the Adam-style update, per-tensor norms
    give the trust ratio, stage 2 applies it.  ``global_grad_norm`` is a 1-element tensor (or None)."""
    if not tensor_lists[0]:
        return
    gl, pl, ml, vl = tensor_lists[:4]
    lowp = tensor_lists[4] if len(tensor_lists) == 5 else None
    updates = [torch.empty_like(p, dtype=torch.float32) for p in pl]
    if _is_cuda(tensor_lists):
        dev = pl[0].device
        n, lists, sizes, dtypes, keep = _pack([gl, pl, ml, vl, updates])
        ch = _adaptive_chunk(chunk, tensor_lists)
        N.count_launch()
        N.call("smpk_mt_lamb_stage1", N.stream_ptr(dev), n, lists, sizes, dtypes, ch, float(beta1), float(beta2),
               float(eps), int(step), int(bool(bias_correction)), float(weight_decay), int(bool(grad_averaging)),
               int(mode), N.ptr(global_grad_norm), float(max_grad_norm), float(grad_scale), _noop_ptr(noop_flag))
        _, p_norms = multi_tensor_l2norm(chunk, None, [pl], True)
        _, u_norms = multi_tensor_l2norm(chunk, None, [updates], True)
        second = [updates, pl] + ([lowp] if lowp is not None else [])
        n, lists, sizes, dtypes, keep2 = _pack(second)
        N.count_launch()
        N.call("smpk_mt_lamb_stage2", N.stream_ptr(dev), n, lists, sizes, dtypes, ch, len(second), float(lr),
               float(weight_decay), int(bool(use_nvlamb)), p_norms.data_ptr(), u_norms.data_ptr(),
               _noop_ptr(noop_flag))
        return
    if _skip(noop_flag):
        return
    clip = 1.0
    if global_grad_norm is not None and max_grad_norm > 0:
        gn = float(global_grad_norm.item()) * grad_scale
        clip = max_grad_norm / gn if gn > max_grad_norm else 1.0
    bc1 = 1 - beta1 ** step if bias_correction else 1.0
    bc2 = 1 - beta2 ** step if bias_correction else 1.0
    beta3 = 1 - beta1 if grad_averaging else 1.0
    for i in range(len(gl)):
        g = gl[i].float() * grad_scale * clip
        p, m, v = pl[i], ml[i], vl[i]
        pf = p.float()
        if mode == 0:
            g = g + weight_decay * pf
        m.copy_(beta1 * m + beta3 * g)
        v.copy_(beta2 * v + (1 - beta2) * g * g)
        upd = (m / bc1) / ((v / bc2).sqrt() + eps)
        if mode == 1:
            upd = upd + weight_decay * pf
        pn, un = pf.norm(), upd.norm()
        ratio = lr
        if (use_nvlamb or weight_decay != 0) and pn > 0 and un > 0:
            ratio = lr * float(pn / un)
        p.copy_((pf - ratio * upd).to(p.dtype))
        if lowp is not None:
            lowp[i].copy_(p.to(lowp[i].dtype))


def multi_tensor_novograd(chunk, noop_flag, tensor_lists, grad_norms, lr, beta1, beta2, eps, step, bias_correction,
                          weight_decay, grad_averaging, moment_mode, norm_type=2, grad_scale=1.0):
    """NovoGrad over [g, p, m] (+ p_lowprec); ``grad_norms`` (fp32 [n]) holds the per-tensor second moment,
    blended in place here: v = beta2 * v + (1 - beta2) * ||g||  (v = ||g|| on the first step)."""
    if not tensor_lists[0]:
        return
    assert norm_type == 2, "only the L2 variant is implemented"
    gl = tensor_lists[0]
    _, cur = multi_tensor_l2norm(chunk, None, [gl], True)
    cur = cur * grad_scale
    if step == 1:
        grad_norms.copy_(cur)
    else:
        # blend in squared space then take the root, as the Apex kernel does for the L2 norm type
        grad_norms.copy_(torch.sqrt(beta2 * grad_norms * grad_norms + (1 - beta2) * cur * cur))
    if _is_cuda(tensor_lists):
        n, lists, sizes, dtypes, keep = _pack(tensor_lists)
        N.count_launch()
        N.call("smpk_mt_novograd", N.stream_ptr(gl[0].device), n, lists, sizes, dtypes,
               _adaptive_chunk(chunk, tensor_lists), len(tensor_lists), float(lr), float(beta1), float(beta2),
               float(eps), int(step), int(bool(bias_correction)), float(weight_decay), int(bool(grad_averaging)),
               int(moment_mode), grad_norms.data_ptr(), float(grad_scale), _noop_ptr(noop_flag))
        return
    if _skip(noop_flag):
        return
    bc1 = 1 - beta1 ** step if bias_correction else 1.0
    bc2 = math.sqrt(1 - beta2 ** step) if bias_correction else 1.0
    beta3 = 1 - beta1 if grad_averaging else 1.0
    for i in range(len(gl)):
        g = gl[i].float() * grad_scale
        p, m = tensor_lists[1][i], tensor_lists[2][i]
        pf = p.float()
        denom = float(grad_norms[i]) / bc2 + eps
        if moment_mode == 0:
            g = g / denom + weight_decay * pf
            m.copy_(beta1 * m + beta3 * g)
            pf = pf - lr * (m / bc1)
        else:
            m.copy_(beta1 * m + beta3 * g)
            pf = pf - lr * ((m / bc1) / denom + weight_decay * pf)
        p.copy_(pf.to(p.dtype))
        if len(tensor_lists) == 4:
            tensor_lists[3][i].copy_(p.to(tensor_lists[3][i].dtype))
