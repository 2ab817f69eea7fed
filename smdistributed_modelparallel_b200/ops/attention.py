"""Attention building blocks on the sm_100a kernels: the flash-style fused attention (forward and backward,
``csrc/kernels/sm100/attention.cu``), QKV split with fused rotary embedding and head transposes.

The reference computes attention as ``baddbmm`` -> fused softmax -> dropout -> ``matmul`` and keeps the
``[B, heads, Sq, Sk]`` scores and probabilities for backward (reference smdistributed/modelparallel/torch/nn/
transformer.py:1617-1657,1769-1835), and rotary embeddings with eager ops (:114-163,1565-1615).  Here one kernel
per direction keeps the score tiles in tensor memory: nothing of size ``Sq x Sk`` is ever written to global memory,
the backward pass recomputes the probabilities from the saved log-sum-exp rows, and causal / byte / local-window
masks and dropout (counter-based, so a checkpoint recomputation sees the same mask) are applied in registers.
Rotary is fused into the QKV split.
"""

from __future__ import annotations

import math

import torch

from . import _native as N


# ----------------------------------------------------------------------------------------- rotary
_table_cache: dict = {}


def rotary_tables(seq_len: int, rot_dim: int, base: float, device) -> tuple[torch.Tensor, torch.Tensor]:
    """fp32 cos / sin tables of shape [seq_len, rot_dim // 2]."""
    key = (seq_len, rot_dim, float(base), str(device))
    hit = _table_cache.get(key)
    if hit is None:
        inv = 1.0 / (base ** (torch.arange(0, rot_dim, 2, dtype=torch.float32, device=device) / rot_dim))
        ang = torch.outer(torch.arange(seq_len, dtype=torch.float32, device=device), inv)
        hit = (ang.cos().contiguous(), ang.sin().contiguous())
        _table_cache[key] = hit
    return hit


def _rotate_ref(x, cos, sin, rot, style, inverse):
    """x [B, h, S, d]; rotate the first ``rot`` features. style 1 = interleaved pairs, 2 = half split."""
    if style == 0 or rot == 0:
        return x
    S = x.shape[2]
    c = cos[:S][None, None].float()
    s = sin[:S][None, None].float()
    if inverse:
        s = -s
    xr, xp = x[..., :rot].float(), x[..., rot:]
    if style == 1:
        a, b = xr[..., 0::2], xr[..., 1::2]
        out = torch.stack([a * c - b * s, a * s + b * c], dim=-1).flatten(-2)
    else:
        h = rot // 2
        a, b = xr[..., :h], xr[..., h:]
        out = torch.cat([a * c - b * s, a * s + b * c], dim=-1)
    return torch.cat([out.to(x.dtype), xp], dim=-1)


class _QKVRotary(torch.autograd.Function):
    @staticmethod
    def forward(ctx, fused, B, S, heads, d, rot, style, cos, sin, pos_offset):
        ctx.meta = (B, S, heads, d, rot, style, pos_offset)
        ctx.save_for_backward(cos, sin)
        ctx.fused_shape = fused.shape
        f2 = fused.reshape(B * S, 3 * heads * d)
        if fused.is_cuda and f2.stride(-1) == 1 and fused.dtype in N._DTYPE_CODE:
            q = torch.empty(B, heads, S, d, dtype=fused.dtype, device=fused.device)
            k = torch.empty_like(q)
            v = torch.empty_like(q)
            N.count_launch()
            N.call("smpk_qkv_rotary", N.stream_ptr(fused.device), f2.data_ptr(), q.data_ptr(), k.data_ptr(),
                   v.data_ptr(), N.dtype_code(fused.dtype), N.ptr(cos), N.ptr(sin), B, S, heads, d, rot, style,
                   pos_offset, 0, f2.stride(0))
            return q, k, v
        f5 = f2.reshape(B, S, 3, heads, d).permute(2, 0, 3, 1, 4)
        q, k, v = f5[0], f5[1], f5[2]
        if style:
            cs, sn = cos[pos_offset:], sin[pos_offset:]
            q, k = _rotate_ref(q, cs, sn, rot, style, False), _rotate_ref(k, cs, sn, rot, style, False)
        return q.contiguous(), k.contiguous(), v.contiguous()

    @staticmethod
    def backward(ctx, dq, dk, dv):
        B, S, heads, d, rot, style, pos_offset = ctx.meta
        cos, sin = ctx.saved_tensors
        if dq.is_cuda and dq.dtype in N._DTYPE_CODE:
            dq, dk, dv = dq.contiguous(), dk.contiguous(), dv.contiguous()
            df = torch.empty(B * S, 3 * heads * d, dtype=dq.dtype, device=dq.device)
            N.count_launch()
            N.call("smpk_qkv_rotary", N.stream_ptr(dq.device), df.data_ptr(), dq.data_ptr(), dk.data_ptr(),
                   dv.data_ptr(), N.dtype_code(dq.dtype), N.ptr(cos), N.ptr(sin), B, S, heads, d, rot, style,
                   pos_offset, 1, df.stride(0))
        else:
            if style:
                cs, sn = cos[pos_offset:], sin[pos_offset:]
                dq, dk = _rotate_ref(dq, cs, sn, rot, style, True), _rotate_ref(dk, cs, sn, rot, style, True)
            df = torch.stack([dq, dk, dv], dim=0).permute(1, 3, 0, 2, 4).reshape(B * S, 3 * heads * d)
        return df.reshape(ctx.fused_shape), None, None, None, None, None, None, None, None, None


def qkv_split_rotary(fused, B, S, heads, d, rot=0, style=0, cos=None, sin=None, pos_offset=0):
    """fused [B*S, 3*heads*d] (q | k | v blocks, head-major) -> q, k, v each [B, heads, S, d]."""
    return _QKVRotary.apply(fused, B, S, heads, d, rot, style, cos, sin, pos_offset)


class _HeadTranspose(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, to_heads_first):
        ctx.flag = to_heads_first
        return _head_transpose(x, to_heads_first)

    @staticmethod
    def backward(ctx, dy):
        return _head_transpose(dy.contiguous(), not ctx.flag), None


def _head_transpose(x, to_heads_first):
    if to_heads_first:  # [B, S, h, d] -> [B, h, S, d]
        B, S, h, d = x.shape
        out_shape = (B, h, S, d)
    else:  # [B, h, S, d] -> [B, S, h, d]
        B, h, S, d = x.shape
        out_shape = (B, S, h, d)
    if x.is_cuda and x.is_contiguous() and d % 8 == 0 and x.dtype in N._DTYPE_CODE:
        out = torch.empty(out_shape, dtype=x.dtype, device=x.device)
        N.count_launch()
        N.call("smpk_head_transpose", N.stream_ptr(x.device), x.data_ptr(), out.data_ptr(), N.dtype_code(x.dtype), B,
               S, h, d, int(to_heads_first))
        return out
    return x.transpose(1, 2).contiguous()


def heads_to_tokens(x):
    """[B, heads, S, d] -> [B, S, heads*d]."""
    B, h, S, d = x.shape
    return _HeadTranspose.apply(x, False).reshape(B, S, h * d)


def tokens_to_heads(x, heads):
    """[B, S, heads*d] -> [B, heads, S, d]."""
    B, S, hd = x.shape
    return _HeadTranspose.apply(x.reshape(B, S, heads, hd // heads), True)


def _operand(t):
    """TMA needs a unit inner stride, 16-byte aligned outer strides and base; copy only when a view breaks that."""
    if t.stride(-1) != 1 or any(st % 8 for st in t.stride()[:-1]) or t.data_ptr() % 16:
        return t.contiguous()
    return t


def _mask_args(mask):
    if mask is None:
        return None, 0, 0, 0
    return (mask.data_ptr(), mask.stride(0) if mask.shape[0] > 1 else 0, mask.stride(1) if mask.shape[1] > 1 else 0,
            mask.stride(2) if mask.shape[2] > 1 else 0)


class _FlashAttention(torch.autograd.Function):
    """softmax(scale q k^T + masks) v on the fused tcgen05 kernels.  q, k, v are [B, heads, S, d] views; the context
    comes back either in the same layout or token-major ([B, Sq, heads, d]) so the output projection can consume it
    without a transpose.  Saved for backward: q, k, v, the context and one fp32 log-sum-exp per row."""

    @staticmethod
    def forward(ctx, q, k, v, mask, scale, causal, window, p_drop, seed, offset, token_major):
        B, h, Sq, d = q.shape
        Sk = k.shape[2]
        q, k, v = _operand(q), _operand(k), _operand(v)
        if token_major:
            o = torch.empty(B, Sq, h, d, dtype=q.dtype, device=q.device)
            o_str = (o.stride(0), o.stride(2), o.stride(1))
        else:
            o = torch.empty(B, h, Sq, d, dtype=q.dtype, device=q.device)
            o_str = (o.stride(0), o.stride(1), o.stride(2))
        lse = torch.empty(B * h, Sq, dtype=torch.float32, device=q.device)
        N.count_launch()
        N.call("smpk_attn_fwd", N.stream_ptr(q.device), N.dtype_code(q.dtype), B, h, Sq, Sk, d,
               q.data_ptr(), *q.stride()[:3], k.data_ptr(), *k.stride()[:3], v.data_ptr(), *v.stride()[:3],
               o.data_ptr(), *o_str, lse.data_ptr(), float(scale), int(causal), int(window), *_mask_args(mask),
               float(p_drop), int(seed), int(offset))
        ctx.save_for_backward(q, k, v, o, lse, mask)
        ctx.meta = (float(scale), int(causal), int(window), float(p_drop), int(seed), int(offset), token_major)
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse, mask = ctx.saved_tensors
        scale, causal, window, p_drop, seed, offset, token_major = ctx.meta
        B, h, Sq, d = q.shape
        Sk = k.shape[2]
        do = _operand(do)
        if token_major:
            o_str = (o.stride(0), o.stride(2), o.stride(1))
            g_str = (do.stride(0), do.stride(2), do.stride(1))
        else:
            o_str, g_str = o.stride()[:3], do.stride()[:3]
        dq = torch.empty(B, h, Sq, d, dtype=q.dtype, device=q.device)
        dk = torch.empty(B, h, Sk, d, dtype=q.dtype, device=q.device)
        dv = torch.empty(B, h, Sk, d, dtype=q.dtype, device=q.device)
        delta = torch.empty(B * h, Sq, dtype=torch.float32, device=q.device)
        N.count_launch(4 if d > 128 else 3)
        N.call("smpk_attn_bwd", N.stream_ptr(q.device), N.dtype_code(q.dtype), B, h, Sq, Sk, d,
               q.data_ptr(), *q.stride()[:3], k.data_ptr(), *k.stride()[:3], v.data_ptr(), *v.stride()[:3],
               o.data_ptr(), *o_str, do.data_ptr(), *g_str, lse.data_ptr(), delta.data_ptr(),
               dq.data_ptr(), *dq.stride()[:3], dk.data_ptr(), *dk.stride()[:3], dv.data_ptr(), *dv.stride()[:3],
               scale, causal, window, *_mask_args(mask), p_drop, seed, offset)
        return dq, dk, dv, None, None, None, None, None, None, None, None


def flash_applicable(q, k, v) -> bool:
    d = q.shape[-1]
    return (q.is_cuda and q.dtype in (torch.bfloat16, torch.float16) and k.dtype == q.dtype and v.dtype == q.dtype
            and d % 32 == 0 and 32 <= d <= 256 and v.shape[-1] == d)


def _dropout_counter():
    """(seed, offset) for the counter-based dropout of one attention call, drawn from the CPU generator: no device
    synchronisation, and an activation-checkpoint recomputation (which restores the generator) redraws the same."""
    r = torch.randint(0, 2 ** 31 - 1, (2,), dtype=torch.int64)
    return int(torch.initial_seed()) & (2 ** 63 - 1), (int(r[0]) << 31) | int(r[1])


# --------------------------------------------------------------------------------- core attention
def core_attention(q, k, v, *, causal: bool = True, mask=None, scale: float | None = None, dropout_p: float = 0.0,
                   training: bool = True, window: int | None = None, gemm_scale: float = 1.0,
                   token_major: bool = False, bias=None):
    """softmax(scale * q k^T + mask) v for q,k,v [B, heads, S, d]; returns [B, heads, Sq, d], or [B, Sq, heads, d]
    with ``token_major``.

    ``mask`` is a uint8/bool [B,1,Sq|1,Sk] tensor whose non-zero entries are excluded; ``causal`` applies the
    lower-triangular mask (aligned to the last key when Sq != Sk); ``window`` restricts every query to the last
    ``window`` visible keys (local attention).  ``gemm_scale`` is accepted for the reference's
    ``query_key_layer_scaling`` / ``attention_in_fp32`` options (reference torch/nn/transformer.py:1769-1835): their
    purpose - keeping 16-bit scores in range - is moot here because scores only exist as fp32 accumulators in
    tensor memory, so it does not change the result.  ``bias`` is an additive score bias; calls that carry one run on the
    library's scaled-dot-product attention instead of the fused kernels.
    """
    B, h, Sq, d = q.shape
    Sk = k.shape[2]
    scale = (1.0 / math.sqrt(d)) if scale is None else scale
    p_drop = float(dropout_p) if training else 0.0
    if mask is not None:
        mask = mask.to(torch.uint8) if mask.dtype != torch.uint8 else mask
        if mask.dim() != 4 or mask.stride(-1) != 1:
            mask = mask.reshape(mask.shape[0], 1, -1, mask.shape[-1]).contiguous()
    if bias is None and flash_applicable(q, k, v):
        seed, offset = _dropout_counter() if p_drop > 0.0 else (0, 0)
        return _FlashAttention.apply(q, k, v, mask, float(scale), bool(causal), int(window or 0), p_drop, seed, offset,
                                     token_major)
    # CPU tensors, fp32 or head sizes the kernels do not tile: the library's scaled-dot-product attention
    dead = None
    if causal or window:
        i = torch.arange(Sq, device=q.device)[:, None] + (Sk - Sq)
        j = torch.arange(Sk, device=q.device)[None, :]
        dead = (j > i) if causal else torch.zeros(Sq, Sk, dtype=torch.bool, device=q.device)
        if window:
            dead = dead | (j <= i - window)
        dead = dead[None, None]
    if mask is not None:
        dead = mask.bool() if dead is None else (dead | mask.bool())
    keep = None if dead is None else ~dead
    if bias is not None:
        # additive score bias ([B|1, heads|1, Sq, Sk], e.g. T5's relative position bias): only the library path takes one
        add = bias.to(q.dtype)
        keep = add if dead is None else add.masked_fill(dead, float("-inf"))
    out = torch.nn.functional.scaled_dot_product_attention(q, k, v, attn_mask=keep, dropout_p=p_drop, scale=scale)
    if keep is not None:
        out = torch.nan_to_num(out, nan=0.0)            # rows without a visible key
    return out.transpose(1, 2).contiguous() if token_major else out
