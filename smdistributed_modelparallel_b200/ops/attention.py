"""Attention building blocks on the sm_100a kernels: autograd batched GEMMs (tcgen05), QKV split with
fused rotary embedding, head transposes and the composed core attention.

The reference computes attention as ``baddbmm`` -> fused softmax -> dropout -> ``matmul`` (reference
smdistributed/modelparallel/torch/nn/transformer.py:1617-1657,1769-1835) and rotary embeddings with
eager ops (:114-163,1565-1615).  Here both batched GEMMs run on the tcgen05 kernel in the layouts the
data already has (no transposed copies), the softmax is the native causal/masked kernel and rotary is
fused into the QKV split.
"""

from __future__ import annotations

import math

import torch

from . import _native as N
from .gemm import gemm
from .softmax import scaled_masked_softmax, scaled_upper_triang_masked_softmax


class _BmmNT(torch.autograd.Function):
    """out[b] = alpha * a[b] @ b[b]^T   (a [B,M,K], b [B,N,K])."""

    @staticmethod
    def forward(ctx, a, b, alpha):
        ctx.alpha = alpha
        ctx.save_for_backward(a, b)
        return gemm(a, b, alpha=alpha)

    @staticmethod
    def backward(ctx, dout):
        a, b = ctx.saved_tensors
        dout = dout.contiguous()
        da = gemm(dout, b, b_mn_major=True, alpha=ctx.alpha) if ctx.needs_input_grad[0] else None
        db = gemm(dout, a, a_mn_major=True, b_mn_major=True, alpha=ctx.alpha) if ctx.needs_input_grad[1] else None
        return da, db, None


class _BmmNN(torch.autograd.Function):
    """out[b] = a[b] @ b[b]   (a [B,M,K], b [B,K,N])."""

    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        return gemm(a, b, b_mn_major=True)

    @staticmethod
    def backward(ctx, dout):
        a, b = ctx.saved_tensors
        dout = dout.contiguous()
        da = gemm(dout, b) if ctx.needs_input_grad[0] else None
        db = gemm(a, dout, a_mn_major=True, b_mn_major=True) if ctx.needs_input_grad[1] else None
        return da, db


def bmm_nt(a, b, alpha: float = 1.0):
    return _BmmNT.apply(a.contiguous(), b.contiguous(), alpha)


def bmm_nn(a, b):
    return _BmmNN.apply(a.contiguous(), b.contiguous())


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


class _CausalAttention(torch.autograd.Function):
    """softmax(scale q k^T + causal mask) v with every GEMM told about the triangle: score tiles above the
    diagonal are never computed or written, and the products with P / dS (lower triangular) only reduce over the
    non-zero part - about 44 % of the attention FLOPs and score traffic at S = 2048."""

    @staticmethod
    def forward(ctx, q, k, v, scale, gemm_scale=1.0):
        from .gemm import gemm

        BH, S, d = q.shape
        scores = torch.empty(BH, S, S, dtype=q.dtype, device=q.device)
        gemm(q, k, out=scores, causal="lower_out", alpha=gemm_scale)    # stored scores = gemm_scale * q k^T
        probs = torch.empty_like(scores)
        N.count_launch()
        N.call("smpk_softmax_fwd", N.stream_ptr(q.device), scores.data_ptr(), probs.data_ptr(), N.dtype_code(q.dtype),
               None, 3, float(scale), BH, 1, S, S)       # mode 3: causal, zero fill only up to the next 256 columns
        del scores
        out = gemm(probs, v, b_mn_major=True, causal="lower_a")
        ctx.save_for_backward(q, k, v, probs)
        ctx.scale, ctx.gemm_scale = scale, gemm_scale
        return out

    @staticmethod
    def backward(ctx, dout):
        from .gemm import gemm

        q, k, v, probs = ctx.saved_tensors
        dout = dout.contiguous()
        BH, S, d = q.shape
        dv = gemm(probs, dout, a_mn_major=True, b_mn_major=True, causal="lower_at")      # P^T dO
        dp = torch.empty_like(probs)
        gemm(dout, v, out=dp, causal="lower_out")                                        # dO V^T (lower tiles)
        ds = torch.empty_like(probs)
        N.count_launch()
        N.call("smpk_softmax_bwd", N.stream_ptr(q.device), dp.data_ptr(), probs.data_ptr(), ds.data_ptr(),
               N.dtype_code(q.dtype), 2, float(ctx.scale), BH, S, S)
        del dp
        a = ctx.gemm_scale
        dq = gemm(ds, k, b_mn_major=True, causal="lower_a", alpha=a)                     # dS K
        dk = gemm(ds, q, a_mn_major=True, b_mn_major=True, causal="lower_at", alpha=a)   # dS^T Q
        return dq, dk, dv, None, None


def _causal_fast_path(q, k, mask, window, dropout_p, training) -> bool:
    return (q.is_cuda and mask is None and window is None and q.shape[2] == k.shape[2] and q.shape[2] % 8 == 0 and q.shape[2] <= 8192
            and q.shape[3] % 8 == 0 and q.dtype in (torch.bfloat16, torch.float16) and not (dropout_p > 0.0 and training)
            and N.available())


# --------------------------------------------------------------------------------- core attention
def core_attention(q, k, v, *, causal: bool = True, mask=None, scale: float | None = None, dropout_p: float = 0.0,
                   training: bool = True, window: int | None = None, gemm_scale: float = 1.0):
    """softmax(scale * q k^T + mask) v for q,k,v [B, heads, S, d]; returns [B, heads, Sq, d].

    ``mask`` is a uint8/bool [B,1,Sq,Sk] tensor whose non-zero entries are excluded; ``causal`` applies
    the lower-triangular mask; ``window`` (local attention) is folded into the byte mask.  ``gemm_scale`` is the
    part of ``scale`` applied by the score GEMM itself (the stored 16-bit scores are ``gemm_scale * q k^T``, the
    softmax applies the rest) - the reference's ``query_key_layer_scaling`` / ``attention_in_fp32`` guard against
    fp16 overflow of the raw scores (reference torch/nn/transformer.py:1769-1835).
    """
    B, h, Sq, d = q.shape
    Sk = k.shape[2]
    scale = (1.0 / math.sqrt(d)) if scale is None else scale
    if causal and _causal_fast_path(q, k, mask, window, dropout_p, training):
        ctx = _CausalAttention.apply(q.reshape(B * h, Sq, d).contiguous(), k.reshape(B * h, Sk, d).contiguous(),
                                     v.reshape(B * h, Sk, d).contiguous(), float(scale / gemm_scale), float(gemm_scale))
        return ctx.reshape(B, h, Sq, d)
    scores = bmm_nt(q.reshape(B * h, Sq, d), k.reshape(B * h, Sk, d), alpha=gemm_scale)
    scale = scale / gemm_scale
    if window is not None:
        i = torch.arange(Sq, device=q.device)[:, None] + (Sk - Sq)
        j = torch.arange(Sk, device=q.device)[None, :]
        wmask = ((j > i) if causal else torch.zeros(Sq, Sk, dtype=torch.bool, device=q.device)) | (j <= i - window)
        wmask = wmask[None, None].expand(B, 1, Sq, Sk)
        mask = wmask if mask is None else (mask.bool() | wmask)
        causal = False
    if mask is not None:
        if causal:
            tri = torch.ones(Sq, Sk, dtype=torch.bool, device=q.device).triu(diagonal=1 + Sk - Sq)
            mask = mask.bool() | tri[None, None]
        probs = scaled_masked_softmax(scores.reshape(B, h, Sq, Sk), mask, scale).reshape(B * h, Sq, Sk)
    elif causal:
        probs = scaled_upper_triang_masked_softmax(scores, scale)
    else:
        probs = scaled_masked_softmax(scores.reshape(B, h, Sq, Sk), None, scale).reshape(B * h, Sq, Sk)
    if dropout_p > 0.0 and training:
        probs = torch.nn.functional.dropout(probs, dropout_p, True)
    ctx = bmm_nn(probs, v.reshape(B * h, Sk, d))
    return ctx.reshape(B, h, Sq, d)
