"""Fused tcgen05 attention (csrc/kernels/sm100/attention.cu) against an fp32 PyTorch restatement: output, log-sum-exp
and dq / dk / dv over head sizes 64-256, causal / byte / padding / local-window masks, Sq != Sk, ragged lengths, fp16 and
counter-based dropout (mask recovered through V = I), plus the property the kernel exists for - no Sq x Sk tensor.

The reference's tier compares its fused softmax with F.softmax at [4,16,1024,1024] (reference
smdistributed/modelparallel/test/torch/test_kernels.py:14-86); its attention itself is an eager composition
(torch/nn/transformer.py:1617-1657).
"""

import math

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda"


def _ref(q, k, v, do, scale, causal, window, mask, keep=None, p_drop=0.0):
    qf, kf, vf = (t.float().detach().requires_grad_(True) for t in (q, k, v))
    s = torch.einsum("bhqd,bhkd->bhqk", qf, kf) * scale
    Sq, Sk = q.shape[2], k.shape[2]
    i = torch.arange(Sq, device=q.device)[:, None] + (Sk - Sq)
    j = torch.arange(Sk, device=q.device)[None, :]
    dead = torch.zeros(Sq, Sk, dtype=torch.bool, device=q.device)
    if causal:
        dead |= j > i
    if window:
        dead |= j <= i - window
    s = s.masked_fill(dead[None, None], float("-inf"))
    if mask is not None:
        s = s.masked_fill(mask.bool(), float("-inf"))
    p = torch.nan_to_num(torch.softmax(s, dim=-1), nan=0.0)
    if keep is not None:
        p = p * keep / (1.0 - p_drop)
    out = torch.einsum("bhqk,bhkd->bhqd", p, vf)
    out.backward(do.float())
    return out.detach(), qf.grad, kf.grad, vf.grad


def _rel(a, r):
    return (a.float() - r).abs().max().item() / (r.abs().max().item() + 1e-6)


CASES = [
    # B, H, Sq, Sk, d, causal, window, mask, dtype
    (1, 1, 128, 128, 64, False, 0, None, torch.bfloat16),
    (2, 3, 256, 256, 64, True, 0, None, torch.bfloat16),
    (2, 2, 512, 512, 128, True, 0, None, torch.bfloat16),
    (2, 2, 512, 512, 96, True, 0, None, torch.bfloat16),
    (1, 2, 1024, 1024, 256, True, 0, None, torch.bfloat16),
    (1, 2, 300, 300, 64, True, 0, None, torch.float16),
    (2, 2, 200, 333, 64, False, 0, "pad", torch.bfloat16),
    (1, 2, 384, 384, 128, False, 0, "full", torch.bfloat16),
    (1, 2, 640, 640, 128, True, 256, None, torch.bfloat16),
    (1, 2, 128, 512, 64, True, 0, None, torch.bfloat16),
    (1, 1, 192, 192, 192, True, 0, None, torch.bfloat16),
    (1, 2, 2304, 2304, 64, True, 0, None, torch.bfloat16),
]


@pytest.mark.parametrize("B,H,Sq,Sk,d,causal,window,mt,dt", CASES)
@pytest.mark.parametrize("token_major", [False, True])
def test_flash_attention_matches_fp32(B, H, Sq, Sk, d, causal, window, mt, dt, token_major):
    from smdistributed_modelparallel_b200.ops import attention as A

    torch.manual_seed(Sq * 7 + d)
    q, k, v = [torch.randn(B, H, s, d, device=DEV, dtype=dt).requires_grad_(True) for s in (Sq, Sk, Sk)]
    mask = None
    if mt == "pad":
        mask = torch.zeros(B, 1, 1, Sk, dtype=torch.uint8, device=DEV)
        mask[0, :, :, Sk - 50:] = 1
        mask[-1, :, :, Sk - 7:] = 1
    elif mt == "full":
        mask = (torch.rand(B, 1, Sq, Sk, device=DEV) < 0.3).to(torch.uint8)
    assert A.flash_applicable(q, k, v)
    scale = 1.0 / math.sqrt(d)
    o = A.core_attention(q, k, v, causal=causal, mask=mask, scale=scale, window=window or None, token_major=token_major)
    assert o.shape == ((B, Sq, H, d) if token_major else (B, H, Sq, d))
    go = torch.randn_like(o)
    dq, dk, dv = torch.autograd.grad(o, (q, k, v), go)
    o_bh = o.transpose(1, 2) if token_major else o
    go_bh = go.transpose(1, 2) if token_major else go
    ro, rq, rk, rv = _ref(q, k, v, go_bh, scale, causal, window, mask)
    tol = 2e-2
    assert _rel(o_bh, ro) < tol
    for name, a, r in (("dq", dq, rq), ("dk", dk, rk), ("dv", dv, rv)):
        assert _rel(a, r) < tol, (name, _rel(a, r))


@pytest.mark.parametrize("d", [64, 256])
def test_flash_attention_dropout(d):
    """The keep mask is a pure function of (seed, offset, batch*head, query, key): recover it with V = I (then a
    row of the output is the row of dropped probabilities), check the drop rate, determinism across calls (what an
    activation-checkpoint recomputation relies on) and forward / backward against fp32 with that mask."""
    from smdistributed_modelparallel_b200.ops.attention import _FlashAttention

    torch.manual_seed(3)
    B, H, S, p_drop, seed, offset = 1, 2, d, 0.2, 1234, 77
    dt = torch.bfloat16
    q, k, v = [torch.randn(B, H, S, d, device=DEV, dtype=dt).requires_grad_(True) for _ in range(3)]
    eye = torch.eye(S, device=DEV, dtype=dt)[None, None].expand(B, H, S, d).contiguous()
    scale = 1.0 / math.sqrt(d)
    run = lambda vv, off=offset: _FlashAttention.apply(q, k, vv, None, scale, True, 0, p_drop, seed, off, False)
    o_eye = run(eye)
    tri = torch.ones(S, S, device=DEV, dtype=torch.bool).tril()
    keep = (o_eye != 0) | ~tri[None, None]
    frac = 1.0 - keep[:, :, tri].float().mean().item()
    assert abs(frac - p_drop) < 0.03, frac
    o = run(v)
    assert torch.equal(o, run(v))                                      # same counter -> same mask
    assert not torch.equal(o, run(v, offset + 1))                      # next call -> another mask
    go = torch.randn_like(o)
    dq, dk, dv = torch.autograd.grad(o, (q, k, v), go)
    ro, rq, rk, rv = _ref(q, k, v, go, scale, True, 0, None, keep.float(), p_drop)
    assert _rel(o, ro) < 2e-2
    for name, a, r in (("dq", dq, rq), ("dk", dk, rk), ("dv", dv, rv)):
        assert _rel(a, r) < 2e-2, (name, _rel(a, r))


def test_flash_attention_allocates_nothing_of_size_sq_by_sk():
    """Forward + backward at S = 4096 must stay within a few operand-sized buffers: one [B*H, S, S] bf16 tensor
    would be 1 GiB here."""
    from smdistributed_modelparallel_b200.ops import attention as A

    B, H, S, d = 2, 16, 4096, 128
    q, k, v = [torch.randn(B, H, S, d, device=DEV, dtype=torch.bfloat16).requires_grad_(True) for _ in range(3)]
    operand = q.numel() * 2
    torch.cuda.synchronize()
    torch.cuda.reset_peak_memory_stats()
    base = torch.cuda.memory_allocated()
    o = A.core_attention(q, k, v, causal=True)
    torch.autograd.grad(o, (q, k, v), torch.ones_like(o))
    torch.cuda.synchronize()
    extra = torch.cuda.max_memory_allocated() - base
    assert extra < 8 * operand, (extra, operand)
    assert extra < (B * H * S * S * 2) // 4


def test_dropout_counter_is_restored_by_activation_checkpointing():
    """A checkpointed layer with attention dropout recomputes with the mask of its first forward pass."""
    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.ops import attention as A

    smp.init({"microbatches": 1})
    torch.manual_seed(11)
    q, k, v = [torch.randn(1, 2, 256, 64, device=DEV, dtype=torch.bfloat16).requires_grad_(True) for _ in range(3)]

    def attn(q, k, v):
        return A.core_attention(q, k, v, causal=True, dropout_p=0.3, training=True)

    torch.manual_seed(5)
    o1 = attn(q, k, v)
    g1 = torch.autograd.grad(o1.float().pow(2).sum(), (q, k, v))
    torch.manual_seed(5)
    o2 = smp.checkpoint(attn, q, k, v)
    g2 = torch.autograd.grad(o2.float().pow(2).sum(), (q, k, v))
    assert torch.equal(o1, o2)
    for a, b in zip(g1, g2):
        assert torch.equal(a, b)
