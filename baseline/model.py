"""Reference-layout tensor-parallel GPT (GPT-2 / GPT-J / GPT-NeoX shapes) in plain PyTorch + NCCL."""

from __future__ import annotations

import math

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


# ------------------------------------------------------------------------------------------- tp collectives
def _size(group):
    return 1 if group is None else dist.get_world_size(group)


class _AllReduceFwd(torch.autograd.Function):
    """all-reduce in forward, identity in backward (after a row-parallel linear)."""

    @staticmethod
    def forward(ctx, x, group):
        if _size(group) > 1:
            x = x.contiguous()
            dist.all_reduce(x, group=group)
        return x

    @staticmethod
    def backward(ctx, g):
        return g, None


class _AllReduceBwd(torch.autograd.Function):
    """identity in forward, all-reduce of the gradient in backward (before a column-parallel linear)."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return x

    @staticmethod
    def backward(ctx, g):
        if _size(ctx.group) > 1:
            g = g.contiguous()
            dist.all_reduce(g, group=ctx.group)
        return g, None


class _GatherBatch(torch.autograd.Function):
    """[b, ...] -> [b * t, ...] over the tp group; backward keeps this rank's slice of the (already reduced) gradient."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        n = _size(group)
        if n == 1:
            return x
        x = x.contiguous()
        out = torch.empty((x.shape[0] * n,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(out, x, group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        n = _size(ctx.group)
        if n == 1:
            return g, None
        r = dist.get_rank(ctx.group)
        per = g.shape[0] // n
        return g.narrow(0, r * per, per).contiguous(), None


class _NarrowBatch(torch.autograd.Function):
    """[b * t, ...] -> this rank's [b, ...]; backward all-gathers."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        n = _size(group)
        if n == 1:
            return x
        r = dist.get_rank(group)
        per = x.shape[0] // n
        return x.narrow(0, r * per, per).contiguous()

    @staticmethod
    def backward(ctx, g):
        n = _size(ctx.group)
        if n == 1:
            return g, None
        g = g.contiguous()
        out = torch.empty((g.shape[0] * n,) + tuple(g.shape[1:]), dtype=g.dtype, device=g.device)
        dist.all_gather_into_tensor(out, g, group=ctx.group)
        return out, None


# ------------------------------------------------------------------------------------------- rotary
def _rotary_tables(seq, rot, base, device):
    inv = 1.0 / (base ** (torch.arange(0, rot, 2, dtype=torch.float32, device=device) / rot))
    ang = torch.outer(torch.arange(seq, dtype=torch.float32, device=device), inv)
    return ang.cos(), ang.sin()


def _apply_rotary(x, cos, sin, rot, neox):
    """x [B, h, S, d]; eager ops as in the reference (nn/transformer.py:114-163,1565-1615)."""
    xr, xp = x[..., :rot], x[..., rot:]
    c, s = cos[None, None].to(x.dtype), sin[None, None].to(x.dtype)
    if neox:
        half = rot // 2
        a, b = xr[..., :half], xr[..., half:]
        out = torch.cat([a * c - b * s, a * s + b * c], dim=-1)
    else:
        a, b = xr[..., 0::2], xr[..., 1::2]
        out = torch.stack([a * c - b * s, a * s + b * c], dim=-1).flatten(-2)
    return torch.cat([out, xp], dim=-1)


def _gelu(x, kind):
    if kind in ("gelu_new", "gelu_fast", "gelu_pytorch_tanh"):
        return F.gelu(x, approximate="tanh")
    return F.gelu(x)


# ------------------------------------------------------------------------------------------- layer
class Layer(nn.Module):
    def __init__(self, cfg, group, dtype, device):
        super().__init__()
        H, I, h, d = cfg["hidden_size"], cfg["intermediate_size"], cfg["num_attention_heads"], cfg["attention_head_size"]
        t = _size(group)
        assert h % t == 0 and I % t == 0
        self.cfg, self.group = cfg, group
        self.hl, self.d = h // t, d
        kw = dict(dtype=dtype, device=device)
        qkv_bias = cfg.get("use_qkv_bias", True)
        self.ln1 = nn.LayerNorm(H, eps=cfg["layernorm_epsilon"], **kw)
        self.ln2 = None if cfg.get("single_pre_layernorm", False) else nn.LayerNorm(H, eps=cfg["layernorm_epsilon"], **kw)
        self.q = nn.Linear(H, self.hl * d, bias=qkv_bias, **kw)
        self.k = nn.Linear(H, self.hl * d, bias=qkv_bias, **kw)
        self.v = nn.Linear(H, self.hl * d, bias=qkv_bias, **kw)
        self.dense = nn.Linear(self.hl * d, H, bias=False, **kw)          # row-parallel; bias added after the all-reduce
        self.dense_bias = nn.Parameter(torch.zeros(H, **kw)) if cfg.get("use_attn_dense_bias", True) else None
        self.fc1 = nn.Linear(H, I // t, **kw)
        self.fc2 = nn.Linear(I // t, H, bias=False, **kw)
        self.fc2_bias = nn.Parameter(torch.zeros(H, **kw))
        std = cfg.get("initializer_range", 0.02)
        for lin in (self.q, self.k, self.v, self.dense, self.fc1, self.fc2):
            nn.init.normal_(lin.weight, std=std)
        self.rot = cfg.get("rotary_dim") or 0
        self.neox = bool(cfg.get("gpt_neox_type_rotary", False))
        self.parallel = bool(cfg.get("parallel_attn_output", False))
        self.p_attn, self.p_hidden = cfg.get("attention_dropout_prob", 0.0), cfg.get("hidden_dropout_prob", 0.0)

    def _attention(self, x, rotary):
        B, S, _ = x.shape
        g = self.group
        xin = _AllReduceBwd.apply(x, g)
        q = self.q(xin).view(B, S, self.hl, self.d).transpose(1, 2)
        k = self.k(xin).view(B, S, self.hl, self.d).transpose(1, 2)
        v = self.v(xin).view(B, S, self.hl, self.d).transpose(1, 2)
        if self.rot:
            cos, sin = rotary
            q, k = _apply_rotary(q, cos[:S], sin[:S], self.rot, self.neox), _apply_rotary(k, cos[:S], sin[:S], self.rot, self.neox)
        q3, k3, v3 = (t.reshape(B * self.hl, S, self.d) for t in (q, k, v))
        scores = torch.baddbmm(torch.empty(B * self.hl, S, S, dtype=q.dtype, device=q.device), q3, k3.transpose(1, 2),
                               beta=0.0, alpha=1.0 / math.sqrt(self.d))
        causal = torch.ones(S, S, dtype=torch.bool, device=q.device).triu(1)
        probs = torch.softmax(scores.masked_fill(causal, float("-inf")), dim=-1)
        probs = F.dropout(probs, self.p_attn, self.training)
        ctx = torch.bmm(probs, v3).view(B, self.hl, S, self.d).transpose(1, 2).reshape(B, S, self.hl * self.d)
        out = _AllReduceFwd.apply(self.dense(ctx), g)
        return out if self.dense_bias is None else out + self.dense_bias

    def _mlp(self, x):
        g = self.group
        y = _gelu(self.fc1(_AllReduceBwd.apply(x, g)), self.cfg.get("activation", "gelu"))
        return _AllReduceFwd.apply(self.fc2(y), g) + self.fc2_bias

    def forward(self, x, rotary):
        if self.parallel:            # GPT-J / GPT-NeoX: x + attn(ln1 x) + mlp(ln x)
            h1 = self.ln1(x)
            h2 = h1 if self.ln2 is None else self.ln2(x)
            return x + F.dropout(self._attention(h1, rotary), self.p_hidden, self.training) + \
                F.dropout(self._mlp(h2), self.p_hidden, self.training)
        x = x + F.dropout(self._attention(self.ln1(x), rotary), self.p_hidden, self.training)
        return x + F.dropout(self._mlp(self.ln2(x)), self.p_hidden, self.training)


class _VocabParallelCE(torch.autograd.Function):
    """Cross-entropy over vocabulary-sharded logits: MAX, SUM, SUM all-reduces of one value per token."""

    @staticmethod
    def forward(ctx, logits, target, v0, group):
        n = _size(group)
        lf = logits.float()
        mx = lf.max(dim=-1).values
        if n > 1:
            dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
        lf = lf - mx[:, None]
        local = target - v0
        inside = (local >= 0) & (local < lf.shape[-1])
        picked = lf.gather(1, local.clamp(0, lf.shape[-1] - 1)[:, None])[:, 0] * inside
        ex = lf.exp()
        se = ex.sum(dim=-1)
        if n > 1:
            dist.all_reduce(picked, group=group)
            dist.all_reduce(se, group=group)
        ctx.save_for_backward(ex / se[:, None], local.clamp(0, lf.shape[-1] - 1), inside)
        ctx.dt = logits.dtype
        return se.log() - picked

    @staticmethod
    def backward(ctx, g):
        sm, idx, inside = ctx.saved_tensors
        grad = sm.clone()
        grad.scatter_add_(1, idx[:, None], -inside.float()[:, None])
        return (grad * g[:, None]).to(ctx.dt), None, None, None


class RefGPT(nn.Module):
    """``layers`` selects the slice of the stack this pipeline stage owns; ``first`` / ``last`` add the embedding and
    the final LayerNorm + LM head + loss."""

    def __init__(self, cfg, group, dtype, device, layers=None, first=True, last=True):
        super().__init__()
        self.cfg, self.group = cfg, group
        H, V = cfg["hidden_size"], cfg["vocab_size"]
        t = _size(group)
        r = dist.get_rank(group) if t > 1 else 0
        self.vpr = (V + t - 1) // t
        self.v0 = r * self.vpr
        kw = dict(dtype=dtype, device=device)
        lo, hi = layers if layers is not None else (0, cfg["num_layers"])
        self.first, self.last = first, last
        if first:
            self.wte = nn.Parameter(torch.empty(self.vpr, H, **kw).normal_(std=cfg.get("initializer_range", 0.02)))
            self.wpe = nn.Parameter(torch.empty(cfg["num_positions"], H, **kw).normal_(std=0.02)) \
                if cfg.get("use_positional_embedding", True) else None
        self.layers = nn.ModuleList([Layer(cfg, group, dtype, device) for _ in range(lo, hi)])
        if last:
            self.ln_f = nn.LayerNorm(H, eps=cfg["layernorm_epsilon"], **kw)
            self.head = nn.Parameter(torch.empty(self.vpr, H, **kw).normal_(std=cfg.get("initializer_range", 0.02)))
        self._rot = None
        self.ckpt = False

    def _rotary(self, S, device):
        rot = self.cfg.get("rotary_dim") or 0
        if not rot:
            return None
        if self._rot is None or self._rot[0].shape[0] < S:
            self._rot = _rotary_tables(max(S, self.cfg["num_positions"]), rot, float(self.cfg.get("rotary_emb_base", 10000)), device)
        return self._rot

    def embed(self, ids):
        g = self.group
        ids = _GatherBatch.apply(ids, g)                       # the tp group works on the concatenated batch
        local = ids - self.v0
        inside = (local >= 0) & (local < self.vpr)
        x = F.embedding(local.clamp(0, self.vpr - 1), self.wte) * inside[..., None].to(self.wte.dtype)
        x = _AllReduceFwd.apply(x, g)
        if self.wpe is not None:
            x = x + self.wpe[: ids.shape[1]][None]
        return x

    def run_layers(self, x):
        rotary = self._rotary(x.shape[1], x.device)
        for layer in self.layers:
            if self.ckpt and torch.is_grad_enabled():
                x = torch.utils.checkpoint.checkpoint(layer, x, rotary, use_reentrant=False)
            else:
                x = layer(x, rotary)
        return x

    def loss(self, x, ids):
        g = self.group
        labels = _GatherBatch.apply(ids, g)
        x = self.ln_f(x)
        B, S, H = x.shape
        logits = F.linear(_AllReduceBwd.apply(x[:, :-1].reshape(-1, H), g), self.head)
        per_tok = _VocabParallelCE.apply(logits, labels[:, 1:].reshape(-1), self.v0, g)
        return per_tok.mean()

    def forward(self, ids):
        return self.loss(self.run_layers(self.embed(ids)), ids)
