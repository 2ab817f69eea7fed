"""Tensor-parallel transformer stack: DistributedTransformerLMHead / DistributedTransformer /
DistributedTransformerLayer / DistributedAttentionLayer / DistributedTransformerOutputLayer.

Constructor keywords, module tree and parameter names follow the reference so configurations and
checkpoints carry over (reference smdistributed/modelparallel/torch/nn/transformer.py:189-236 _KEYS,
:368-548 LM-head forward, :854-962 layer forward, :1077-1173 output layer, :1403-1546 attention layer;
names ``word_embedding``, ``position_embedding``, ``transformer.seq_layers.{i}.attention.{query,key,
value,dense,pre_layernorm,layernorm}``, ``...output.{dense1,dense2,pre_layernorm,layernorm}``,
``layernorm``, ``lm_head``: reference torch/nn/huggingface/gptneox.py:163-243).

Execution differs from the reference by design (B200-first):
* token-parallel layout - each tp rank keeps only its own tokens between GEMMs; column-parallel GEMMs
  all-gather tokens first, row-parallel GEMMs reduce-scatter after (``parallel/tp_ops.py``), so
  LayerNorm / dropout / residual touch 1/t of the tokens and no all-reduce remains;
* q, k, v are one fused GEMM on the tcgen05 kernel (the three reference Linear modules survive as
  state-dict names only), bias+GELU is a GEMM epilogue, rotary is fused into the QKV split, and the attention core
  (scores, masks, softmax, dropout, context) is one flash-style kernel pair with no length limit and no
  ``S x S`` tensor (``ops/attention.py``, ``csrc/kernels/sm100/attention.cu``);
* ``optimize: memory`` selects the reference's other layout instead (hidden-sharded activations, ``_forward_memory``).
"""

from __future__ import annotations

import math
import os
from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..backend.exceptions import DistTransformerConfigError
from ..ops import attention as A
from ..ops import attention as A_ops
from ..ops.cross_entropy import vocab_parallel_cross_entropy
from ..parallel import tp_ops
from .dist_module import DistributedModule, parse_args, update_copy_dict
from .layer_norm import FusedLayerNorm
from .utils import (get_local_channels, get_start_pos_for_slicing, initialize_with_input_partition,
                    initialize_with_output_partition, mark_one_rank, parameter_creation_scope, tp_group, tp_rank,
                    tp_size)

_TANH_ACTS = ("gelu_new", "gelu_fast", "gelu_pytorch_tanh", "gelu_tanh")


def _state():
    from ..state_mod import state
    return state


def _param_dtype():
    s = _state()
    if getattr(s, "initialized", False) and s.cfg is not None:
        if s.cfg._fp16_param_init:
            return torch.float16
    return None


def _group():
    return tp_group() if tp_size() > 1 else None


def _memory_layout() -> bool:
    """``optimize: memory`` (reference torch/nn/transformer.py:1429-1472,1519,1112,1138): between the GEMMs the
    activations are ``[full batch, S, H / t]`` - every rank sees every token but only its slice of the hidden
    dimension; q / k / v and the first MLP projection are partitioned along their *input* features and followed by a
    reduce-scatter over the output features, LayerNorms are hidden-sharded (``DistributedLayerNorm``), and an all-to-all
    converts from / to the ``[local batch, S, H]`` layout at the ends of a layer stack."""
    s = _state()
    return bool(getattr(s, "initialized", False) and s.cfg is not None and s.cfg.optimize == "memory" and tp_size() > 1)


def _checkpoint_attentions() -> bool:
    s = _state()
    return bool(getattr(s, "initialized", False) and s.cfg is not None and s.cfg.checkpoint_attentions)


def _plain_linear(x2, weight):
    """x2 [rows, K] @ weight[N, K]^T on the library GEMM (autograd, wgrad into main_grad), no collective."""
    return tp_ops.ColumnParallelLinear.apply(x2, weight, None, None, None, True)


def _gelu_mode(activation: str):
    # SMP_USE_HF_GELU=1 (reference torch/nn/transformer.py:994): plain torch GELU after the GEMM instead of the
    # GELU folded into the GEMM epilogue - a debugging switch for bit-comparisons with HuggingFace
    if os.environ.get("SMP_USE_HF_GELU") == "1":
        return None
    if activation == "gelu":
        return "erf"
    if activation in _TANH_ACTS:
        return "tanh"
    return None


class _NamedSlices:
    """State-dict translation for a fused parameter stored as consecutive row blocks.

    ``attention.qkv_weight`` ([3*local, H]) is saved as ``attention.query.weight``, ``key.weight`` and
    ``value.weight`` so checkpoints keep the reference's names and per-rank shapes."""

    @staticmethod
    def install(module, fused_names, part_names):
        def save_hook(mod, state_dict, prefix, local_metadata):
            for fused, suffix in fused_names:
                key = prefix + fused
                if key not in state_dict:
                    continue
                t = state_dict.pop(key)
                n = t.shape[0] // len(part_names)
                for i, p in enumerate(part_names):
                    state_dict[f"{prefix}{p}.{suffix}"] = t[i * n:(i + 1) * n]
            return state_dict

        def load_hook(state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
            for fused, suffix in fused_names:
                keys = [f"{prefix}{p}.{suffix}" for p in part_names]
                if all(k in state_dict for k in keys):
                    state_dict[prefix + fused] = torch.cat([state_dict.pop(k) for k in keys], dim=0)

        module._register_state_dict_hook(save_hook)
        module._register_load_state_dict_pre_hook(load_hook)


class DistributedAttentionLayer(DistributedModule):
    """Self / cross attention block (reference torch/nn/transformer.py:1044-1420).

    Keys that select between numerically equivalent implementations in the reference and have one implementation
    here: ``fused_softmax`` (reference :1617-1657 chooses its softmax kernel or three torch ops; here the softmax
    lives inside the attention kernel, csrc/kernels/sm100/attention.cu, whatever the flag says) and, on the MLP,
    ``fused_bias_gelu`` (bias + GELU are the epilogue of the first GEMM).  ``mask_value`` is the finite number the
    reference adds to masked scores; the kernel drops masked keys from the softmax exactly, which is the limit of
    that as mask_value -> -inf and differs only for rows in which every key is masked (zero output here, the mean
    of V there).  ``checkpoint_attentions`` is read in ``_attend``."""

    _KEYS = OrderedDict([
        ("num_attention_heads", 32), ("attention_head_size", 32), ("hidden_size", 1024),
        ("attention_dropout_prob", 0.1), ("hidden_dropout_prob", 0.1), ("layernorm_epsilon", 1e-5),
        ("mask_value", -1e4), ("initializer_range", 0.02), ("use_normal_initialization", False),
        ("cross_attention", False), ("causal_mask_size", None), ("pre_layernorm", False), ("post_layernorm", True),
        ("attention_in_fp32", False), ("query_key_layer_scaling", False), ("fp32_residual_addition", False),
        ("fused_softmax", True), ("input_layer", True), ("output_layer", True),
        ("full_attention_mask_and_cross_states", False), ("_scale_qkv_fan_out", False), ("layer_idx", 0),
        ("rotary_dim", None), ("rotary_emb_base", None), ("gpt_neox_type_rotary", False),
        ("parallel_attn_output", False), ("window_size", None), ("use_qkv_bias", True),
        ("use_attn_dense_bias", True), ("scale_attention_scores", True), ("scale_attn_by_layer_idx", False),
        ("single_pre_layernorm", False), ("num_positions", 1024),
    ])

    def __init__(self, *args, **kwargs):
        super().__init__()
        parse_args(args, kwargs, self._KEYS, self)
        if not (self.pre_layernorm or self.post_layernorm or self.single_pre_layernorm or self.output_layer):
            raise DistTransformerConfigError(
                "DistributedAttentionLayer must include LayerNorm if called as part of another distributed module.")
        if self.fp32_residual_addition and not (self.pre_layernorm or self.single_pre_layernorm):
            raise DistTransformerConfigError("fp32 residual addition requires pre-layernorm to be true.")
        H, d = self.hidden_size, self.attention_head_size
        self.local_attention_heads = get_local_channels(self.num_attention_heads)
        self.local_attention_size = la = self.local_attention_heads * d
        self.all_local_attention_sizes = [d * get_local_channels(self.num_attention_heads, r) for r in range(tp_size())]
        if self.local_attention_heads == 0:
            raise DistTransformerConfigError("fewer attention heads than tensor-parallel ranks")
        if self.rotary_dim is not None and self.rotary_emb_base is None:
            self.rotary_emb_base = 10000
        dtype = _param_dtype()
        self.memory_layout = _memory_layout()
        self.hidden_split = [get_local_channels(H, r) for r in range(tp_size())]
        Hl = get_local_channels(H)
        if self.memory_layout:
            A = self.num_attention_heads * d
            with parameter_creation_scope(self, scaled_batch=True, dtype=dtype):
                with initialize_with_input_partition(self):
                    self.qkv_weight = nn.Parameter(torch.empty(3 * A, Hl))      # every output feature, H/t inputs
                    self.dense_weight = nn.Parameter(torch.empty(H, la))
                with initialize_with_output_partition(self):
                    self.qkv_bias = nn.Parameter(torch.zeros(3 * la)) if self.use_qkv_bias else None
                    self.dense_bias = nn.Parameter(torch.zeros(Hl)) if self.use_attn_dense_bias else None
                self._init_weights()
            from .layer_norm import DistributedLayerNorm
            with parameter_creation_scope(self, scaled_batch=True, dtype=dtype):
                self.pre_layernorm_mod = DistributedLayerNorm(H, self.layernorm_epsilon) if self.pre_layernorm else None
                self.layernorm = DistributedLayerNorm(H, self.layernorm_epsilon) if self.post_layernorm else None
            self.weight_split_shapes = {"qkv_weight": self.hidden_split, "dense_weight": self.all_local_attention_sizes}
        else:
            with parameter_creation_scope(self, scaled_batch=True, dtype=dtype):
                with initialize_with_output_partition(self):
                    self.qkv_weight = nn.Parameter(torch.empty(3 * la, H))
                    self.qkv_bias = nn.Parameter(torch.zeros(3 * la)) if self.use_qkv_bias else None
                with initialize_with_input_partition(self, exclude_from_dist=["dense_bias"]):
                    self.dense_weight = nn.Parameter(torch.empty(H, la))
                    # the row-parallel bias lives on tp_rank 0 only and is added before the reduce-scatter
                    self.dense_bias = nn.Parameter(torch.zeros(H)) if (self.use_attn_dense_bias and tp_rank() == 0) else None
                self._init_weights()
            with parameter_creation_scope(self, scaled_batch=False, dtype=dtype):
                mixed = bool(self.fp32_residual_addition)
                self.pre_layernorm_mod = FusedLayerNorm(H, self.layernorm_epsilon, mixed_dtype=mixed) \
                    if self.pre_layernorm else None
                self.layernorm = FusedLayerNorm(H, self.layernorm_epsilon) if self.post_layernorm else None
            self.weight_split_shapes = {"qkv_weight": self.all_local_attention_sizes,
                                        "dense_weight": self.all_local_attention_sizes}
        _NamedSlices.install(self, [("qkv_weight", "weight"), ("qkv_bias", "bias")], ("query", "key", "value"))
        self._install_name_hooks()

    # parameter names in checkpoints: dense.weight / dense.bias / pre_layernorm.* (reference names)
    def _install_name_hooks(self):
        ren = {"dense_weight": "dense.weight", "dense_bias": "dense.bias",
               "pre_layernorm_mod.weight": "pre_layernorm.weight", "pre_layernorm_mod.bias": "pre_layernorm.bias"}

        def save_hook(mod, sd, prefix, meta):
            for a, b in ren.items():
                if prefix + a in sd:
                    sd[prefix + b] = sd.pop(prefix + a)
            return sd

        def load_hook(sd, prefix, meta, strict, missing, unexpected, errors):
            for a, b in ren.items():
                if prefix + b in sd:
                    sd[prefix + a] = sd.pop(prefix + b)

        self._register_state_dict_hook(save_hook)
        self._register_load_state_dict_pre_hook(load_hook)

    def state_dict_layout(self):
        sz = self.all_local_attention_sizes
        if self.memory_layout:
            hs = self.hidden_split
            out = {f"{n}.weight": (1, hs) for n in ("query", "key", "value")}
            out.update({f"{n}.bias": (0, sz) for n in ("query", "key", "value")})
            out.update({"dense.weight": (1, sz), "dense.bias": (0, hs), "pre_layernorm.weight": (0, hs),
                        "pre_layernorm.bias": (0, hs), "layernorm.weight": (0, hs), "layernorm.bias": (0, hs)})
            return out
        out = {f"{n}.weight": (0, sz) for n in ("query", "key", "value")}
        out.update({f"{n}.bias": (0, sz) for n in ("query", "key", "value")})
        out["dense.weight"] = (1, sz)
        out["dense.bias"] = ("one_rank", None)
        return out

    def _init_weights(self):
        std = self.initializer_range
        if self.use_normal_initialization:
            nn.init.normal_(self.qkv_weight, 0.0, std)
            nn.init.normal_(self.dense_weight, 0.0, std)
            if self.qkv_bias is not None:
                nn.init.zeros_(self.qkv_bias)
            if self.dense_bias is not None:
                nn.init.zeros_(self.dense_bias)
        else:
            # nn.Linear default (kaiming uniform with a=sqrt(5)) seen at full, unsplit size
            la, H = self.local_attention_size, self.hidden_size
            blk = self.qkv_weight.shape[0] // 3
            for i in range(3):
                nn.init.kaiming_uniform_(self.qkv_weight.data[i * blk:(i + 1) * blk], a=math.sqrt(5))
            nn.init.kaiming_uniform_(self.dense_weight, a=math.sqrt(5))
            if self.qkv_bias is not None:
                bound = 1.0 / math.sqrt(H)
                nn.init.uniform_(self.qkv_bias, -bound, bound)
            if self.dense_bias is not None:
                bound = 1.0 / math.sqrt(self.num_attention_heads * self.attention_head_size)
                nn.init.uniform_(self.dense_bias, -bound, bound)

    def sdp_first_weight(self):
        """The weight this module's forward multiplies first (sharded data parallelism gathers the unit from
        inside that GEMM)."""
        return self.qkv_weight

    # ---- pieces (also used by DistributedTransformerLayer for the parallel attention+MLP form)
    def attention_core(self, qkv, B, S, attention_mask):
        """qkv [T, 3*local] for all T = B*S tokens of the tp group's batch -> context [T, local]."""
        la, hl, d = self.local_attention_size, self.local_attention_heads, self.attention_head_size
        rot = self.rotary_dim or 0
        style = 0 if not rot else (2 if self.gpt_neox_type_rotary else 1)
        cos = sin = None
        if style:
            cos, sin = A.rotary_tables(max(S, self.num_positions or S), rot, float(self.rotary_emb_base), qkv.device)
        q, k, v = A.qkv_split_rotary(qkv, B, S, hl, d, rot, style, cos, sin, 0)
        causal = self.causal_mask_size is not None
        return self._attend(q, k, v, B, S, causal, attention_mask)

    def _attend(self, q, k, v, B, S, causal, attention_mask):
        d, la = self.attention_head_size, self.local_attention_size
        bias = None
        if isinstance(attention_mask, tuple):        # (boolean mask or None, additive per-head bias): see _full_mask
            attention_mask, bias = attention_mask
        scale = 1.0 / math.sqrt(d) if self.scale_attention_scores else 1.0
        if self.scale_attn_by_layer_idx and not self.query_key_layer_scaling:
            scale = scale / float(self.layer_idx + 1)
        mask = attention_mask if (attention_mask is not None and not causal) else None
        # query_key_layer_scaling / attention_in_fp32: keep the stored 16-bit scores small - the score GEMM applies
        # 1/sqrt(d) (and 1/(layer+1)); the softmax, which works in fp32, undoes the layer factor.  The accumulation
        # itself is always fp32 (TMEM).
        gemm_scale = 1.0
        if self.query_key_layer_scaling or self.attention_in_fp32:
            gemm_scale = 1.0 / math.sqrt(d) if self.scale_attention_scores else 1.0
            if self.query_key_layer_scaling and self.scale_attn_by_layer_idx:
                gemm_scale /= float(self.layer_idx + 1)
        def core(q, k, v):
            return A.core_attention(q, k, v, causal=causal, mask=mask, scale=scale,
                                    dropout_p=self.attention_dropout_prob, training=self.training,
                                    window=self.window_size if causal else None, gemm_scale=gemm_scale,
                                    token_major=True, bias=bias)

        # checkpoint_attentions (reference torch/nn/transformer.py:1487-1493): do not keep the scores / probabilities
        # for backward.  The fused kernel never stores them (it keeps q, k, v, the output and one log-sum-exp per
        # row and recomputes the probabilities tile by tile), so the key only changes the library fallback, which
        # would otherwise hold a [B, heads, Sq, Sk] tensor.
        if _checkpoint_attentions() and torch.is_grad_enabled() and \
                (bias is not None or not A.flash_applicable(q, k, v)) and (q.requires_grad or k.requires_grad):
            from torch.utils.checkpoint import checkpoint
            ctx = checkpoint(core, q, k, v, use_reentrant=False, preserve_rng_state=True)
        else:
            ctx = core(q, k, v)
        return ctx.reshape(B * S, la)        # [B, S, heads, d] straight from the attention kernel

    def forward(self, inp):
        """inp: (hidden [b, S, H], attention_mask) or (hidden, cross_states, attention_mask)."""
        if self.cross_attention:
            hidden, cross_states, attention_mask = inp
        else:
            hidden, attention_mask = inp
            cross_states = None
        if self.memory_layout:
            return self._forward_memory(hidden, cross_states, attention_mask)
        b, S, H = hidden.shape
        g = _group()
        t = tp_size()
        la, hl = self.local_attention_size, self.local_attention_heads
        x2 = hidden.reshape(b * S, H)
        h = self.pre_layernorm_mod(x2) if self.pre_layernorm_mod is not None else x2
        if h.dtype != self.qkv_weight.dtype:
            h = h.to(self.qkv_weight.dtype)
        mask = _full_mask(attention_mask, b * t, g)
        if cross_states is None:
            qkv = tp_ops.ag_linear(h, self.qkv_weight, self.qkv_bias, None, g)
            ctx2 = self.attention_core(qkv, b * t, S, mask)
        else:
            clen = cross_states.shape[1]
            c2 = cross_states.reshape(-1, H).to(self.qkv_weight.dtype)
            qb = None if self.qkv_bias is None else self.qkv_bias[:la]
            kvb = None if self.qkv_bias is None else self.qkv_bias[la:]
            q2 = tp_ops.ag_linear(h, self.qkv_weight[:la], qb, None, g)
            kv = tp_ops.ag_linear(c2, self.qkv_weight[la:], kvb, None, g)
            q = A.tokens_to_heads(q2.reshape(b * t, S, la), hl)
            k = A.tokens_to_heads(kv[:, :la].reshape(b * t, clen, la), hl)
            v = A.tokens_to_heads(kv[:, la:].reshape(b * t, clen, la), hl)
            ctx2 = self._attend(q, k, v, b * t, S, False, mask)
        out = tp_ops.linear_rs(ctx2, self.dense_weight, self.dense_bias, g)
        out = F.dropout(out, self.hidden_dropout_prob, self.training)
        out = out.to(x2.dtype) + x2
        if self.layernorm is not None:
            out = self.layernorm(out)
        out = out.reshape(b, S, H)
        if self.cross_attention:
            return out, cross_states, attention_mask
        return out, attention_mask


    def _forward_memory(self, hidden, cross_states, attention_mask):
        """hidden [B, S, H/t] (full batch, hidden-sharded) -> same layout.  q / k / v: partial products over this
        rank's input features, reduce-scattered over the output features so that each rank ends up with its own
        heads; the output projection reduce-scatters back to H/t."""
        from .utils import reduce_scatter_for_tp

        B, S, Hl = hidden.shape
        g = _group()
        la, hl, d = self.local_attention_size, self.local_attention_heads, self.attention_head_size
        A = self.num_attention_heads * d
        x2 = hidden.reshape(B * S, Hl)
        h = self.pre_layernorm_mod(x2) if self.pre_layernorm_mod is not None else x2
        if h.dtype != self.qkv_weight.dtype:
            h = h.to(self.qkv_weight.dtype)
        mask = _full_mask(attention_mask, B, g)
        sizes = self.all_local_attention_sizes
        if cross_states is None:
            part = _plain_linear(h, self.qkv_weight).reshape(B * S, 3, A)
            qkv = reduce_scatter_for_tp(part, 2, split_shapes=sizes)                 # [T, 3, la]
            if self.qkv_bias is not None:
                qkv = qkv + self.qkv_bias.reshape(3, la)
            ctx2 = self.attention_core(qkv.reshape(B * S, 3 * la), B, S, mask)
        else:
            Sc = cross_states.shape[1]
            c2 = cross_states.reshape(-1, Hl).to(self.qkv_weight.dtype)
            q2 = reduce_scatter_for_tp(_plain_linear(h, self.qkv_weight[:A]), 1, split_shapes=sizes)
            kv = reduce_scatter_for_tp(_plain_linear(c2, self.qkv_weight[A:]).reshape(-1, 2, A), 2, split_shapes=sizes)
            if self.qkv_bias is not None:
                q2 = q2 + self.qkv_bias[:la]
                kv = kv + self.qkv_bias[la:].reshape(2, la)
            q = A_ops.tokens_to_heads(q2.reshape(B, S, la), hl)
            k = A_ops.tokens_to_heads(kv[:, 0].reshape(B, Sc, la), hl)
            v = A_ops.tokens_to_heads(kv[:, 1].reshape(B, Sc, la), hl)
            ctx2 = self._attend(q, k, v, B, S, False, mask)
        out = reduce_scatter_for_tp(_plain_linear(ctx2, self.dense_weight), 1, split_shapes=self.hidden_split)
        if self.dense_bias is not None:
            out = out + self.dense_bias
        out = F.dropout(out, self.hidden_dropout_prob, self.training)
        out = out.to(x2.dtype) + x2
        if self.layernorm is not None:
            out = self.layernorm(out)
        out = out.reshape(B, S, Hl)
        if self.cross_attention:
            return out, cross_states, attention_mask
        return out, attention_mask


def _full_mask(attention_mask, full_batch, group):
    """Bring a per-rank padding mask to the tp group's full batch as a uint8 'masked-out' tensor
    [B, 1, 1|Sq, Sk] -> expanded lazily by the softmax wrapper.  Accepts 0/1 keep-masks [b, S] and
    additive masks [b, 1, 1, S] (negative = masked), as the reference does."""
    if attention_mask is None:
        return None
    m = attention_mask
    if m.dim() == 4 and m.is_floating_point() and m.shape[1] > 1:
        # additive bias with a head dimension (the reference's T5 hooks add the relative position bias to the mask,
        # torch/nn/huggingface/t5.py:62-82): this rank's heads, kept additive; entries at or below -1e4 also mask
        heads = m.shape[1]
        local = get_local_channels(heads)
        start = sum(get_local_channels(heads, r) for r in range(tp_rank()))
        b_local = m.narrow(1, start, local)
        if group is not None and b_local.shape[0] not in (1, full_batch):
            b_local = tp_ops.gather_tokens(b_local.contiguous(), group)
        return (None, b_local)
    if m.dim() == 2:
        m = (m == 0)
    else:
        m = (m < 0) if m.is_floating_point() else (m == 0)
        m = m.reshape(m.shape[0], -1, m.shape[-1])[:, -1, :] if m.dim() == 4 and m.shape[2] == 1 else m
    m = m.to(torch.uint8)
    if group is not None and m.shape[0] != full_batch:
        m = tp_ops.gather_tokens(m.contiguous(), group)
    if m.dim() == 2:
        m = m[:, None, None, :]
    return m


class DistributedTransformerOutputLayer(DistributedModule):
    """The MLP block: pre-LN -> dense1 (column-parallel, bias+GELU fused in the GEMM epilogue) ->
    dense2 (row-parallel) -> dropout -> residual -> post-LN."""

    _KEYS = OrderedDict([
        ("hidden_size", 1024), ("intermediate_size", 4096), ("hidden_dropout_prob", 0.1), ("activation", "gelu"),
        ("layernorm_epsilon", 1e-5), ("initializer_range", 0.02), ("use_normal_initialization", False),
        ("pre_layernorm", False), ("post_layernorm", True), ("fp32_residual_addition", False),
        ("fused_bias_gelu", False), ("input_layer", True), ("output_layer", True), ("parallel_attn_output", False),
        ("single_pre_layernorm", False),
    ])

    def __init__(self, *args, **kwargs):
        super().__init__()
        parse_args(args, kwargs, self._KEYS, self)
        H, I = self.hidden_size, self.intermediate_size
        self.local_intermediate_size = li = get_local_channels(I)
        self.all_local_intermediate_sizes = [get_local_channels(I, r) for r in range(tp_size())]
        dtype = _param_dtype()
        self.memory_layout = _memory_layout()
        self.hidden_split = [get_local_channels(H, r) for r in range(tp_size())]
        if self.memory_layout:
            Hl = get_local_channels(H)
            with parameter_creation_scope(self, scaled_batch=True, dtype=dtype):
                with initialize_with_input_partition(self):
                    self.dense1_weight = nn.Parameter(torch.empty(I, Hl))        # every output feature, H/t inputs
                    self.dense2_weight = nn.Parameter(torch.empty(H, li))
                with initialize_with_output_partition(self):
                    self.dense1_bias = nn.Parameter(torch.zeros(li))
                    self.dense2_bias = nn.Parameter(torch.zeros(Hl))
                self._init_weights()
            from .layer_norm import DistributedLayerNorm
            with parameter_creation_scope(self, scaled_batch=True, dtype=dtype):
                self.pre_layernorm_mod = DistributedLayerNorm(H, self.layernorm_epsilon) if self.pre_layernorm else None
                self.layernorm = DistributedLayerNorm(H, self.layernorm_epsilon) if self.post_layernorm else None
            self.weight_split_shapes = {"dense1_weight": self.hidden_split,
                                        "dense2_weight": self.all_local_intermediate_sizes}
        else:
            with parameter_creation_scope(self, scaled_batch=True, dtype=dtype):
                with initialize_with_output_partition(self):
                    self.dense1_weight = nn.Parameter(torch.empty(li, H))
                    self.dense1_bias = nn.Parameter(torch.zeros(li))
                with initialize_with_input_partition(self, exclude_from_dist=["dense2_bias"]):
                    self.dense2_weight = nn.Parameter(torch.empty(H, li))
                    self.dense2_bias = nn.Parameter(torch.zeros(H)) if tp_rank() == 0 else None
                self._init_weights()
            with parameter_creation_scope(self, scaled_batch=False, dtype=dtype):
                mixed = bool(self.fp32_residual_addition)
                self.pre_layernorm_mod = FusedLayerNorm(H, self.layernorm_epsilon, mixed_dtype=mixed) \
                    if self.pre_layernorm else None
                self.layernorm = FusedLayerNorm(H, self.layernorm_epsilon) if self.post_layernorm else None
            self.weight_split_shapes = {"dense1_weight": self.all_local_intermediate_sizes,
                                        "dense2_weight": self.all_local_intermediate_sizes}
        ren = {"dense1_weight": "dense1.weight", "dense1_bias": "dense1.bias", "dense2_weight": "dense2.weight",
               "dense2_bias": "dense2.bias", "pre_layernorm_mod.weight": "pre_layernorm.weight",
               "pre_layernorm_mod.bias": "pre_layernorm.bias"}

        def save_hook(mod, sd, prefix, meta):
            for a, b in ren.items():
                if prefix + a in sd:
                    sd[prefix + b] = sd.pop(prefix + a)
            return sd

        def load_hook(sd, prefix, meta, strict, missing, unexpected, errors):
            for a, b in ren.items():
                if prefix + b in sd:
                    sd[prefix + a] = sd.pop(prefix + b)

        self._register_state_dict_hook(save_hook)
        self._register_load_state_dict_pre_hook(load_hook)

    def _init_weights(self):
        H, I = self.hidden_size, self.intermediate_size
        if self.use_normal_initialization:
            nn.init.normal_(self.dense1_weight, 0.0, self.initializer_range)
            nn.init.normal_(self.dense2_weight, 0.0, self.initializer_range)
            nn.init.zeros_(self.dense1_bias)
            if self.dense2_bias is not None:
                nn.init.zeros_(self.dense2_bias)
        else:
            nn.init.kaiming_uniform_(self.dense1_weight, a=math.sqrt(5))
            nn.init.kaiming_uniform_(self.dense2_weight, a=math.sqrt(5))
            nn.init.uniform_(self.dense1_bias, -1.0 / math.sqrt(H), 1.0 / math.sqrt(H))
            if self.dense2_bias is not None:
                nn.init.uniform_(self.dense2_bias, -1.0 / math.sqrt(I), 1.0 / math.sqrt(I))

    def state_dict_layout(self):
        sz = self.all_local_intermediate_sizes
        if self.memory_layout:
            hs = self.hidden_split
            return {"dense1.weight": (1, hs), "dense1.bias": (0, sz), "dense2.weight": (1, sz), "dense2.bias": (0, hs),
                    "pre_layernorm.weight": (0, hs), "pre_layernorm.bias": (0, hs), "layernorm.weight": (0, hs),
                    "layernorm.bias": (0, hs)}
        return {"dense1.weight": (0, sz), "dense1.bias": (0, sz), "dense2.weight": (1, sz),
                "dense2.bias": ("one_rank", None)}

    def sdp_first_weight(self):
        return self.dense1_weight

    def _act_other(self, h):
        """Activations that do not ride a GEMM epilogue (relu, or any GELU when SMP_USE_HF_GELU=1)."""
        if self.activation == "relu":
            return F.relu(h)
        if self.activation == "gelu":
            return F.gelu(h)
        if self.activation in _TANH_ACTS:
            return F.gelu(h, approximate="tanh")
        return getattr(F, self.activation)(h)

    def _forward_memory(self, hidden):
        """[B, S, H/t] -> [B, S, H/t]: both projections produce partial sums over this rank's input features and are
        reduce-scattered over their output features."""
        from .utils import reduce_scatter_for_tp

        shape = hidden.shape
        x2 = hidden.reshape(-1, shape[-1])
        h = self.pre_layernorm_mod(x2) if self.pre_layernorm_mod is not None else x2
        if h.dtype != self.dense1_weight.dtype:
            h = h.to(self.dense1_weight.dtype)
        inter = reduce_scatter_for_tp(_plain_linear(h, self.dense1_weight), 1,
                                      split_shapes=self.all_local_intermediate_sizes) + self.dense1_bias
        inter = self._act_other(inter)
        out = reduce_scatter_for_tp(_plain_linear(inter, self.dense2_weight), 1, split_shapes=self.hidden_split)
        out = out + self.dense2_bias
        out = F.dropout(out, self.hidden_dropout_prob, self.training)
        out = out.to(x2.dtype) + x2
        if self.layernorm is not None:
            out = self.layernorm(out)
        return out.reshape(shape)

    def forward(self, hidden):
        if self.memory_layout:
            return self._forward_memory(hidden)
        shape = hidden.shape
        H = shape[-1]
        g = _group()
        x2 = hidden.reshape(-1, H)
        h = self.pre_layernorm_mod(x2) if self.pre_layernorm_mod is not None else x2
        if h.dtype != self.dense1_weight.dtype:
            h = h.to(self.dense1_weight.dtype)
        mode = _gelu_mode(self.activation)
        if mode is not None and tp_ops.fused_mlp_applicable(h, self.dense1_weight, g):
            out = tp_ops.FusedMLP.apply(h, self.dense1_weight, self.dense1_bias, self.dense2_weight, self.dense2_bias, mode)
        else:
            inter = tp_ops.ag_linear(h, self.dense1_weight, self.dense1_bias, mode, g)
            if mode is None:
                inter = self._act_other(inter)
            out = tp_ops.linear_rs(inter, self.dense2_weight, self.dense2_bias, g)
        out = F.dropout(out, self.hidden_dropout_prob, self.training)
        out = out.to(x2.dtype) + x2
        if self.layernorm is not None:
            out = self.layernorm(out)
        return out.reshape(shape)


class DistributedTransformerLayer(DistributedModule):
    _KEYS = OrderedDict([
        ("num_attention_heads", 32), ("attention_head_size", 32), ("hidden_size", 1024), ("intermediate_size", 4096),
        ("attention_dropout_prob", 0.1), ("hidden_dropout_prob", 0.1), ("activation", "gelu"),
        ("layernorm_epsilon", 1e-5), ("initializer_range", 0.02), ("use_normal_initialization", False),
        ("causal_mask_size", None), ("add_cross_attention", False), ("pre_layernorm", False), ("post_layernorm", True),
        ("attention_in_fp32", False), ("query_key_layer_scaling", False), ("fp32_residual_addition", False),
        ("fused_softmax", True), ("fused_bias_gelu", False), ("input_layer", True), ("output_layer", True),
        ("mask_value", -1e4), ("layer_idx", 0), ("rotary_dim", None), ("rotary_emb_base", None),
        ("gpt_neox_type_rotary", False), ("parallel_attn_output", False), ("window_size", None),
        ("use_qkv_bias", True), ("use_attn_dense_bias", True), ("single_pre_layernorm", False),
        ("scale_attention_scores", True), ("scale_attn_by_layer_idx", False), ("num_positions", 1024),
        ("_scale_qkv_fan_out", False), ("full_attention_mask_and_cross_states", False),
    ])

    def __init__(self, *args, **kwargs):
        super().__init__()
        cfg = parse_args(args, kwargs, self._KEYS, self)
        if self.parallel_attn_output and self.add_cross_attention:
            raise DistTransformerConfigError("parallel_attn_output is not supported with cross attention")
        self.memory_layout = _memory_layout()
        self.hidden_split = [get_local_channels(self.hidden_size, r) for r in range(tp_size())]
        if self.memory_layout and self.parallel_attn_output:
            raise DistTransformerConfigError("optimize='memory' does not support parallel_attn_output (GPT-J / "
                                             "GPT-NeoX style blocks); use optimize='speed'")
        # GPT-J: one LayerNorm feeds both branches (single_pre_layernorm); GPT-NeoX: each branch has its own
        attn_cfg = update_copy_dict(cfg, cross_attention=False)
        out_cfg = dict(cfg)
        if self.parallel_attn_output and self.single_pre_layernorm:
            out_cfg = update_copy_dict(cfg, pre_layernorm=False, post_layernorm=False)
            attn_cfg = update_copy_dict(attn_cfg, pre_layernorm=True, post_layernorm=False)
        self.attention = DistributedAttentionLayer(**attn_cfg)
        self.cross_attention = DistributedAttentionLayer(**update_copy_dict(cfg, cross_attention=True)) \
            if self.add_cross_attention else None
        self.output = DistributedTransformerOutputLayer(**out_cfg)

    def forward(self, inp):
        if self.add_cross_attention:
            hidden, attention_mask, cross_states, cross_mask = inp
        else:
            hidden, attention_mask = inp[0], inp[1]
            cross_states = cross_mask = None
        if self.memory_layout:
            from .utils import scatter_and_merge_for_tp
            if self.input_layer:      # [local batch, S, H] -> [full batch, S, H/t]  (all-to-all, reference :1429-1434)
                hidden = scatter_and_merge_for_tp(hidden, split_dim=2, merge_dim=0, split_shapes=self.hidden_split)
            sharded_cross = cross_states
            if cross_states is not None and (self.input_layer or cross_states.shape[-1] == self.hidden_size):
                sharded_cross = scatter_and_merge_for_tp(cross_states, split_dim=2, merge_dim=0,
                                                         split_shapes=self.hidden_split)
            hidden, _ = self.attention((hidden, attention_mask))
            if self.cross_attention is not None:
                hidden, _, _ = self.cross_attention((hidden, sharded_cross, cross_mask))
            hidden = self.output(hidden)
            if self.output_layer:     # back to [local batch, S, H]
                hidden = scatter_and_merge_for_tp(hidden, split_dim=0, merge_dim=2, merge_shapes=self.hidden_split)
        elif self.parallel_attn_output:
            hidden = self._parallel_forward(hidden, attention_mask)
        else:
            hidden, _ = self.attention((hidden, attention_mask))
            if self.cross_attention is not None:
                hidden, _, _ = self.cross_attention((hidden, cross_states, cross_mask))
            hidden = self.output(hidden)
        if self.add_cross_attention:
            return hidden, attention_mask, cross_states, cross_mask
        return hidden, attention_mask

    def sdp_first_weight(self):
        return self.attention.qkv_weight

    def _parallel_forward(self, hidden, attention_mask):
        """x + attn(ln_a(x)) + mlp(ln_m(x)).  The attention-dense partial product rides into the MLP's
        second GEMM as an epilogue residual so a single reduce-scatter carries both branches; when the
        LayerNorm is shared (GPT-J) one all-gather - fused into the QKV GEMM - feeds both branches."""
        b, S, H = hidden.shape
        g, t = _group(), tp_size()
        att, out = self.attention, self.output
        x2 = hidden.reshape(b * S, H)
        wdtype = att.qkv_weight.dtype
        ha = att.pre_layernorm_mod(x2) if att.pre_layernorm_mod is not None else x2
        ha = ha.to(wdtype) if ha.dtype != wdtype else ha
        mask = _full_mask(attention_mask, b * t, g)
        mode = _gelu_mode(out.activation)
        shared_ln = out.pre_layernorm_mod is None
        if mode is not None and tp_ops.fused_mlp_applicable(ha, out.dense1_weight, g):
            # one rank: QKV first (sharded data parallelism gathers the unit inside that GEMM), then the whole MLP
            # as one fused function with the attention branch added by its last GEMM's epilogue
            qkv = tp_ops.ag_linear(ha, att.qkv_weight, att.qkv_bias, None, g)
            if shared_ln:
                hm = ha
            else:
                hm = out.pre_layernorm_mod(x2)
                hm = hm.to(wdtype) if hm.dtype != wdtype else hm
            ctx2 = att.attention_core(qkv, b * t, S, mask)
            p1 = tp_ops.RowParallelLinear.apply(ctx2, att.dense_weight, att.dense_bias, g, False)
            y = tp_ops.FusedMLP.apply(hm, out.dense1_weight, out.dense1_bias, out.dense2_weight, out.dense2_bias, mode, p1)
            y = F.dropout(y, self.hidden_dropout_prob, self.training)
            y = y.to(x2.dtype) + x2
            return y.reshape(b, S, H)
        if shared_ln:
            qkv, xg = tp_ops.ag_linear(ha, att.qkv_weight, att.qkv_bias, None, g, want_gathered=True)
            inter = tp_ops.ColumnParallelLinear.apply(xg, out.dense1_weight, out.dense1_bias, mode, g, True)
        else:
            qkv = tp_ops.ag_linear(ha, att.qkv_weight, att.qkv_bias, None, g)
            hm = out.pre_layernorm_mod(x2)
            hm = hm.to(wdtype) if hm.dtype != wdtype else hm
            inter = tp_ops.ag_linear(hm, out.dense1_weight, out.dense1_bias, mode, g)
        if mode is None:
            inter = out._act_other(inter)
        ctx2 = att.attention_core(qkv, b * t, S, mask)
        p1 = tp_ops.RowParallelLinear.apply(ctx2, att.dense_weight, att.dense_bias, g, False)   # partial, not reduced
        y = tp_ops.linear_rs(inter, out.dense2_weight, out.dense2_bias, g, residual_full=p1)
        y = F.dropout(y, self.hidden_dropout_prob, self.training)
        y = y.to(x2.dtype) + x2
        return y.reshape(b, S, H)


class DistributedTransformer(DistributedModule):
    _KEYS = OrderedDict([("num_layers", 12)] + list(DistributedTransformerLayer._KEYS.items())
                        + [("attention_layers_type", None), ("_output_full_batch", False)])

    def __init__(self, *args, **kwargs):
        super().__init__()
        cfg = parse_args(args, kwargs, self._KEYS, self)
        layers = []
        for i in range(self.num_layers):
            lcfg = {k: v for k, v in cfg.items() if k in DistributedTransformerLayer._KEYS}
            lcfg.update(layer_idx=i, input_layer=(i == 0), output_layer=(i == self.num_layers - 1))
            if self.attention_layers_type is not None:          # GPT-Neo: alternate global / local attention
                kind = self.attention_layers_type[i % len(self.attention_layers_type)]
                lcfg["window_size"] = self.window_size if kind == "local" else None
            layers.append(DistributedTransformerLayer(**lcfg))
        self.seq_layers = nn.Sequential(*layers)

    def forward(self, inp):
        return self.seq_layers(inp)


class DistributedTransformerLMHead(DistributedModule):
    """Embeddings + transformer stack + (tied) LM head + loss.

    ``forward((input_ids, attention_mask, token_type_ids, position_ids, labels))`` returns
    ``(loss, logits)`` when labels are given, else logits - the reference's calling convention
    (reference torch/nn/transformer.py:368-548)."""

    _KEYS = OrderedDict([
        ("num_layers", 12), ("num_attention_heads", 32), ("attention_head_size", 32), ("hidden_size", 1024),
        ("intermediate_size", 4096), ("vocab_size", 30522), ("num_positions", 1024), ("attention_dropout_prob", 0.1),
        ("hidden_dropout_prob", 0.1), ("embedding_dropout_prob", 0.1), ("activation", "gelu"),
        ("layernorm_epsilon", 1e-5), ("mask_value", -1e4), ("num_token_types", 0), ("causal_mask_size", None),
        ("add_cross_attention", False), ("add_lm_head", True), ("initializer_range", 0.02),
        ("use_normal_initialization", False), ("pre_layernorm", False), ("post_layernorm", True),
        ("attention_in_fp32", False), ("query_key_layer_scaling", False), ("fp32_residual_addition", False),
        ("fused_softmax", True), ("fused_bias_gelu", False), ("distribute_embedding", False),
        ("_scale_qkv_fan_out", False), ("_precision_test", False), ("rotary_dim", None), ("rotary_emb_base", None),
        ("gpt_neox_type_rotary", False), ("use_positional_embedding", True), ("parallel_attn_output", False),
        ("use_lm_head_bias", False), ("attention_layers_type", None), ("use_qkv_bias", True),
        ("use_attn_dense_bias", True), ("window_size", None), ("final_layernorm", False),
        ("tie_input_output_embedding", True), ("single_pre_layernorm", False), ("scale_attention_scores", True),
        ("scale_attn_by_layer_idx", False),
    ])

    def __init__(self, *args, **kwargs):
        super().__init__()
        cfg = parse_args(args, kwargs, self._KEYS, self)
        if not (self.pre_layernorm or self.single_pre_layernorm or self.post_layernorm):
            raise DistTransformerConfigError(
                "At least one of pre_layernorm, single_pre_layernorm and post_layernorm must be True.")
        if self.causal_mask_size is not None and self.num_positions > self.causal_mask_size:
            raise DistTransformerConfigError(
                f"If causal mask size ({self.causal_mask_size}) is provided, it must be larger than or equal to the "
                f"number of positions ({self.num_positions}).")
        dtype = _param_dtype()
        s = _state()
        self.prescaled_batch = bool(getattr(s, "initialized", False) and s.cfg.prescaled_batch)
        if self.distribute_embedding:
            from .embedding import DistributedEmbedding
            self.word_embedding = DistributedEmbedding(self.vocab_size, self.hidden_size,
                                                       initializer_range=self.initializer_range, vocab_parallel=True,
                                                       _skip_allgather=self.prescaled_batch, dtype=dtype)
        else:
            with parameter_creation_scope(self, scaled_batch=False, dtype=dtype,
                                          use_normal=self.use_normal_initialization,
                                          initializer_range=self.initializer_range):
                self.word_embedding = nn.Embedding(self.vocab_size, self.hidden_size)
        with parameter_creation_scope(self, scaled_batch=False, dtype=dtype, use_normal=self.use_normal_initialization,
                                      initializer_range=self.initializer_range):
            if self.use_positional_embedding:
                self.position_embedding = nn.Embedding(self.num_positions, self.hidden_size)
            if self.num_token_types > 0:
                self.token_type_embedding = nn.Embedding(self.num_token_types, self.hidden_size)
        self.dropout = nn.Dropout(self.embedding_dropout_prob)
        tcfg = {k: v for k, v in cfg.items() if k in DistributedTransformer._KEYS}
        self.transformer = DistributedTransformer(**tcfg)
        if self.final_layernorm:
            with parameter_creation_scope(self, scaled_batch=False, dtype=dtype):
                self.layernorm = FusedLayerNorm(self.hidden_size, self.layernorm_epsilon)
        if self.add_lm_head:
            if self.tie_input_output_embedding:
                self.lm_head_weight = None
            else:
                with parameter_creation_scope(self, scaled_batch=self.distribute_embedding, dtype=dtype):
                    rows = self.word_embedding.weight.shape[0]
                    w = nn.Parameter(torch.empty(rows, self.hidden_size))
                    nn.init.normal_(w, 0.0, self.initializer_range)
                    if self.distribute_embedding:
                        from .utils import mark_distributed
                        mark_distributed(w, 0)
                    self.lm_head_weight = w
            with parameter_creation_scope(self, scaled_batch=self.distribute_embedding, dtype=dtype):
                self.lm_head_bias = nn.Parameter(torch.zeros(self.word_embedding.weight.shape[0])) \
                    if self.use_lm_head_bias else None

        ren = {"lm_head_weight": "lm_head.weight", "lm_head_bias": "lm_head.bias"}

        def save_hook(mod, sd, prefix, meta):
            for a, b in ren.items():
                if prefix + a in sd:
                    sd[prefix + b] = sd.pop(prefix + a)
            if mod.add_lm_head and mod.tie_input_output_embedding and prefix + "word_embedding.weight" in sd:
                sd[prefix + "lm_head.weight"] = sd[prefix + "word_embedding.weight"]
            return sd

        def load_hook(sd, prefix, meta, strict, missing, unexpected, errors):
            for a, b in ren.items():
                if prefix + b in sd:
                    if a == "lm_head_weight" and self.tie_input_output_embedding:
                        sd.pop(prefix + b)
                    else:
                        sd[prefix + a] = sd.pop(prefix + b)

        self._register_state_dict_hook(save_hook)
        self._register_load_state_dict_pre_hook(load_hook)

    def state_dict_layout(self):
        if self.distribute_embedding and self.add_lm_head:
            return {"lm_head.weight": (0, None), "lm_head.bias": (0, None)}
        return {}

    def sdp_directly_used_parameters(self):
        """Parameters of child modules that this module's own forward touches (tied LM head)."""
        if self.add_lm_head and self.tie_input_output_embedding:
            return [self.word_embedding.weight]
        return []

    # ------------------------------------------------------------------
    def _head_weight(self):
        return self.word_embedding.weight if self.tie_input_output_embedding else self.lm_head_weight

    def forward(self, inputs):
        if self.add_cross_attention:
            input_ids, attention_mask, token_type_ids, position_ids, cross_states, cross_mask, labels = inputs
        else:
            input_ids, attention_mask, token_type_ids, position_ids, labels = inputs
            cross_states = cross_mask = None
        t, r = tp_size(), tp_rank()
        g = _group()
        Bin, S = input_ids.shape
        H = self.hidden_size
        dev = input_ids.device
        prescaled = self.prescaled_batch and t > 1

        # ---- which tokens does this rank own?  local batch (default) or a 1/t token slice (prescaled_batch)
        if prescaled:
            T_full = Bin * S
            if T_full % t != 0:
                from ..backend.exceptions import InvalidSeqLenPrescaledBatchError
                raise InvalidSeqLenPrescaledBatchError()
            T_l = T_full // t
            tok = torch.arange(r * T_l, (r + 1) * T_l, device=dev)
            full_B = Bin
        else:
            T_l = Bin * S
            tok = None
            full_B = Bin * t
        pos_local = (torch.arange(T_l, device=dev) + (r * T_l if prescaled else 0)) % S

        # ---- embeddings on local tokens
        if self.distribute_embedding:
            emb = self.word_embedding(input_ids).reshape(T_l, H)
        else:
            ids_l = input_ids.reshape(-1)[tok] if prescaled else input_ids.reshape(-1)
            emb = self.word_embedding(ids_l)
        if self.use_positional_embedding:
            emb = emb + self.position_embedding(pos_local)
        if token_type_ids is not None:
            if self.num_token_types <= 0:
                raise DistTransformerConfigError("token_type_ids provided, but num_token_types was not!")
            tt = token_type_ids.reshape(-1)[tok] if prescaled else token_type_ids.reshape(-1)
            emb = emb + self.token_type_embedding(tt)
        hidden = self.dropout(emb)

        # ---- transformer stack (hidden travels as [local sequences or 1, local tokens, H])
        b_l = Bin if not prescaled else None
        if prescaled:
            # sequences are split across ranks: tell the layers the true (B, S) through a view they can undo
            hidden3 = _TokenShard(hidden, full_B, S)
        else:
            hidden3 = hidden.reshape(b_l, S, H)
        mask = attention_mask
        if self.causal_mask_size is not None:
            mask = None          # the causal kernel ignores padding masks, as the reference's fused path does
        if self.add_cross_attention:
            out = self.transformer((hidden3, mask, cross_states, cross_mask))
        else:
            out = self.transformer((hidden3, mask))
        hidden = out[0].reshape(T_l, H)
        if self.final_layernorm:
            hidden = self.layernorm(hidden)
        if not self.add_lm_head:
            return hidden.reshape(-1, S, H) if not prescaled else hidden

        # ---- LM head and loss
        W = self._head_weight()
        if hidden.dtype != W.dtype:
            hidden = hidden.to(W.dtype)
        if self.distribute_embedding and t > 1:
            logits = tp_ops.column_parallel_linear(hidden, W, self.lm_head_bias, None, g, False)   # [T, V/t]
            vocab_start = self.word_embedding.vocab_start_idx
        else:
            logits = tp_ops.column_parallel_linear(hidden, W, self.lm_head_bias, None, None, True)  # [T_l, V]
            vocab_start = 0
        if labels is None:
            if self.distribute_embedding and t > 1:
                from .utils import allgather_for_tp, get_split_shapes
                logits = allgather_for_tp(logits, 1, get_split_shapes(self.vocab_size))
                logits = tp_ops.NarrowTokens.apply(logits, g)
            return logits.reshape(-1, S, logits.shape[-1]) if not prescaled else logits

        shift = torch.full_like(labels, -100)
        shift[:, :-1] = labels[:, 1:]
        shift = shift.reshape(-1)
        n_valid_per_rank = Bin * (S - 1) if not prescaled else (Bin * (S - 1)) / t
        if self.distribute_embedding and t > 1:
            full_labels = shift if prescaled else tp_ops.gather_tokens(shift.contiguous(), g)
            loss_tok = vocab_parallel_cross_entropy(logits, full_labels, vocab_start, g, -100, inplace_grad=True)
            # every rank computed the full batch's loss; its own tokens define its local objective
            loss = _LocalOfShared.apply(loss_tok, r, t).sum() / n_valid_per_rank
            ret_logits = logits
        else:
            lab_l = shift[tok] if prescaled else shift
            loss_tok = vocab_parallel_cross_entropy(logits, lab_l, 0, None, -100, inplace_grad=True)
            loss = loss_tok.sum() / n_valid_per_rank
            ret_logits = logits.reshape(-1, S, logits.shape[-1]) if not prescaled else logits
        return loss, ret_logits


class _LocalOfShared(torch.autograd.Function):
    """Select this rank's token slice of a per-token loss that every tp rank computed identically.

    Backward scatters the incoming gradient into the full vector and sums it over the tp group, so the
    (shared) upstream computation receives every rank's contribution."""

    @staticmethod
    def forward(ctx, full, r, t):
        ctx.r, ctx.t, ctx.n = r, t, full.shape[0]
        per = full.shape[0] // t
        return full.narrow(0, r * per, per).clone()

    @staticmethod
    def backward(ctx, g):
        import torch.distributed as dist
        full = g.new_zeros(ctx.n)
        per = ctx.n // ctx.t
        full.narrow(0, ctx.r * per, per).copy_(g)
        grp = _group()
        if grp is not None:
            dist.all_reduce(full, group=grp)
        return full, None, None


class _TokenShard:
    """Placeholder for prescaled_batch hidden states (token slices that cut through sequences)."""

    def __new__(cls, hidden, full_B, S):
        # layers only need (b*t, S) to rebuild full sequences after the all-gather: express the local shard as
        # [full_B / t, S, H] when sequences divide evenly, else fall back to a flat single-row view
        T_l, H = hidden.shape
        if T_l % S == 0:
            return hidden.reshape(T_l // S, S, H)
        raise DistTransformerConfigError(
            "prescaled_batch needs batch_size divisible by the tensor-parallel degree in this implementation")
