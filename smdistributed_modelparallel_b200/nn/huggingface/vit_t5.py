"""ViT encoder and T5 block hooks.  As in the reference these exist but are NOT registered automatically
(reference smdistributed/modelparallel/torch/nn/predefined_hooks.py:140-147 leaves T5 commented out pending
relative position bias support; huggingface/vit.py and t5.py ship the hooks only).  They can be registered by
hand with ``smp.tp_register_with_module``.
"""

from __future__ import annotations

from ...backend.exceptions import HFT5ConfigError, HFViTConfigError
from .common import Rule, Translator, gelu_name


def vit_encoder_init_hook(config, *args, **kwargs):
    if config.hidden_size % config.num_attention_heads != 0:
        raise HFViTConfigError("hidden_size must be divisible by num_attention_heads")
    act = config.hidden_act if isinstance(config.hidden_act, str) else "gelu"
    return (), dict(
        num_layers=config.num_hidden_layers, num_attention_heads=config.num_attention_heads,
        attention_head_size=config.hidden_size // config.num_attention_heads, hidden_size=config.hidden_size,
        intermediate_size=config.intermediate_size, attention_dropout_prob=config.attention_probs_dropout_prob,
        hidden_dropout_prob=config.hidden_dropout_prob, activation=gelu_name(act),
        layernorm_epsilon=config.layer_norm_eps, initializer_range=config.initializer_range,
        use_normal_initialization=True, causal_mask_size=None, pre_layernorm=True, post_layernorm=False,
        use_qkv_bias=getattr(config, "qkv_bias", True))


def vit_encoder_forward_hook(hidden_states, head_mask=None, output_attentions=False, output_hidden_states=False,
                             return_dict=True, **extra):
    if output_attentions or output_hidden_states:
        raise HFViTConfigError("output_attentions / output_hidden_states are not supported")
    return ((hidden_states, None),), {}


def vit_encoder_return_hook(outputs):
    from transformers.modeling_outputs import BaseModelOutput

    return BaseModelOutput(last_hidden_state=outputs[0])


def vit_translator(config=None):
    L = "seq_layers.{i}"
    rules = []
    for s in ("weight", "bias"):
        for n in ("query", "key", "value"):
            rules.append(Rule(f"layer.{{i}}.attention.attention.{n}.{s}", f"{L}.attention.{n}.{s}"))
        rules.append(Rule(f"layer.{{i}}.attention.output.dense.{s}", f"{L}.attention.dense.{s}"))
        rules.append(Rule(f"layer.{{i}}.layernorm_before.{s}", f"{L}.attention.pre_layernorm.{s}"))
        rules.append(Rule(f"layer.{{i}}.layernorm_after.{s}", f"{L}.output.pre_layernorm.{s}"))
        rules.append(Rule(f"layer.{{i}}.intermediate.dense.{s}", f"{L}.output.dense1.{s}"))
        rules.append(Rule(f"layer.{{i}}.output.dense.{s}", f"{L}.output.dense2.{s}"))
    return Translator(rules, (), config)


def t5_block_init_hook(config, has_relative_attention_bias=False, *args, **kwargs):
    """``T5Block(config, has_relative_attention_bias)`` -> DistributedTransformerLayer arguments (reference
    torch/nn/huggingface/t5.py:11-33).  The block that owns the relative-attention-bias table (the first of each
    stack) stays a HuggingFace module: returning None tells the registry not to distribute it."""
    if has_relative_attention_bias:
        return None
    if config.d_kv * config.num_heads != config.d_model:
        raise HFT5ConfigError(f"d_kv ({config.d_kv}) x num_heads ({config.num_heads}) must equal d_model "
                              f"({config.d_model}) for a distributed T5 block")
    return (), dict(num_attention_heads=config.num_heads, attention_head_size=config.d_kv,
                    hidden_size=config.d_model, intermediate_size=config.d_ff,
                    attention_dropout_prob=config.dropout_rate, hidden_dropout_prob=config.dropout_rate,
                    add_cross_attention=bool(config.is_decoder))


def _bias_plus_mask(position_bias, mask):
    """HuggingFace hands every block the relative position bias [1|B, heads, Sq, Sk] and an additive mask
    [B, 1, 1|Sq, Sk]; the distributed layer takes their sum as one additive per-head bias (the layer keeps this
    rank's heads).  Unlike the reference (t5.py:62-72, which adds in place and so stacks the mask onto the shared bias
    again at every layer), the sum is a new tensor."""
    if position_bias is None:
        return mask
    return position_bias if mask is None else position_bias + mask


def t5_block_forward_hook(hidden_states, attention_mask=None, position_bias=None, encoder_hidden_states=None,
                          encoder_attention_mask=None, encoder_decoder_position_bias=None, layer_head_mask=None,
                          cross_attn_layer_head_mask=None, past_key_value=None, use_cache=False,
                          output_attentions=False, return_dict=True, **extra):
    """T5Block.forward arguments -> the distributed layer's input tuple (reference t5.py:35-98)."""
    past = past_key_value if past_key_value is not None else extra.get("past_key_values")
    head_masks = (layer_head_mask, cross_attn_layer_head_mask, extra.get("encoder_layer_head_mask"))
    if any(m is not None for m in head_masks) or past is not None or use_cache or output_attentions:
        raise HFT5ConfigError("head masks, cached keys / values, use_cache and output_attentions are not supported by "
                              "the distributed T5 block")
    if attention_mask is None and position_bias is None:
        raise HFT5ConfigError("attention_mask is a required argument of the distributed T5 block")
    self_mask = _bias_plus_mask(position_bias, attention_mask)
    if encoder_hidden_states is not None:
        cross_mask = _bias_plus_mask(encoder_decoder_position_bias, encoder_attention_mask)
        return ((hidden_states, self_mask, encoder_hidden_states, cross_mask),), {}
    return ((hidden_states, self_mask),), {}


def t5_block_return_hook(outputs):
    """The HuggingFace stack indexes a block's result as (hidden_states, position_bias[, cross position bias]); the
    biases come from the undistributed first block, so later blocks return None for them (reference t5.py:100-101)."""
    return (outputs[0], None, None)


def get_t5_block_hooks():
    return t5_block_init_hook, t5_block_forward_hook, t5_block_return_hook
