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
    if has_relative_attention_bias:
        raise HFT5ConfigError("T5 blocks that own the relative position bias cannot be distributed")
    return None
