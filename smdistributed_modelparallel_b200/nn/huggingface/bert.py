"""HuggingFace BertEncoder / RobertaEncoder -> DistributedTransformer (post-LayerNorm encoder stack).

Reference: smdistributed/modelparallel/torch/nn/huggingface/bert.py:171-197 (init hook) and roberta.py;
translators map ``layer.{i}.attention.self.{query,key,value}`` / ``attention.output.{dense,LayerNorm}`` /
``intermediate.dense`` / ``output.{dense,LayerNorm}`` onto ``seq_layers.{i}.attention.*`` / ``output.*``.
"""

from __future__ import annotations

from ...backend.exceptions import HFBertConfigError, HFRobertaConfigError
from .common import Rule, Translator, check_unsupported, gelu_name

L = "seq_layers.{i}"


def _encoder_init_hook(exc):
    def hook(config, *args, **kwargs):
        if config.hidden_size % config.num_attention_heads != 0:
            raise exc("hidden_size must be divisible by num_attention_heads")
        if getattr(config, "position_embedding_type", "absolute") != "absolute":
            return None
        if getattr(config, "is_decoder", False) or getattr(config, "add_cross_attention", False):
            return None
        act = config.hidden_act if isinstance(config.hidden_act, str) else "gelu"
        return (), dict(
            num_layers=config.num_hidden_layers, num_attention_heads=config.num_attention_heads,
            attention_head_size=config.hidden_size // config.num_attention_heads, hidden_size=config.hidden_size,
            intermediate_size=config.intermediate_size, attention_dropout_prob=config.attention_probs_dropout_prob,
            hidden_dropout_prob=config.hidden_dropout_prob, activation=gelu_name(act),
            layernorm_epsilon=config.layer_norm_eps, initializer_range=config.initializer_range,
            use_normal_initialization=True, causal_mask_size=None, add_cross_attention=False, pre_layernorm=False,
            post_layernorm=True)
    return hook


def _encoder_forward_hook(name, exc):
    def hook(hidden_states, attention_mask=None, head_mask=None, encoder_hidden_states=None,
             encoder_attention_mask=None, past_key_values=None, use_cache=None, output_attentions=False,
             output_hidden_states=False, return_dict=True, **extra):
        hm = head_mask
        if isinstance(hm, (list, tuple)) and all(h is None for h in hm):
            hm = None
        check_unsupported(name, exc, head_mask=hm, encoder_hidden_states=encoder_hidden_states,
                          past_key_values=past_key_values)
        if output_attentions or output_hidden_states:
            raise exc(f"output_attentions / output_hidden_states are not supported for {name}")
        return ((hidden_states, attention_mask),), {}
    return hook


def _encoder_return_hook(outputs):
    from transformers.modeling_outputs import BaseModelOutputWithPastAndCrossAttentions

    return BaseModelOutputWithPastAndCrossAttentions(last_hidden_state=outputs[0])


def encoder_translator(config=None):
    rules = []
    for s in ("weight", "bias"):
        for n in ("query", "key", "value"):
            rules.append(Rule(f"layer.{{i}}.attention.self.{n}.{s}", f"{L}.attention.{n}.{s}"))
        rules.append(Rule(f"layer.{{i}}.attention.output.dense.{s}", f"{L}.attention.dense.{s}"))
        rules.append(Rule(f"layer.{{i}}.attention.output.LayerNorm.{s}", f"{L}.attention.layernorm.{s}"))
        rules.append(Rule(f"layer.{{i}}.intermediate.dense.{s}", f"{L}.output.dense1.{s}"))
        rules.append(Rule(f"layer.{{i}}.output.dense.{s}", f"{L}.output.dense2.{s}"))
        rules.append(Rule(f"layer.{{i}}.output.LayerNorm.{s}", f"{L}.output.layernorm.{s}"))
    return Translator(rules, (), config)


bert_init_hook = _encoder_init_hook(HFBertConfigError)
roberta_init_hook = _encoder_init_hook(HFRobertaConfigError)
bert_forward_hook = _encoder_forward_hook("BertEncoder", HFBertConfigError)
roberta_forward_hook = _encoder_forward_hook("RobertaEncoder", HFRobertaConfigError)
encoder_return_hook = _encoder_return_hook
