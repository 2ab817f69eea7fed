"""HuggingFace GPT-2 / GPT-J / GPT-Neo / GPT-NeoX: init hooks (config -> DistributedTransformerLMHead
keywords), forward / return hooks (HF call signature and output classes), state-dict translators.

Reference: smdistributed/modelparallel/torch/nn/huggingface/gpt2.py, gptj.py:34-83 (+ translators),
gptneo.py, gptneox.py:35-90 (init hook), :194-222 / :300-350 (per-head QKV de-/re-interleave).
"""

from __future__ import annotations

import torch

from ...backend.exceptions import (HFGPT2ConfigError, HFGPTJConfigError, HFGPTNeoConfigError, HFGPTNeoxConfigError)
from .common import FusedQKVRule, Rule, Translator, _t, check_unsupported, gelu_name

L = "transformer.seq_layers.{i}"


def _lm_forward_hook(name, exc):
    def hook(input_ids=None, past_key_values=None, attention_mask=None, token_type_ids=None, position_ids=None,
             head_mask=None, inputs_embeds=None, encoder_hidden_states=None, encoder_attention_mask=None, labels=None,
             use_cache=None, output_attentions=None, output_hidden_states=None, return_dict=None, **extra):
        check_unsupported(name, exc, past_key_values=past_key_values, inputs_embeds=inputs_embeds, head_mask=head_mask,
                          output_attentions=output_attentions, output_hidden_states=output_hidden_states)
        if return_dict is not None and not return_dict:
            raise exc(f"return_dict=False is not supported for {name}")
        if encoder_hidden_states is not None:
            return ((input_ids, attention_mask, token_type_ids, position_ids, encoder_hidden_states,
                     encoder_attention_mask, labels),), {}
        return ((input_ids, attention_mask, token_type_ids, position_ids, labels),), {}
    return hook


def _lm_return_hook(outputs):
    from transformers.modeling_outputs import CausalLMOutputWithPast

    if isinstance(outputs, tuple):
        loss, logits = outputs
        return CausalLMOutputWithPast(loss=loss, logits=logits)
    return CausalLMOutputWithPast(loss=None, logits=outputs)


# ------------------------------------------------------------------------------------------- GPT-2
def gpt2_init_hook(config):
    if config.n_embd % config.n_head != 0:
        raise HFGPT2ConfigError(f"n_embd ({config.n_embd}) must be divisible by n_head ({config.n_head})")
    if getattr(config, "add_cross_attention", False) or getattr(config, "reorder_and_upcast_attn", False):
        return None
    return (), dict(
        num_layers=config.n_layer, num_attention_heads=config.n_head, attention_head_size=config.n_embd // config.n_head,
        hidden_size=config.n_embd, intermediate_size=config.n_inner if config.n_inner is not None else 4 * config.n_embd,
        vocab_size=config.vocab_size, num_positions=config.n_positions, attention_dropout_prob=config.attn_pdrop,
        hidden_dropout_prob=config.resid_pdrop, embedding_dropout_prob=config.embd_pdrop,
        activation=gelu_name(config.activation_function), layernorm_epsilon=config.layer_norm_epsilon,
        causal_mask_size=config.n_positions, add_lm_head=True, initializer_range=config.initializer_range,
        use_normal_initialization=True, pre_layernorm=True, post_layernorm=False, final_layernorm=True,
        scale_attention_scores=getattr(config, "scale_attn_weights", True),
        scale_attn_by_layer_idx=getattr(config, "scale_attn_by_inverse_layer_idx", False),
        tie_input_output_embedding=getattr(config, "tie_word_embeddings", True))


def _split3(w, cfg):          # GPT-2 Conv1D c_attn: weight [H, 3H] (in, out) / bias [3H]
    if w.dim() == 2:
        q, k, v = w.t().chunk(3, dim=0)
        return q.contiguous(), k.contiguous(), v.contiguous()
    return tuple(c.contiguous() for c in w.chunk(3, dim=0))


def _merge3(q, k, v, cfg):
    if q.dim() == 2:
        return torch.cat([q, k, v], dim=0).t().contiguous()
    return torch.cat([q, k, v], dim=0)


def gpt2_translator(config=None):
    rules = [
        Rule("transformer.wte.weight", "word_embedding.weight"), Rule("transformer.wpe.weight", "position_embedding.weight"),
        Rule("transformer.ln_f.weight", "layernorm.weight"), Rule("transformer.ln_f.bias", "layernorm.bias"),
        Rule("lm_head.weight", "lm_head.weight"),
    ]
    for hf, smp in (("ln_1", "attention.pre_layernorm"), ("ln_2", "output.pre_layernorm")):
        for s in ("weight", "bias"):
            rules.append(Rule(f"transformer.h.{{i}}.{hf}.{s}", f"{L}.{smp}.{s}"))
    for hf, smp in (("attn.c_proj", "attention.dense"), ("mlp.c_fc", "output.dense1"), ("mlp.c_proj", "output.dense2")):
        rules.append(Rule(f"transformer.h.{{i}}.{hf}.weight", f"{L}.{smp}.weight", _t, _t))
        rules.append(Rule(f"transformer.h.{{i}}.{hf}.bias", f"{L}.{smp}.bias"))
    qkv = [FusedQKVRule("transformer.h.{i}.attn.c_attn.weight", f"{L}.attention", "weight", _split3, _merge3),
           FusedQKVRule("transformer.h.{i}.attn.c_attn.bias", f"{L}.attention", "bias", _split3, _merge3)]
    return Translator(rules, qkv, config, drop_hf=[r"\.attn\.bias$", r"\.attn\.masked_bias$"])


def gpt2_layer_init_hook(config, layer_idx=None):
    out = gpt2_init_hook(config)
    if out is None:
        return None
    kw = out[1]
    keep = ("num_attention_heads", "attention_head_size", "hidden_size", "intermediate_size", "attention_dropout_prob",
            "hidden_dropout_prob", "activation", "layernorm_epsilon", "initializer_range", "use_normal_initialization",
            "causal_mask_size", "pre_layernorm", "post_layernorm", "scale_attention_scores", "scale_attn_by_layer_idx")
    kw = {k: kw[k] for k in keep}
    kw["layer_idx"] = layer_idx or 0
    kw["num_positions"] = config.n_positions
    return (), kw


def gpt2_layer_forward_hook(hidden_states, layer_past=None, attention_mask=None, head_mask=None,
                            encoder_hidden_states=None, encoder_attention_mask=None, use_cache=False,
                            output_attentions=False, **extra):
    check_unsupported("GPT2Block", HFGPT2ConfigError, layer_past=layer_past, head_mask=head_mask,
                      encoder_hidden_states=encoder_hidden_states)
    return ((hidden_states, attention_mask),), {}


def gpt2_layer_return_hook(outputs):
    return (outputs[0],)


# ------------------------------------------------------------------------------------------- GPT-J
def gptj_init_hook(config):
    if config.n_embd % config.n_head != 0:
        raise HFGPTJConfigError(f"n_embd ({config.n_embd}) must be divisible by n_head ({config.n_head})")
    if config.activation_function not in ("gelu_new", "relu"):
        raise HFGPTJConfigError("Only 'gelu_new' and 'relu' activations are supported.")
    return (), dict(
        num_layers=config.n_layer, num_attention_heads=config.n_head, attention_head_size=config.n_embd // config.n_head,
        hidden_size=config.n_embd, rotary_dim=config.rotary_dim, mask_value=-1e9, use_positional_embedding=False,
        parallel_attn_output=True, use_lm_head_bias=True, tie_input_output_embedding=config.tie_word_embeddings,
        use_attn_dense_bias=False, use_qkv_bias=False, final_layernorm=True, single_pre_layernorm=True,
        vocab_size=config.vocab_size, activation=gelu_name(config.activation_function), add_lm_head=True,
        intermediate_size=config.n_inner if config.n_inner is not None else 4 * config.n_embd,
        attention_dropout_prob=config.attn_pdrop, hidden_dropout_prob=config.resid_pdrop,
        embedding_dropout_prob=config.embd_pdrop, layernorm_epsilon=config.layer_norm_epsilon,
        initializer_range=config.initializer_range, use_normal_initialization=True, pre_layernorm=True,
        post_layernorm=False, causal_mask_size=config.n_positions, num_positions=config.n_positions,
        scale_attention_scores=getattr(config, "scale_attn_weights", True))


def gptj_translator(config=None):
    rules = [Rule("transformer.wte.weight", "word_embedding.weight"), Rule("transformer.ln_f.weight", "layernorm.weight"),
             Rule("transformer.ln_f.bias", "layernorm.bias"), Rule("lm_head.weight", "lm_head.weight"),
             Rule("lm_head.bias", "lm_head.bias")]
    for s in ("weight", "bias"):
        rules.append(Rule(f"transformer.h.{{i}}.ln_1.{s}", f"{L}.attention.pre_layernorm.{s}"))
        rules.append(Rule(f"transformer.h.{{i}}.mlp.fc_in.{s}", f"{L}.output.dense1.{s}"))
        rules.append(Rule(f"transformer.h.{{i}}.mlp.fc_out.{s}", f"{L}.output.dense2.{s}"))
    for hf, smp in (("q_proj", "query"), ("k_proj", "key"), ("v_proj", "value"), ("out_proj", "dense")):
        rules.append(Rule(f"transformer.h.{{i}}.attn.{hf}.weight", f"{L}.attention.{smp}.weight"))
    return Translator(rules, (), config, drop_hf=[r"\.attn\.bias$", r"\.attn\.masked_bias$", r"embed_positions"])


# ------------------------------------------------------------------------------------------- GPT-Neo
def gptneo_init_hook(config):
    if config.hidden_size % config.num_heads != 0:
        raise HFGPTNeoConfigError("hidden_size must be divisible by num_heads")
    return (), dict(
        num_layers=config.num_layers, num_attention_heads=config.num_heads,
        attention_head_size=config.hidden_size // config.num_heads, hidden_size=config.hidden_size,
        intermediate_size=config.intermediate_size if config.intermediate_size is not None else 4 * config.hidden_size,
        vocab_size=config.vocab_size, num_positions=config.max_position_embeddings,
        attention_dropout_prob=config.attention_dropout, hidden_dropout_prob=config.resid_dropout,
        embedding_dropout_prob=config.embed_dropout, activation=gelu_name(config.activation_function),
        layernorm_epsilon=config.layer_norm_epsilon, causal_mask_size=config.max_position_embeddings, add_lm_head=True,
        initializer_range=config.initializer_range, use_normal_initialization=True, pre_layernorm=True,
        post_layernorm=False, final_layernorm=True, use_qkv_bias=False, scale_attention_scores=False,
        attention_layers_type=list(config.attention_layers), window_size=config.window_size,
        tie_input_output_embedding=getattr(config, "tie_word_embeddings", True))


def gptneo_translator(config=None):
    rules = [Rule("transformer.wte.weight", "word_embedding.weight"), Rule("transformer.wpe.weight", "position_embedding.weight"),
             Rule("transformer.ln_f.weight", "layernorm.weight"), Rule("transformer.ln_f.bias", "layernorm.bias"),
             Rule("lm_head.weight", "lm_head.weight")]
    for s in ("weight", "bias"):
        rules.append(Rule(f"transformer.h.{{i}}.ln_1.{s}", f"{L}.attention.pre_layernorm.{s}"))
        rules.append(Rule(f"transformer.h.{{i}}.ln_2.{s}", f"{L}.output.pre_layernorm.{s}"))
        rules.append(Rule(f"transformer.h.{{i}}.attn.attention.out_proj.{s}", f"{L}.attention.dense.{s}"))
        rules.append(Rule(f"transformer.h.{{i}}.mlp.c_fc.{s}", f"{L}.output.dense1.{s}"))
        rules.append(Rule(f"transformer.h.{{i}}.mlp.c_proj.{s}", f"{L}.output.dense2.{s}"))
    for hf, smp in (("q_proj", "query"), ("k_proj", "key"), ("v_proj", "value")):
        rules.append(Rule(f"transformer.h.{{i}}.attn.attention.{hf}.weight", f"{L}.attention.{smp}.weight"))
    return Translator(rules, (), config, drop_hf=[r"attention\.bias$", r"attention\.masked_bias$"])


# ------------------------------------------------------------------------------------------- GPT-NeoX
def gptneox_init_hook(config):
    if config.hidden_size % config.num_attention_heads != 0:
        raise HFGPTNeoxConfigError("hidden_size must be divisible by num_attention_heads")
    if not getattr(config, "use_parallel_residual", True):
        return None
    d = config.hidden_size // config.num_attention_heads
    rp = getattr(config, "rope_parameters", None) or {}
    rope = rp.get("rope_theta") or getattr(config, "rotary_emb_base", None) or getattr(config, "rope_theta", 10000)
    pct = rp.get("partial_rotary_factor")
    if pct is None:
        pct = getattr(config, "rotary_pct", None)
    if pct is None:
        pct = getattr(config, "partial_rotary_factor", 0.25)
    if rp.get("rope_type", "default") != "default":
        return None
    return (), dict(
        num_layers=config.num_hidden_layers, num_attention_heads=config.num_attention_heads, attention_head_size=d,
        hidden_size=config.hidden_size, intermediate_size=config.intermediate_size, vocab_size=config.vocab_size,
        num_positions=config.max_position_embeddings, attention_dropout_prob=getattr(config, "attention_dropout", 0.0),
        hidden_dropout_prob=getattr(config, "hidden_dropout", 0.0), embedding_dropout_prob=0.0,
        activation=gelu_name(config.hidden_act), layernorm_epsilon=config.layer_norm_eps,
        causal_mask_size=config.max_position_embeddings, add_lm_head=True, initializer_range=config.initializer_range,
        use_normal_initialization=True, pre_layernorm=True, post_layernorm=False, final_layernorm=True,
        rotary_dim=int(d * pct), rotary_emb_base=rope, gpt_neox_type_rotary=True, use_positional_embedding=False,
        parallel_attn_output=True, tie_input_output_embedding=getattr(config, "tie_word_embeddings", False),
        use_lm_head_bias=False)


def _neox_split(w, cfg):
    """query_key_value is laid out per head as [heads, 3, d, ...]: pull q, k, v apart."""
    heads = cfg.num_attention_heads
    d = cfg.hidden_size // heads
    if w.dim() == 2:
        x = w.view(heads, 3, d, w.shape[1])
        return tuple(x[:, j].reshape(heads * d, w.shape[1]).contiguous() for j in range(3))
    x = w.view(heads, 3, d)
    return tuple(x[:, j].reshape(heads * d).contiguous() for j in range(3))


def _neox_merge(q, k, v, cfg):
    heads = cfg.num_attention_heads
    d = cfg.hidden_size // heads
    if q.dim() == 2:
        x = torch.stack([t.view(heads, d, t.shape[1]) for t in (q, k, v)], dim=1)
        return x.reshape(3 * heads * d, q.shape[1]).contiguous()
    x = torch.stack([t.view(heads, d) for t in (q, k, v)], dim=1)
    return x.reshape(3 * heads * d).contiguous()


def gptneox_translator(config=None):
    P = "gpt_neox.layers.{i}"
    rules = [Rule("gpt_neox.embed_in.weight", "word_embedding.weight"),
             Rule("gpt_neox.final_layer_norm.weight", "layernorm.weight"),
             Rule("gpt_neox.final_layer_norm.bias", "layernorm.bias"), Rule("embed_out.weight", "lm_head.weight")]
    for s in ("weight", "bias"):
        rules.append(Rule(f"{P}.input_layernorm.{s}", f"{L}.attention.pre_layernorm.{s}"))
        rules.append(Rule(f"{P}.post_attention_layernorm.{s}", f"{L}.output.pre_layernorm.{s}"))
        rules.append(Rule(f"{P}.attention.dense.{s}", f"{L}.attention.dense.{s}"))
        rules.append(Rule(f"{P}.mlp.dense_h_to_4h.{s}", f"{L}.output.dense1.{s}"))
        rules.append(Rule(f"{P}.mlp.dense_4h_to_h.{s}", f"{L}.output.dense2.{s}"))
    qkv = [FusedQKVRule(f"{P}.attention.query_key_value.weight", f"{L}.attention", "weight", _neox_split, _neox_merge),
           FusedQKVRule(f"{P}.attention.query_key_value.bias", f"{L}.attention", "bias", _neox_split, _neox_merge)]
    return Translator(rules, qkv, config, drop_hf=[r"attention\.bias$", r"masked_bias$", r"rotary_emb\.inv_freq$"])


FORWARD = {"gpt2": _lm_forward_hook("GPT2LMHeadModel", HFGPT2ConfigError),
           "gptj": _lm_forward_hook("GPTJForCausalLM", HFGPTJConfigError),
           "gptneo": _lm_forward_hook("GPTNeoForCausalLM", HFGPTNeoConfigError),
           "gptneox": _lm_forward_hook("GPTNeoXForCausalLM", HFGPTNeoxConfigError)}
RETURN = _lm_return_hook
