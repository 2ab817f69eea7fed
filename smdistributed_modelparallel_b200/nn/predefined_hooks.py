"""Registration of the built-in HuggingFace mappings (reference smdistributed/modelparallel/torch/nn/
predefined_hooks.py:81-151): BertEncoder, RobertaEncoder -> DistributedTransformer; GPT2LMHeadModel,
GPTJForCausalLM, GPTNeoForCausalLM, GPTNeoXForCausalLM -> DistributedTransformerLMHead; GPT2Block ->
DistributedTransformerLayer.  The reference accepts transformers 4.17 / 4.20 / 4.21 only; here any
version whose model classes import is accepted, with config fields read defensively.
"""

from __future__ import annotations

from ..backend.logger import get_logger


def _translators(make, config_holder):
    """(smp -> hf, hf -> smp) closures; the model config is captured when the init hook runs."""
    def smp_to_hf(sd):
        return make(config_holder.get("config")).smp_to_hf(sd)

    def hf_to_smp(sd):
        return make(config_holder.get("config")).hf_to_smp(sd)

    return (smp_to_hf, hf_to_smp)


def _capturing(init_hook, holder):
    def hook(config, *args, **kwargs):
        holder["config"] = config
        return init_hook(config, *args, **kwargs)
    return hook


def register_predefined(registry) -> list:
    try:
        import transformers  # noqa: F401
    except Exception as e:  # pragma: no cover
        get_logger().debug(f"transformers not importable: {e}")
        return []
    from .huggingface import bert as B
    from .huggingface import gpt as G
    from .transformer import DistributedTransformer, DistributedTransformerLayer, DistributedTransformerLMHead

    done = []

    def add(key, import_path, cls_name, dist_cls, init, fwd, ret, make_translator):
        try:
            mod = __import__(import_path, fromlist=[cls_name])
            cls = getattr(mod, cls_name)
        except Exception as e:
            get_logger().debug(f"{key}: {cls_name} unavailable ({e})")
            return
        holder = {}
        registry.register_with_module(cls, dist_cls, _capturing(init, holder), fwd, ret,
                                      _translators(make_translator, holder) if make_translator else None)
        done.append(key)

    add("huggingface-bert", "transformers.models.bert.modeling_bert", "BertEncoder", DistributedTransformer,
        B.bert_init_hook, B.bert_forward_hook, B.encoder_return_hook, B.encoder_translator)
    add("huggingface-roberta", "transformers.models.roberta.modeling_roberta", "RobertaEncoder", DistributedTransformer,
        B.roberta_init_hook, B.roberta_forward_hook, B.encoder_return_hook, B.encoder_translator)
    add("huggingface-gpt-2", "transformers.models.gpt2.modeling_gpt2", "GPT2LMHeadModel", DistributedTransformerLMHead,
        G.gpt2_init_hook, G.FORWARD["gpt2"], G.RETURN, G.gpt2_translator)
    add("huggingface-gpt-2-layer", "transformers.models.gpt2.modeling_gpt2", "GPT2Block", DistributedTransformerLayer,
        G.gpt2_layer_init_hook, G.gpt2_layer_forward_hook, G.gpt2_layer_return_hook, None)
    add("huggingface-gpt-j", "transformers.models.gptj.modeling_gptj", "GPTJForCausalLM", DistributedTransformerLMHead,
        G.gptj_init_hook, G.FORWARD["gptj"], G.RETURN, G.gptj_translator)
    add("huggingface-gpt-neo", "transformers.models.gpt_neo.modeling_gpt_neo", "GPTNeoForCausalLM",
        DistributedTransformerLMHead, G.gptneo_init_hook, G.FORWARD["gptneo"], G.RETURN, G.gptneo_translator)
    add("huggingface-gpt-neox", "transformers.models.gpt_neox.modeling_gpt_neox", "GPTNeoXForCausalLM",
        DistributedTransformerLMHead, G.gptneox_init_hook, G.FORWARD["gptneox"], G.RETURN, G.gptneox_translator)
    return done
