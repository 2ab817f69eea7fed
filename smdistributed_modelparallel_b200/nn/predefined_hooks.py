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


_TABLE = (
    # key, modeling module, class, distributed class name, hook names in the family module, translator factory
    ("huggingface-bert", "transformers.models.bert.modeling_bert", "BertEncoder", "DistributedTransformer",
     ("B", "bert_init_hook", "bert_forward_hook", "encoder_return_hook", "encoder_translator")),
    ("huggingface-roberta", "transformers.models.roberta.modeling_roberta", "RobertaEncoder", "DistributedTransformer",
     ("B", "roberta_init_hook", "roberta_forward_hook", "encoder_return_hook", "encoder_translator")),
    ("huggingface-gpt-2", "transformers.models.gpt2.modeling_gpt2", "GPT2LMHeadModel", "DistributedTransformerLMHead",
     ("G", "gpt2_init_hook", ("FORWARD", "gpt2"), "RETURN", "gpt2_translator")),
    ("huggingface-gpt-2-layer", "transformers.models.gpt2.modeling_gpt2", "GPT2Block", "DistributedTransformerLayer",
     ("G", "gpt2_layer_init_hook", "gpt2_layer_forward_hook", "gpt2_layer_return_hook", None)),
    ("huggingface-gpt-j", "transformers.models.gptj.modeling_gptj", "GPTJForCausalLM", "DistributedTransformerLMHead",
     ("G", "gptj_init_hook", ("FORWARD", "gptj"), "RETURN", "gptj_translator")),
    ("huggingface-gpt-neo", "transformers.models.gpt_neo.modeling_gpt_neo", "GPTNeoForCausalLM",
     "DistributedTransformerLMHead", ("G", "gptneo_init_hook", ("FORWARD", "gptneo"), "RETURN", "gptneo_translator")),
    ("huggingface-gpt-neox", "transformers.models.gpt_neox.modeling_gpt_neox", "GPTNeoXForCausalLM",
     "DistributedTransformerLMHead", ("G", "gptneox_init_hook", ("FORWARD", "gptneox"), "RETURN", "gptneox_translator")),
)


def register_predefined(registry) -> list:
    """Register the HuggingFace mappings.  Importing seven modeling modules costs ~10 s, so each mapping is
    installed by a post-import hook on its modeling module: immediately when that module is already loaded,
    otherwise the moment the user's script first imports the model class (which necessarily precedes
    constructing it, the point at which the registry has to see the constructor arguments)."""
    import importlib.util
    import sys

    if importlib.util.find_spec("transformers") is None:
        return []
    done = []

    def install(mod, key, cls_name, dist_name, spec):
        from . import transformer as T
        from .huggingface import bert as B
        from .huggingface import gpt as G

        cls = getattr(mod, cls_name, None)
        if cls is None or cls in registry.tp_modules:
            return
        fam = {"B": B, "G": G}[spec[0]]

        def pick(name):
            if name is None:
                return None
            if isinstance(name, tuple):
                return getattr(fam, name[0])[name[1]]
            return getattr(fam, name)

        init, fwd, ret, make = pick(spec[1]), pick(spec[2]), pick(spec[3]), pick(spec[4])
        holder = {}
        registry.register_with_module(cls, getattr(T, dist_name), _capturing(init, holder), fwd, ret,
                                      _translators(make, holder) if make else None)
        done.append(key)

    try:
        import wrapt
    except ImportError:                                           # pragma: no cover - eager fallback
        wrapt = None
    for key, path, cls_name, dist_name, spec in _TABLE:
        hook = (lambda mod, _a=(key, cls_name, dist_name, spec): install(mod, *_a))
        if path in sys.modules:
            hook(sys.modules[path])
        elif wrapt is not None:
            wrapt.register_post_import_hook(hook, path)
        else:
            try:
                hook(importlib.import_module(path))
            except Exception as e:
                get_logger().debug(f"{key}: unavailable ({e})")
    return done
