"""Random HuggingFace configurations (GPT-2, GPT-J, GPT-NeoX, GPT-Neo: head counts that do not divide the degree, odd
rotary dimensions, local-attention windows, layer counts) under tensor parallelism on 2-3 gloo ranks: the replaced model,
loaded from the original's state dict, reproduces the original's loss, and the state dict translated back equals the
original (reference test/torch/mpi/test_translate_state_dict.py, mpi_hybrid/test_gpt_grad.py fix one shape per
family).  ``SMP_RANDOM_HF_SEEDS=a-b`` widens the search."""

import os
import random

import pytest
import torch

transformers = pytest.importorskip("transformers")

from dist_utils import run_distributed


def options(seed):
    rng = random.Random(seed)
    fam = rng.choice(["gpt2", "gptj", "gptneox", "gptneo"])
    tp = rng.choice([2, 3])
    heads = rng.randint(tp, 6)
    d = rng.choice([4, 8])
    return {"family": fam, "tp": tp, "heads": heads, "hidden": heads * d, "layers": rng.randint(1, 3),
            "rotary": rng.choice([2, 4]) if d >= 4 else 2, "window": rng.choice([4, 8]), "vocab": rng.choice([50, 64, 97]),
            "seq": 16}


def make(o):
    fam = o["family"]
    if fam == "gpt2":
        from transformers import GPT2Config, GPT2LMHeadModel as Cls
        cfg = GPT2Config(n_layer=o["layers"], n_head=o["heads"], n_embd=o["hidden"], n_positions=32, vocab_size=o["vocab"],
                         resid_pdrop=0, embd_pdrop=0, attn_pdrop=0)
    elif fam == "gptj":
        from transformers import GPTJConfig, GPTJForCausalLM as Cls
        cfg = GPTJConfig(n_layer=o["layers"], n_head=o["heads"], n_embd=o["hidden"], n_positions=32, vocab_size=o["vocab"],
                         rotary_dim=o["rotary"], resid_pdrop=0, embd_pdrop=0, attn_pdrop=0)
    elif fam == "gptneox":
        from transformers import GPTNeoXConfig, GPTNeoXForCausalLM as Cls
        cfg = GPTNeoXConfig(num_hidden_layers=o["layers"], num_attention_heads=o["heads"], hidden_size=o["hidden"],
                            intermediate_size=2 * o["hidden"], max_position_embeddings=32, vocab_size=o["vocab"],
                            rotary_pct=0.5, hidden_dropout=0.0, attention_dropout=0.0)
    else:
        from transformers import GPTNeoConfig, GPTNeoForCausalLM as Cls
        kinds = [["global", "local"][i % 2] for i in range(o["layers"])]
        cfg = GPTNeoConfig(num_layers=o["layers"], num_heads=o["heads"], hidden_size=o["hidden"], max_position_embeddings=32,
                           vocab_size=o["vocab"], attention_types=[[kinds, 1]], window_size=o["window"], resid_dropout=0,
                           embed_dropout=0, attention_dropout=0)
    return cfg, Cls


def _worker(rank, world, seed):
    import smdistributed_modelparallel_b200 as smp

    o = options(seed)
    smp.init({"tensor_parallel_degree": o["tp"], "ddp": True, "microbatches": 1, "ddp_dist_backend": "gloo"})
    cfg, Cls = make(o)
    torch.manual_seed(seed)
    hf = Cls(cfg).eval()
    hf_sd = {k: v.clone() for k, v in hf.state_dict().items()}
    ids_all = torch.randint(0, o["vocab"], (2 * world, o["seq"]), generator=torch.Generator().manual_seed(seed))
    with torch.no_grad():
        ref = hf(input_ids=ids_all, labels=ids_all).loss.item()
    with smp.tensor_parallelism(enabled=True):
        hf2 = Cls(cfg)
    model = smp.DistributedModel(hf2)
    replaced = type(model.module).__name__
    model.load_state_dict(hf_sd)
    model.eval()

    @smp.step
    def fwd(model, ids):
        return model(input_ids=ids, labels=ids).loss

    with torch.no_grad():
        loss = fwd(model, ids_all[rank * 2:(rank + 1) * 2]).reduce_mean().item()
    to_hf = smp.state.tp_registry.active_translate_functions[0][0] if replaced != Cls.__name__ else (lambda sd: sd)
    back = to_hf(model.state_dict())
    worst = max((back[k].float() - v.float()).abs().max().item() for k, v in hf_sd.items()
                if k in back and v.is_floating_point())
    missing = [k for k in hf_sd if k not in back and not k.endswith(("attn.bias", "masked_bias", "inv_freq"))]
    return {"ref": ref, "loss": loss, "replaced": replaced, "worst": worst, "missing": missing}


def _seeds():
    spec = os.environ.get("SMP_RANDOM_HF_SEEDS", "")
    if spec:
        a, b = spec.split("-")
        return list(range(int(a), int(b) + 1))
    return [1, 2, 3, 4, 6, 7, 9, 12]


@pytest.mark.parametrize("seed", _seeds())
def test_random_huggingface_configurations_under_tensor_parallelism(seed):
    o = options(seed)
    res = run_distributed(_worker, o["tp"], seed, timeout=300)
    mean = sum(r["loss"] for r in res) / len(res)
    assert abs(mean - res[0]["ref"]) < 1e-4, (seed, o, mean, res[0]["ref"], res[0]["replaced"])
    for r in res:
        assert r["replaced"] == "DistributedTransformerLMHead", (seed, o, r["replaced"])
        assert r["worst"] == 0.0 and not r["missing"], (seed, o, r["worst"], r["missing"][:4])
