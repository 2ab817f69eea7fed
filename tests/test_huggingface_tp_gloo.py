"""HuggingFace GPT-2 / GPT-J under tensor_parallel_degree=2 on CPU/gloo: the auto-replaced model, loaded from the
original HF state dict through the translator, must reproduce the HF loss on the union batch and hand back an
HF-format state dict (reference tier: test/torch/mpi_hybrid/test_gpt_grad.py, TP registry tests in test/torch/mpi)."""

import pytest
import torch

transformers = pytest.importorskip("transformers")

from dist_utils import run_distributed


def _worker(rank, world, family):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"tensor_parallel_degree": 2, "ddp": True, "microbatches": 1, "ddp_dist_backend": "gloo"})
    torch.manual_seed(0)
    if family == "gpt2":
        from transformers import GPT2Config, GPT2LMHeadModel as Cls
        cfg = GPT2Config(n_layer=2, n_head=4, n_embd=32, n_positions=32, vocab_size=96, resid_pdrop=0, embd_pdrop=0,
                         attn_pdrop=0)
    elif family == "gptneox":
        from transformers import GPTNeoXConfig, GPTNeoXForCausalLM as Cls
        cfg = GPTNeoXConfig(num_hidden_layers=2, num_attention_heads=4, hidden_size=32, intermediate_size=64,
                            max_position_embeddings=32, vocab_size=96, rotary_pct=0.5, hidden_dropout=0.0,
                            attention_dropout=0.0)
    elif family == "gptneo":
        from transformers import GPTNeoConfig, GPTNeoForCausalLM as Cls
        cfg = GPTNeoConfig(num_layers=2, num_heads=4, hidden_size=32, max_position_embeddings=32, vocab_size=96,
                           attention_types=[[["global", "local"], 1]], window_size=8, resid_dropout=0, embed_dropout=0,
                           attention_dropout=0)
    else:
        from transformers import GPTJConfig, GPTJForCausalLM as Cls
        cfg = GPTJConfig(n_layer=2, n_head=4, n_embd=32, n_positions=32, vocab_size=96, rotary_dim=4, resid_pdrop=0,
                         embd_pdrop=0, attn_pdrop=0)
    hf = Cls(cfg).eval()
    hf_sd = {k: v.clone() for k, v in hf.state_dict().items()}
    ids_all = torch.randint(0, cfg.vocab_size, (4, 16))
    with torch.no_grad():
        ref = hf(input_ids=ids_all, labels=ids_all).loss.item()
    with smp.tensor_parallelism(enabled=True):
        hf2 = Cls(cfg)
    model = smp.DistributedModel(hf2)
    replaced = type(model.module).__name__
    model.load_state_dict(hf_sd)                       # full HF checkpoint: translated, then sliced per tp rank
    model.eval()

    @smp.step
    def fwd(model, ids):
        return model(input_ids=ids, labels=ids).loss

    with torch.no_grad():
        loss = fwd(model, ids_all[rank * 2:(rank + 1) * 2]).reduce_mean().item()
    to_hf, _ = smp.state.tp_registry.active_translate_functions[0]
    back = to_hf(model.state_dict())                   # gathered over tp, back in HF naming
    worst = max((back[k].float() - v.float()).abs().max().item() for k, v in hf_sd.items() if k in back)
    missing = [k for k in hf_sd if k not in back]
    local_qkv = [tuple(p.shape) for n, p in model.module.named_parameters() if n.endswith("qkv_weight")][0]
    return {"ref": ref, "loss": loss, "replaced": replaced, "worst": worst, "missing": missing, "qkv": local_qkv}


@pytest.mark.parametrize("family", ["gpt2", "gptj", "gptneox", "gptneo"])
def test_hf_model_under_tp2(family):
    res = run_distributed(_worker, 2, family, timeout=300)
    assert all(r["replaced"] == "DistributedTransformerLMHead" for r in res)
    assert all(r["qkv"] == (48, 32) for r in res), res[0]["qkv"]       # 3 x (32 / 2) rows per rank
    # token-mean loss over the union batch = mean of the two ranks' losses (equal token counts)
    mean = sum(r["loss"] for r in res) / 2
    assert abs(mean - res[0]["ref"]) < 5e-5, (mean, res[0]["ref"])
    for r in res:
        assert r["worst"] == 0.0 and not r["missing"], (r["worst"], r["missing"][:4])


def _hf_pp_worker(rank, world):
    import copy

    import smdistributed_modelparallel_b200 as smp
    from transformers import GPT2Config, GPT2LMHeadModel

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 2, "ddp": False})
    torch.manual_seed(1)
    cfg = GPT2Config(n_layer=4, n_head=4, n_embd=32, vocab_size=64, n_positions=16, resid_pdrop=0.0, embd_pdrop=0.0,
                     attn_pdrop=0.0)
    net = GPT2LMHeadModel(cfg)
    oracle = copy.deepcopy(net)
    model = smp.DistributedModel(net)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))

    @smp.step
    def step(model, ids):
        out = model(input_ids=ids, labels=ids)            # keyword arguments, ModelOutput result
        model.backward(out.loss)
        return out.loss

    ids = torch.randint(0, 64, (4, 16), generator=torch.Generator().manual_seed(3))
    opt.zero_grad()
    loss = step(model, ids).reduce_mean().item()
    ol = sum(oracle(input_ids=ids[i:i + 2], labels=ids[i:i + 2]).loss / 2 for i in (0, 2))
    ol.backward()
    ref = dict(oracle.named_parameters())
    worst = max((p.grad - ref[n].grad).abs().max().item() for n, p in model.local_named_parameters())
    opt.step()
    loss2 = step(model, ids).reduce_mean().item()
    return {"loss": loss, "oracle": ol.item(), "worst": worst, "loss2": loss2,
            "ranks": sorted(set(smp.state.module_manager.partition_info().values()))}


def test_stock_hf_gpt2_under_pp2():
    """An unmodified HuggingFace GPT-2 (keyword arguments, ``ModelOutput`` results, tied embeddings) split over two
    pipeline stages by the auto-partitioner: loss and local gradients match the single-process model."""
    res = run_distributed(_hf_pp_worker, 2, timeout=300)
    for r in res:
        assert r["ranks"] == [0, 1]
        assert abs(r["loss"] - r["oracle"]) < 5e-5 and r["worst"] < 3e-5, r
        assert r["loss2"] < r["loss"]
