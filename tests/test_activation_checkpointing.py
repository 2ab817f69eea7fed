"""Activation checkpointing on one process (reference test/torch/test_activation_checkpointing.py): smp.checkpoint
on modules and functions with nested inputs, checkpoint_sequential strategies, RNG preservation with dropout,
and smp.set_activation_checkpointing applied through DistributedModel."""

import pytest
import torch
import torch.nn as nn

import smdistributed_modelparallel_b200 as smp
from smdistributed_modelparallel_b200.backend.exceptions import CheckpointingConfigError
from smdistributed_modelparallel_b200.runtime.checkpointing import _chunks, checkpoint, checkpoint_sequential


def _mlp(seed, drop=0.0):
    torch.manual_seed(seed)
    return nn.Sequential(nn.Linear(16, 32), nn.Tanh(), nn.Dropout(drop), nn.Linear(32, 16))


def _grads(model):
    return [p.grad.clone() for p in model.parameters()]


def test_checkpoint_module_matches_plain_backward():
    m = _mlp(0)
    x = torch.randn(4, 16, requires_grad=True)
    m(x).square().sum().backward()
    ref, refx = _grads(m), x.grad.clone()
    m.zero_grad()
    x.grad = None
    checkpoint(m, x).square().sum().backward()
    for a, b in zip(ref, _grads(m)):
        assert torch.allclose(a, b)
    assert torch.allclose(refx, x.grad)


def test_checkpoint_reaches_parameters_when_no_input_needs_grad():
    m = _mlp(1)
    x = torch.randn(4, 16)                                   # no grad on the input
    checkpoint(m, x).sum().backward()
    assert all(p.grad is not None and p.grad.abs().sum() > 0 for p in m.parameters())


def test_checkpoint_nested_inputs_and_outputs():
    lin = nn.Linear(8, 8)

    def fn(pair, extra=None, flag=True):
        a, b = pair
        return {"s": lin(a) + b * extra["w"], "t": (a.sum(), flag)}

    a, b = torch.randn(3, 8, requires_grad=True), torch.randn(3, 8, requires_grad=True)
    w = torch.randn(3, 8)
    out = checkpoint(fn, (a, b), extra={"w": w}, flag=False)
    assert out["t"][1] is False
    (out["s"].sum() + out["t"][0]).backward()
    ref = fn((a.detach().requires_grad_(True), b.detach().requires_grad_(True)), extra={"w": w})
    assert torch.allclose(out["s"], ref["s"])
    assert torch.allclose(b.grad, w)


def test_rng_state_is_replayed_for_dropout():
    m = _mlp(2, drop=0.5)
    m.train()
    x = torch.randn(64, 16, requires_grad=True)
    torch.manual_seed(123)
    y = checkpoint(m, x, preserve_rng_state=True)
    y.sum().backward()                                       # recomputation must draw the same dropout mask
    g = x.grad.clone()
    x.grad = None
    m.zero_grad()
    torch.manual_seed(123)
    m(x).sum().backward()
    assert torch.allclose(g, x.grad)


@pytest.mark.parametrize("strategy,expected", [("each", [(0, 1), (1, 2), (2, 3), (3, 4), (4, 5)]),
                                               ("contiguous", [(0, 5)]), ("group_2", [(0, 2), (2, 4), (4, 5)])])
def test_sequential_strategies(strategy, expected):
    assert _chunks(5, strategy) == expected
    torch.manual_seed(3)
    seq = nn.Sequential(*[nn.Sequential(nn.Linear(8, 8), nn.Tanh()) for _ in range(5)])
    x = torch.randn(2, 8, requires_grad=True)
    seq(x).sum().backward()
    ref = _grads(seq)
    seq.zero_grad()
    checkpoint_sequential(seq, x, strategy=strategy).sum().backward()
    for a, b in zip(ref, _grads(seq)):
        assert torch.allclose(a, b)


def test_bad_strategy():
    with pytest.raises(CheckpointingConfigError):
        _chunks(4, "group_0")
    with pytest.raises(CheckpointingConfigError):
        _chunks(4, "every_other")
    with pytest.raises(CheckpointingConfigError):
        checkpoint_sequential(nn.Linear(2, 2), torch.zeros(1, 2))


def test_set_activation_checkpointing_through_distributed_model():
    smp.reset()
    smp.init({"microbatches": 2})
    torch.manual_seed(4)
    net = nn.Sequential(nn.Linear(8, 8), nn.Sequential(nn.Linear(8, 8), nn.Tanh(), nn.Linear(8, 8)), nn.Linear(8, 2))
    ref_net = nn.Sequential(nn.Linear(8, 8), nn.Sequential(nn.Linear(8, 8), nn.Tanh(), nn.Linear(8, 8)), nn.Linear(8, 2))
    ref_net.load_state_dict(net.state_dict())
    smp.set_activation_checkpointing(net[1], strategy="contiguous")
    model = smp.DistributedModel(net)
    calls = {"n": 0}
    net[1][0].register_forward_hook(lambda *a: calls.__setitem__("n", calls["n"] + 1))

    @smp.step
    def train_step(model, x):
        loss = model(x).square().mean()
        model.backward(loss)
        return loss

    x = torch.randn(4, 8)
    train_step(model, x)
    assert calls["n"] == 4                                   # 2 microbatches x (forward + recomputation)
    ref_net(x).square().mean().backward()                    # mean over the batch = mean of the microbatch means
    for a, b in zip(ref_net.parameters(), model.module.parameters()):
        assert torch.allclose(a.grad, b.grad, atol=1e-6)


def test_checkpointed_hf_blocks_ignore_the_kv_cache():
    """A stock HuggingFace GPT-2 hands every block a mutable key/value cache (``use_cache`` defaults to true); the
    recomputation would append to it a second time and attend over doubled keys.  Checkpointed calls drop the cache,
    so gradients equal those of the plain model."""
    import copy

    from transformers import GPT2Config, GPT2LMHeadModel

    smp.reset()
    smp.init({"microbatches": 2})
    torch.manual_seed(1)
    cfg = GPT2Config(n_layer=2, n_head=4, n_embd=32, vocab_size=64, n_positions=16, resid_pdrop=0.0, embd_pdrop=0.0,
                     attn_pdrop=0.0)
    net = GPT2LMHeadModel(cfg)
    oracle = copy.deepcopy(net)
    for blk in net.transformer.h:
        smp.set_activation_checkpointing(blk)
    model = smp.DistributedModel(net)

    @smp.step
    def step(model, ids):
        out = model(input_ids=ids, labels=ids)             # use_cache left at its default
        model.backward(out.loss)
        return out.loss

    ids = torch.randint(0, 64, (4, 16), generator=torch.Generator().manual_seed(3))
    loss = step(model, ids).reduce_mean().item()
    ol = sum(oracle(input_ids=ids[i:i + 2], labels=ids[i:i + 2]).loss / 2 for i in (0, 2))
    ol.backward()
    ref = dict(oracle.named_parameters())
    worst = max((p.grad - ref[n].grad).abs().max().item() for n, p in net.named_parameters())
    assert abs(loss - ol.item()) < 5e-6 and worst < 5e-6, (loss, ol.item(), worst)


def test_checkpoint_attentions_wraps_the_library_attention_path(monkeypatch):
    """``checkpoint_attentions`` (reference torch/nn/transformer.py:1487-1493): where attention runs as separate
    library ops (CPU here; additive position bias on the GPU) the core is recomputed in backward instead of keeping the
    [B, heads, Sq, Sk] probabilities; gradients are unchanged."""
    import torch.utils.checkpoint as tuc

    import smdistributed_modelparallel_b200 as smp

    kw = dict(num_attention_heads=2, attention_head_size=8, hidden_size=16, intermediate_size=32,
              attention_dropout_prob=0.0, hidden_dropout_prob=0.0, pre_layernorm=True, post_layernorm=False,
              causal_mask_size=8, use_normal_initialization=True)
    x0 = torch.randn(2, 8, 16)
    calls = []
    real = tuc.checkpoint
    monkeypatch.setattr(tuc, "checkpoint", lambda fn, *a, **k: (calls.append(1), real(fn, *a, **k))[1])
    grads = []
    for flag in (False, True):
        smp.reset()
        smp.init({"microbatches": 1, "checkpoint_attentions": flag})
        torch.manual_seed(0)
        layer = smp.nn.DistributedTransformerLayer(**kw)
        x = x0.clone().requires_grad_(True)
        layer((x, None))[0].square().sum().backward()
        grads.append((x.grad.clone(), layer.attention.qkv_weight.grad.clone() if layer.attention.qkv_weight.grad is not None
                      else layer.attention.qkv_weight.main_grad.clone()))
        assert len(calls) == (1 if flag else 0)
    assert torch.allclose(grads[0][0], grads[1][0], atol=1e-6)
    assert torch.allclose(grads[0][1], grads[1][1], atol=1e-6)


def test_recomputation_runs_under_the_autocast_mode_of_the_forward_pass():
    """backward() is called outside the autocast region; the recomputed forward must still use it, otherwise the
    recomputed activations (and gradients) are those of another dtype."""
    import smdistributed_modelparallel_b200 as smp

    smp.reset()
    smp.init({"microbatches": 1})
    seen = []

    def fn(x, w):
        y = x @ w
        seen.append(y.dtype)
        return y.float().sum(dim=-1)

    x = torch.randn(4, 8, requires_grad=True)
    w = torch.randn(8, 8, requires_grad=True)
    with torch.autocast("cpu", dtype=torch.bfloat16):
        out = smp.checkpoint(fn, x, w)
    out.sum().backward()
    assert seen == [torch.bfloat16, torch.bfloat16], seen
    assert x.grad is not None and w.grad is not None
