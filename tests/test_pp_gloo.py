"""Pipeline parallelism on CPU/gloo, world size 2: GPT-2-style model, pp=2, 4 microbatches.  Loss and every
parameter gradient must match the dense single-process model (the reference's oracle design,
reference test/torch/smp_test_base.py:603-639; BASELINE config #1)."""

import pytest
import torch

from dist_utils import run_distributed

CFG = dict(num_layers=4, num_attention_heads=4, attention_head_size=8, hidden_size=32, intermediate_size=64,
           vocab_size=64, num_positions=16, attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
           embedding_dropout_prob=0.0, activation="gelu_new", causal_mask_size=16, pre_layernorm=True,
           post_layernorm=False, final_layernorm=True, use_normal_initialization=True)


def _worker(rank, world, pipeline, microbatches):
    import copy

    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "microbatches": microbatches, "pipeline": pipeline, "ddp": False})
    torch.manual_seed(7)
    module = smp.nn.DistributedTransformerLMHead(**CFG)
    oracle = copy.deepcopy(module)
    model = smp.DistributedModel(module)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    torch.manual_seed(11)
    ids = torch.randint(0, CFG["vocab_size"], (8, 16))
    opt.zero_grad()
    out = train_step(model, ids)
    loss = out.reduce_mean().item()
    # oracle: same batch in one piece
    ol, _ = oracle((ids, None, None, None, ids))
    ol.backward()
    ref = {n: p.grad for n, p in oracle.named_parameters()}
    worst = 0.0
    n_local = 0
    for n, p in model.local_named_parameters():
        assert p.grad is not None, f"{n} has no grad on pp_rank {smp.pp_rank()}"
        worst = max(worst, (p.grad - ref[n]).abs().max().item())
        n_local += 1
    part = smp.state.module_manager.partition_info()
    opt.step()
    return {"loss": loss, "oracle": ol.item(), "worst_grad_err": worst, "n_local": n_local,
            "ranks_used": sorted(set(part.values())), "hops": smp.state.module_manager.hops}


@pytest.mark.parametrize("pipeline", ["interleaved", "simple"])
def test_pp2_matches_dense(pipeline):
    res = run_distributed(_worker, 2, pipeline, 4)
    for r in res:
        assert abs(r["loss"] - r["oracle"]) < 1e-5, r
        assert r["worst_grad_err"] < 1e-5, r
        assert r["ranks_used"] == [0, 1]
        assert r["n_local"] > 0
    assert sum(r["n_local"] for r in res) > 0
