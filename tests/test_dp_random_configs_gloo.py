"""Random data-parallel configurations on four gloo ranks: tensor-parallel degree 1 / 2 (reduced-data-parallel groups),
fp32 (per-parameter hooks, native gradient counter, planned buckets) or bf16 (contiguous fp32 gradient buffer, native
bucket reducer), microbatches, bucket cap, overlap on / off, optimizer-state sharding, SGD / AdamW.  Parameters after
three steps are compared with the dense model on the union batch (reference test/torch/mpi_hybrid/test_allreduce_ddp.py,
test_opt_sharding.py).  ``SMP_RANDOM_DP_SEEDS=a-b`` widens the search."""

import os
import random

import pytest
import torch

from dist_utils import run_distributed

CFG = dict(num_layers=3, num_attention_heads=4, attention_head_size=4, hidden_size=16, intermediate_size=40,
           vocab_size=48, num_positions=12, attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
           embedding_dropout_prob=0.0, activation="gelu", causal_mask_size=12, pre_layernorm=True, post_layernorm=False,
           final_layernorm=True, use_normal_initialization=True)


def options(seed):
    rng = random.Random(seed)
    return {"tp": rng.choice([1, 1, 2]), "bf16": rng.random() < 0.5, "microbatches": rng.choice([1, 2, 4]),
            "bucket_mb": rng.choice([0.001, 0.01, 100]), "overlap": rng.random() < 0.7,
            "shard": rng.random() < 0.4, "adam": rng.random() < 0.6, "min_elems": rng.choice([100, 1 << 18])}


def _worker(rank, world, seed):
    import smdistributed_modelparallel_b200 as smp

    o = options(seed)
    os.environ["SMP_DDP_BUCKET_MIN_ELEMS"] = str(o["min_elems"])
    smp.init({"tensor_parallel_degree": o["tp"], "ddp": True, "ddp_dist_backend": "gloo", "microbatches": o["microbatches"],
              "bf16": o["bf16"], "shard_optimizer_state": o["shard"]})
    torch.manual_seed(3)
    with smp.model_creation(tensor_parallelism=o["tp"] > 1, dtype=torch.bfloat16 if o["bf16"] else None):
        module = smp.nn.DistributedTransformerLMHead(**CFG)
    model = smp.DistributedModel(module, overlapping_allreduce=o["overlap"], bucket_cap_mb=o["bucket_mb"])
    make = (lambda p: torch.optim.AdamW(p, lr=1e-2, weight_decay=0.01)) if o["adam"] else \
        (lambda p: torch.optim.SGD(p, lr=0.1))
    opt = smp.DistributedOptimizer(make(model.parameters()))
    start = {k: v.float().clone() for k, v in model.state_dict().items()}

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    per = 4
    all_ids = torch.randint(0, CFG["vocab_size"], (3, per * world, 12), generator=torch.Generator().manual_seed(seed))
    mine = smp.dp_rank()
    for it in range(3):
        opt.zero_grad()
        train_step(model, all_ids[it, mine * per:(mine + 1) * per])
        opt.step()
    return {"start": start, "end": {k: v.float().clone() for k, v in model.state_dict().items()}, "all_ids": all_ids,
            "native": type(getattr(model.reducer, "_native", None)).__name__ if model.reducer is not None else None}


def _seeds():
    spec = os.environ.get("SMP_RANDOM_DP_SEEDS", "")
    if spec:
        a, b = spec.split("-")
        return list(range(int(a), int(b) + 1))
    return [1, 3, 5, 9, 12, 18]


@pytest.mark.parametrize("seed", _seeds())
def test_random_data_parallel_configurations_track_dense(seed):
    import smdistributed_modelparallel_b200  # noqa: F401
    from smdistributed_modelparallel_b200.nn.transformer import DistributedTransformerLMHead

    o = options(seed)
    res = run_distributed(_worker, 4, seed, timeout=300)
    dense = DistributedTransformerLMHead(**CFG)
    dense.load_state_dict(res[0]["start"], strict=False)
    opt = torch.optim.AdamW(dense.parameters(), lr=1e-2, weight_decay=0.01) if o["adam"] else \
        torch.optim.SGD(dense.parameters(), lr=0.1)
    for it in range(3):
        ids = res[0]["all_ids"][it].long()
        opt.zero_grad()
        dense((ids, None, None, None, ids))[0].backward()
        opt.step()
    dsd = dense.state_dict()
    tol = 6e-2 if o["bf16"] else 2e-3
    for r in res:
        worst = max((r["end"][k] - dsd[k]).abs().max().item() for k in dsd if k in r["end"])
        assert worst < tol, (seed, o, worst)
