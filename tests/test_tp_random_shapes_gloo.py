"""Tensor-parallel transformer layers at random (mostly uneven) shapes: heads, hidden and intermediate sizes that do
not divide the tensor-parallel degree, pre / post LayerNorm, both activation layouts (``optimize: speed`` and
``optimize: memory``), degrees 2-4.  Output and input gradient of a two-layer stack are compared with the same stack
built densely from the gathered state dict (the reference pins one such case, H = 845 with 13 heads on 4 ranks:
test/torch/mpi_4ps/test_dist_modules.py:81-130).  ``SMP_RANDOM_SHAPE_SEEDS=a-b`` widens the search."""

import os
import random

import pytest
import torch

from dist_utils import run_distributed


def config(seed):
    rng = random.Random(seed)
    tp = rng.choice([2, 3, 4])
    heads = rng.randint(tp, 9)
    d = rng.choice([2, 4, 6])
    hidden = rng.randint(max(tp, 5), 30)
    kw = dict(num_attention_heads=heads, attention_head_size=d, hidden_size=hidden,
              intermediate_size=rng.randint(tp, 40), attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
              activation=rng.choice(["gelu", "relu"]), pre_layernorm=rng.random() < 0.6, post_layernorm=True,
              causal_mask_size=8 if rng.random() < 0.7 else None, use_normal_initialization=True)
    if not kw["pre_layernorm"]:
        kw["post_layernorm"] = True
    return tp, rng.choice(["speed", "memory"]), kw, rng.randint(1, 3)


def _worker(rank, world, seed):
    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.nn.transformer import DistributedTransformer

    tp, layout, kw, per_rank = config(seed)
    smp.init({"tensor_parallel_degree": tp, "ddp": True, "ddp_dist_backend": "gloo", "optimize": layout})
    torch.manual_seed(seed)
    with smp.tensor_parallelism(enabled=True):
        stack = DistributedTransformer(num_layers=2, **kw)
    data = torch.Generator().manual_seed(1000 + seed)
    h_all = torch.randn(per_rank * tp, 8, kw["hidden_size"], generator=data)
    h = h_all[rank * per_rank:(rank + 1) * per_rank].clone().requires_grad_(True)
    y = stack((h, None))[0]
    y.square().sum().backward()
    model = smp.DistributedModel(stack)
    return {"y": y.detach(), "dx": h.grad.clone(), "h_all": h_all,
            "sd": {k: v.clone() for k, v in model.state_dict().items()}}


def _seeds():
    spec = os.environ.get("SMP_RANDOM_SHAPE_SEEDS", "")
    if spec:
        a, b = spec.split("-")
        return list(range(int(a), int(b) + 1))
    return [1, 2, 5, 8, 13, 21]


@pytest.mark.parametrize("seed", _seeds())
def test_random_uneven_tensor_parallel_stacks(seed):
    import smdistributed_modelparallel_b200  # noqa: F401
    from smdistributed_modelparallel_b200.nn.transformer import DistributedTransformer

    tp, layout, kw, per_rank = config(seed)
    res = run_distributed(_worker, tp, seed, timeout=240)
    dense = DistributedTransformer(num_layers=2, **kw)
    missing = dense.load_state_dict(res[0]["sd"], strict=False)
    assert not missing.missing_keys and not missing.unexpected_keys, (seed, missing)
    ha = res[0]["h_all"].clone().requires_grad_(True)
    yd = dense((ha, None))[0]
    yd.square().sum().backward()
    for r, rr in enumerate(res):
        sl = slice(r * per_rank, (r + 1) * per_rank)
        assert torch.allclose(rr["y"], yd[sl].detach(), atol=5e-5), (seed, tp, layout, kw, (rr["y"] - yd[sl]).abs().max())
        assert torch.allclose(rr["dx"], ha.grad[sl], atol=5e-5), (seed, tp, layout, (rr["dx"] - ha.grad[sl]).abs().max())
