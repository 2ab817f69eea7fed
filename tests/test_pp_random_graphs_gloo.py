"""Randomly generated module trees with random manual placements on three pipeline stages: chains, nested containers,
residual branches, a module used twice, a module with two outputs, a stage that owns nothing.  Loss and every local
gradient are compared with a single-process copy.  The seeds below are fixed; ``SMP_RANDOM_GRAPH_SEEDS=a-b`` widens the
search (the engine places any ``nn.Module`` anywhere - reference torch/module_manager.py, patches/execution.py - so the
structure of the tree, not a layer list, is what has to be exercised)."""

import copy
import os
import random

import pytest
import torch
import torch.nn as nn

from dist_utils import run_distributed

DIM = 6


class Residual(nn.Module):
    def __init__(self, inner):
        super().__init__()
        self.inner = inner

    def forward(self, x):
        return x + self.inner(x)


class TwoOut(nn.Module):
    def __init__(self):
        super().__init__()
        self.a, self.b = nn.Linear(DIM, DIM), nn.Linear(DIM, DIM)

    def forward(self, x):
        return self.a(x), {"aux": torch.tanh(self.b(x))}


class Join(nn.Module):
    def __init__(self, two):
        super().__init__()
        self.two = two
        self.mix = nn.Linear(DIM, DIM)

    def forward(self, x):
        y, d = self.two(x)
        return self.mix(y) + d["aux"]


class Twice(nn.Module):
    def __init__(self, shared):
        super().__init__()
        self.shared = shared

    def forward(self, x):
        return self.shared(torch.relu(self.shared(x)))


def _make(rng, depth=0):
    kind = rng.choice(["lin", "lin", "seq", "res", "join", "twice"] if depth < 2 else ["lin"])
    if kind == "lin":
        return nn.Sequential(nn.Linear(DIM, DIM), nn.Tanh()) if rng.random() < 0.5 else nn.Linear(DIM, DIM)
    if kind == "seq":
        return nn.Sequential(*[_make(rng, depth + 1) for _ in range(rng.randint(2, 3))])
    if kind == "res":
        return Residual(_make(rng, depth + 1))
    if kind == "join":
        return Join(TwoOut())
    return Twice(_make(rng, depth + 1))


def build(seed):
    rng = random.Random(seed)
    torch.manual_seed(seed)
    net = nn.Sequential(*[_make(rng) for _ in range(rng.randint(3, 5))])
    # a placement for a random subset of the modules (children inherit from the closest placed ancestor)
    placement = {}
    for name, m in net.named_modules():
        if name and rng.random() < 0.45:
            placement[name] = rng.randrange(3)
    return net, placement


def _worker(rank, world, seed):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 3, "microbatches": 2, "ddp": False, "auto_partition": False,
              "default_partition": 0})
    net, placement = build(seed)
    oracle = copy.deepcopy(net)
    model = smp.DistributedModel(net)
    mods = dict(net.named_modules())
    for name, part in placement.items():
        smp.set_partition(mods[name], part)

    @smp.step
    def train_step(model, x):
        loss = model(x).square().mean()
        model.backward(loss)
        return loss

    x = torch.randn(8, DIM, generator=torch.Generator().manual_seed(seed + 1))
    loss = train_step(model, x).reduce_mean().item()
    ol = sum(oracle(x[i:i + 4]).square().mean() / 2 for i in (0, 4))
    ol.backward()
    ref = dict(oracle.named_parameters())
    worst, n_local = 0.0, 0
    for n, p in model.local_named_parameters():
        assert p.grad is not None, (seed, n)
        worst = max(worst, (p.grad - ref[n].grad).abs().max().item())
        n_local += 1
    return {"loss": loss, "oracle": ol.item(), "worst": worst, "n_local": n_local,
            "stages": sorted(set(smp.state.module_manager.partition_info().values()))}


def _seeds():
    spec = os.environ.get("SMP_RANDOM_GRAPH_SEEDS", "")
    if spec:
        a, b = spec.split("-")
        return list(range(int(a), int(b) + 1))
    return [3, 11, 17, 23, 42, 57]


@pytest.mark.parametrize("seed", _seeds())
def test_random_module_trees_on_three_stages(seed):
    res = run_distributed(_worker, 3, seed, timeout=180)
    total = sum(r["n_local"] for r in res)
    assert total == len(list(build(seed)[0].parameters())), (total, seed)      # every parameter has exactly one owner
    for r in res:
        assert abs(r["loss"] - r["oracle"]) < 5e-6 and r["worst"] < 5e-6, (seed, r)
