"""Pipeline parallelism end to end on tiny hand-written networks, 2 CPU/gloo ranks, against the same network trained
in one process (reference test/torch/mpi/test_e2e.py: module reuse, nn.Sequential spanning partitions, buffers,
manual and nested partitions, several outputs)."""

import pytest
import torch
import torch.nn as nn

from dist_utils import run_distributed


class Reuse(nn.Module):
    """The same child is called twice; children live on different ranks."""

    def __init__(self):
        super().__init__()
        self.a = nn.Linear(8, 8)
        self.b = nn.Linear(8, 8)
        self.out = nn.Linear(8, 3)

    def forward(self, x):
        h = torch.tanh(self.a(x))
        h = torch.tanh(self.b(h)) + h
        h = torch.tanh(self.a(h))            # reuse of `a`: two gradient contributions per microbatch
        return self.out(h)


class WithBuffers(nn.Module):
    def __init__(self):
        super().__init__()
        self.f = nn.Sequential(nn.Linear(8, 16), nn.BatchNorm1d(16), nn.ReLU(), nn.Linear(16, 16), nn.ReLU(),
                               nn.Linear(16, 8), nn.Tanh(), nn.Linear(8, 3))

    def forward(self, x):
        return self.f(x)


class TwoHeads(nn.Module):
    def __init__(self):
        super().__init__()
        self.trunk = nn.Sequential(nn.Linear(8, 8), nn.Tanh())
        self.h1 = nn.Linear(8, 3)
        self.h2 = nn.Sequential(nn.Linear(8, 8), nn.Tanh(), nn.Linear(8, 3))

    def forward(self, x, scale=1.0):
        t = self.trunk(x)
        return self.h1(t) * scale, {"aux": self.h2(t)}


class Unused(nn.Module):
    """A child that never runs and an input that is ignored (the reference documents both as unsupported graphs,
    test/torch/mpi/xfails; here they simply produce no gradient)."""

    def __init__(self):
        super().__init__()
        self.a = nn.Sequential(nn.Linear(8, 8), nn.Tanh())
        self.dead = nn.Linear(8, 8)
        self.b = nn.Linear(8, 3)

    def forward(self, x, ignored=None):
        return self.b(self.a(x))


def _build(kind, smp=None):
    torch.manual_seed(11)
    if kind == "reuse":
        net = Reuse()
        if smp is not None:
            smp.set_partition(net.a, 0)
            smp.set_partition(net.b, 1)
            smp.set_partition(net.out, 1)
    elif kind == "buffers":
        net = WithBuffers()
    elif kind == "unused":
        net = Unused()
        if smp is not None:
            smp.set_partition(net.a, 0)
            smp.set_partition(net.dead, 1)
            smp.set_partition(net.b, 1)
    else:
        net = TwoHeads()
        if smp is not None:
            smp.set_partition(net.trunk, 0)
            smp.set_partition(net.h1, 0)
            smp.set_partition(net.h2, 1)
    return net


def _loss(kind, model, x, y):
    if kind == "heads":
        a, d = model(x, scale=2.0)
        return ((a - y) ** 2).mean() + (d["aux"] ** 2).mean()
    if kind == "unused":
        return ((model(x, ignored=y) - y) ** 2).mean()
    return ((model(x) - y) ** 2).mean()


def _worker(rank, world, kind, auto):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 2, "ddp": False, "auto_partition": auto,
              "default_partition": 0})
    net = _build(kind, None if auto else smp)
    model = smp.DistributedModel(net)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))

    @smp.step
    def train_step(model, x, y):
        loss = _loss(kind, model, x, y)
        model.backward(loss)
        return loss

    torch.manual_seed(5)
    losses = []
    for _ in range(3):
        x, y = torch.randn(8, 8), torch.randn(8, 3)
        opt.zero_grad()
        out = train_step(model, x, y)
        opt.step()
        losses.append(out.reduce_mean().item())
    sd = {k: v.clone() for k, v in model.state_dict().items()}        # gathered over the pipeline ranks
    return {"losses": losses, "sd": sd, "parts": sorted(set(smp.state.module_manager.partition_info().values()))}


@pytest.mark.parametrize("kind,auto", [("reuse", False), ("buffers", True), ("heads", False), ("unused", False)])
def test_pp2_toy_networks_match_single_process(kind, auto):
    res = run_distributed(_worker, 2, kind, auto)
    ref = _build(kind)
    opt = torch.optim.SGD(ref.parameters(), lr=0.1)
    torch.manual_seed(5)
    want = []
    for _ in range(3):
        x, y = torch.randn(8, 8), torch.randn(8, 3)
        opt.zero_grad()
        # two microbatches of 4: batch statistics (BatchNorm) are per microbatch, losses are averaged
        total = 0.0
        for mb in range(2):
            l = _loss(kind, ref, x[mb * 4:(mb + 1) * 4], y[mb * 4:(mb + 1) * 4]) / 2
            l.backward()
            total += l.item()
        opt.step()
        want.append(total)
    for r in res:
        assert r["parts"] == [0, 1], r["parts"]
        for a, b in zip(r["losses"], want):
            assert abs(a - b) < 5e-5, (r["losses"], want)
    full = res[0]["sd"]
    for k, v in ref.state_dict().items():
        assert k in full, k
        assert torch.allclose(full[k].float(), v.float(), atol=1e-5), k
