"""``with smp.partition(i):`` inside module constructors, the placement patterns of the reference's backward tier
(reference test/torch/model_zoo/backward_models.py:38-204 run by test/torch/mpi/test_backward.py): a remote module
between local ones, a nested module placed as a whole, a remote child inside a local parent, two remote modules in a
row (one hand-over each way per module), a chain, everything on one stage.  Loss, every local gradient and the
placement itself are checked against a single-process copy on two gloo ranks."""

import copy

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from dist_utils import run_distributed


def _build(kind, smp):
    class Leaf(nn.Module):
        def __init__(self):
            super().__init__()
            self.m = nn.Linear(10, 10)

        def forward(self, x):
            return self.m(x)

    class RemoteChild(nn.Module):                       # the parent stays where it is, its child goes to stage 1
        def __init__(self):
            super().__init__()
            with smp.partition(1):
                self.m = nn.Linear(10, 10)

        def forward(self, x):
            return self.m(x)

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.a = nn.Linear(10, 10)
            if kind == "simple":
                with smp.partition(1):
                    self.b = nn.Linear(10, 10)
                self.c = nn.Linear(10, 10)
                self.order = ["a", "relu", "b", "c", "relu"]
                self.want = {"a": 0, "b": 1, "c": 0}
            elif kind == "nest0":
                with smp.partition(1):
                    self.b = Leaf()
                self.order = ["a", "b"]
                self.want = {"a": 0, "b": 1, "b/m": 1}
            elif kind == "nest1":
                with smp.partition(1):
                    self.nest1 = Leaf()
                self.nest2 = RemoteChild()
                self.order = ["a", "nest1", "nest2"]
                self.want = {"a": 0, "nest1": 1, "nest1/m": 1, "nest2": 0, "nest2/m": 1}
            elif kind == "nest2":
                with smp.partition(1):
                    self.b = Leaf()
                with smp.partition(1):
                    self.c = Leaf()
                self.order = ["a", "b", "c"]
                self.want = {"a": 0, "b": 1, "c": 1, "c/m": 1}
            elif kind == "chain":
                with smp.partition(1):
                    self.b = nn.Linear(10, 10)
                with smp.partition(1):
                    self.c = nn.Linear(10, 10)
                self.d = nn.Linear(10, 10)
                self.order = ["a", "b", "c", "d"]
                self.want = {"a": 0, "b": 1, "c": 1, "d": 0}
            else:                                       # everything on the default stage
                self.b, self.c, self.d = nn.Linear(10, 10), nn.Linear(10, 10), nn.Linear(10, 10)
                self.order = ["a", "b", "c", "d"]
                self.want = {"a": 0, "b": 0, "c": 0, "d": 0}

        def forward(self, x):
            for name in self.order:
                x = F.relu(x) if name == "relu" else getattr(self, name)(x)
            return x

    return Net()


def _worker(rank, world, kind):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 2, "ddp": False, "auto_partition": False,
              "default_partition": 0})
    torch.manual_seed(3)
    net = _build(kind, smp)
    oracle = copy.deepcopy(net)
    model = smp.DistributedModel(net)

    @smp.step
    def train_step(model, x):
        loss = model(x).square().mean()
        model.backward(loss)
        return loss

    x = torch.randn(20, 10, generator=torch.Generator().manual_seed(9))
    loss = train_step(model, x).reduce_mean().item()
    ol = sum(oracle(x[i:i + 10]).square().mean() / 2 for i in (0, 10))
    ol.backward()
    ref = dict(oracle.named_parameters())
    worst, n_local = 0.0, 0
    for n, p in model.local_named_parameters():
        assert p.grad is not None, n
        worst = max(worst, (p.grad - ref[n].grad).abs().max().item())
        n_local += 1
    mm = smp.state.module_manager
    placed = {k: mm.get_partition_by_name("main/" + k) for k in net.want}
    return {"loss": loss, "oracle": ol.item(), "worst": worst, "n_local": n_local, "placed": placed, "want": net.want,
            "hops": mm.hops}


@pytest.mark.parametrize("kind", ["simple", "nest0", "nest1", "nest2", "chain", "one_stage"])
def test_partition_contexts_place_modules_and_gradients_match(kind):
    res = run_distributed(_worker, 2, kind, timeout=300)
    for r in res:
        assert r["placed"] == r["want"], (r["placed"], r["want"])
        assert abs(r["loss"] - r["oracle"]) < 5e-6 and r["worst"] < 5e-6, r
    if kind == "one_stage":
        assert res[1]["n_local"] == 0 and res[0]["hops"] <= 1      # only the step's output travels to the idle stage
    else:
        assert res[1]["n_local"] > 0 and res[0]["hops"] > 1
