"""User registrations in the tensor-parallelism registry on two gloo ranks: ``@smp.tp_register``, the same with an
init hook that converts constructor arguments, and ``smp.tp_register_with_module`` for a class one does not own -
the marked instance is replaced by its distributed counterpart, nothing of the original class is left in the model, and
the replaced model still trains (reference test/torch/mpi/test_tp_registry.py:14-93)."""

import pytest
import torch

from dist_utils import run_distributed


def _worker(rank, world, how):
    import torch.nn as nn

    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 1, "tensor_parallel_degree": 2, "ddp": True, "ddp_dist_backend": "gloo"})
    smp.state.tp_registry.reset()

    def body(cls):
        def forward(self, x):
            return self.id2(self.lin(self.id1(x)))
        cls.forward = forward
        return cls

    if how == "decorator":
        @smp.tp_register(smp.nn.DistributedLinear)
        class Custom(nn.Module):
            def __init__(self, in_features, out_features):
                super().__init__()
                self.id1, self.id2, self.lin = nn.Identity(), nn.Identity(), nn.Linear(in_features, out_features)
        make = lambda: body(Custom)(64, 64)                          # noqa: E731
    else:
        class Custom(nn.Module):
            def __init__(self, features):
                super().__init__()
                self.id1, self.id2, self.lin = nn.Identity(), nn.Identity(), nn.Linear(features, features)
        hook = lambda x: ((x, x), {})                                # noqa: E731  features -> (in_features, out_features)
        if how == "decorator_hook":
            Custom = smp.tp_register(smp.nn.DistributedLinear, init_hook=hook)(Custom)
        else:
            smp.tp_register_with_module(Custom, smp.nn.DistributedLinear, init_hook=hook)
        make = lambda: body(Custom)(64)                              # noqa: E731

    class Model(nn.Module):
        def __init__(self):
            super().__init__()
            self.custom_module = make()
            self.lin1, self.lin2 = nn.Linear(64, 64), nn.Linear(64, 64)

        def forward(self, x):
            return self.lin2(self.lin1(self.custom_module(x)))

    torch.manual_seed(0)
    net = Model()
    smp.set_tensor_parallelism(net.custom_module, True)
    model = smp.DistributedModel(net)
    replaced = type(model.module.custom_module).__name__
    leftovers = sum(isinstance(m, Custom) for m in model.modules())
    w = tuple(model.module.custom_module.weight.shape)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.05))

    @smp.step
    def train_step(model, x):
        loss = model(x).square().mean()
        model.backward(loss)
        return loss

    x = torch.randn(4, 64, generator=torch.Generator().manual_seed(rank))
    losses = []
    for _ in range(3):
        opt.zero_grad()
        losses.append(train_step(model, x).reduce_mean().item())
        opt.step()
    return {"replaced": replaced, "leftovers": leftovers, "weight": w, "losses": losses,
            "plain": type(model.module.lin1).__name__}


@pytest.mark.parametrize("how", ["decorator", "decorator_hook", "with_module"])
def test_registered_module_is_distributed(how):
    for r in run_distributed(_worker, 2, how, timeout=300):
        assert r["replaced"] == "DistributedLinear" and r["leftovers"] == 0
        assert r["weight"] == (64, 32)                      # input features split over the two ranks
        assert r["plain"] == "Linear"                       # unmarked modules stay as they are
        assert r["losses"][-1] < r["losses"][0]
