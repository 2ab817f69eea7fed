"""DistributedModel accessors: names as the wrapped module sees them, local / reducer / virtual parameter views,
buffers, partition display (reference model.py:348-349, 668-701, 1098-1108, 1371-1436)."""

import pytest
import torch
import torch.nn as nn

import smdistributed_modelparallel_b200 as smp
from smdistributed_modelparallel_b200.backend.exceptions import SMPUnsupportedError

from dist_utils import run_distributed


class _Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.a = nn.Linear(8, 8)
        self.bn = nn.BatchNorm1d(8)
        self.b = nn.Linear(8, 2)

    def forward(self, x):
        return self.b(self.bn(torch.tanh(self.a(x))))


def test_accessors_single_process():
    smp.reset()
    smp.init({"microbatches": 2})
    torch.manual_seed(0)
    net = _Net()
    model = smp.DistributedModel(net)
    assert [n for n, _ in model.named_parameters()] == [n for n, _ in net.named_parameters()]
    assert [n for n, _ in model.named_buffers()] == [n for n, _ in net.named_buffers()]
    assert len(list(model.parameters())) == 6 and len(list(model.buffers())) == 3
    assert model.get_module_for_param(net.b.weight) is net.b
    with pytest.raises(KeyError):
        model.get_module_for_param(nn.Parameter(torch.zeros(1)))
    assert model.is_local_buffer("bn.running_mean") and model.is_local_buffer(net.bn.running_var)
    assert not model.is_local_buffer("nope")
    with pytest.raises(SMPUnsupportedError):
        model.is_local_buffer(3)
    assert [n for n, _ in model.default_reducer_named_parameters()] == [n for n, _ in net.named_parameters()]
    assert list(model.scaled_batch_reducer_named_parameters()) == []
    assert list(model.virtual_named_parameters()) == []
    assert model.get_ddp_logging_data() == {}
    lines = model.display_partition()
    assert lines == ["main: 0"]


def _pp_worker(rank, world):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 2, "ddp": False, "auto_partition": False,
              "default_partition": 0})
    torch.manual_seed(0)
    net = _Net()
    smp.set_partition(net.b, 1)
    smp.set_partition(net.bn, 1)
    model = smp.DistributedModel(net)

    @smp.step
    def step(model, x):
        out = model(x)
        loss = out.pow(2).mean()
        model.backward(loss)
        return loss

    step(model, torch.randn(4, 8))
    return {"lines": model.display_partition(), "bn_local": model.is_local_buffer("bn.running_mean"),
            "local": [n for n, _ in model.local_named_parameters()],
            "all": [n for n, _ in model.named_parameters()]}


def test_partition_display_and_local_views_under_pp():
    res = run_distributed(_pp_worker, 2, timeout=180)
    assert res[0]["lines"][0].startswith("main: 0 (children on [0, 1])")
    assert "  a: 0" in res[0]["lines"] and "  b: 1" in res[0]["lines"] and "  bn: 1" in res[0]["lines"]
    assert res[0]["bn_local"] is False and res[1]["bn_local"] is True
    assert res[0]["local"] == ["a.weight", "a.bias"]
    assert res[1]["local"] == ["bn.weight", "bn.bias", "b.weight", "b.bias"]
    assert res[0]["all"] == res[1]["all"] and len(res[0]["all"]) == 6


def test_module_manager_accessors_under_the_reference_names():
    import torch.nn as nn

    import smdistributed_modelparallel_b200 as smp

    smp.reset()
    smp.init({"microbatches": 1})
    net = nn.Sequential(nn.Linear(4, 4), nn.Sequential(nn.Linear(4, 4), nn.ReLU()))
    model = smp.DistributedModel(net)
    mm = smp.state.module_manager
    assert mm.get_module_names()[0] == "main" and "main/1/0" in mm.get_module_names()
    assert mm.is_main_module(model.module) and not mm.is_main_module(net[1])
    assert mm.get_parent_module(net[1][0]) is net[1] and mm.get_parent_module(model.module) is None
    assert len(mm.get_parameters(net[1])) == 2 and len(mm.get_parameters(net, recurse=False)) == 0
    p = net[0].weight
    assert not mm.is_distributed(p) and not mm.is_scaled_batch_parameter(p) and not mm.is_one_rank_parameter(p)
    assert mm.get_parameter_distribution_axis(p) is None
    lin = smp.nn.DistributedLinear(4, 4)
    assert mm.is_distributed_parameter(lin.weight) and mm.get_parameter_distribution_axis(lin.weight) == 1
    assert mm.is_scaled_batch_parameter(lin.weight)
    assert set(mm.get_model_partition_info().values()) == {0} and smp.state.current_minibatch() == 0   # pp = 1: all on rank 0
    assert smp.get_dp_group() == [0] and smp.get_pp_group() == [0] and smp.get_tp_group() == [0]
