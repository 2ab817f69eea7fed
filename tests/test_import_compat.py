"""A script written against ``smdistributed.modelparallel.torch`` runs unchanged (import-path shim)."""

import torch


def test_reference_import_path():
    import smdistributed.modelparallel.torch as smp
    import smdistributed_modelparallel_b200 as impl

    assert smp.DistributedModel is impl.DistributedModel and smp.step is impl.step
    assert smp.nn.DistributedTransformerLMHead is impl.nn.DistributedTransformerLMHead
    from smdistributed.modelparallel.torch.nn import DistributedLinear  # noqa: F401
    from smdistributed.modelparallel.torch.optimizers import FusedAdam  # noqa: F401

    smp.reset()
    smp.init({"microbatches": 2})
    model = smp.DistributedModel(torch.nn.Linear(4, 2))
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))

    @smp.step
    def train_step(model, x):
        loss = model(x).sum()
        model.backward(loss)
        return loss

    out = train_step(model, torch.ones(4, 4))
    opt.step()
    assert smp.rank() == 0 and smp.size() == 1 and len(out) == 2
