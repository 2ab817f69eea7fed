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


def test_reference_keyword_names_of_the_distributed_modules():
    """Constructor keywords as the reference spells them (nn/linear.py:25, nn/layer_norm.py:108,
    nn/embedding.py:40, fp16/fp16.py Bit16_Optimizer)."""
    import smdistributed.modelparallel.torch as smp
    from smdistributed_modelparallel_b200.precision.bit16 import Bit16_Optimizer

    smp.reset()
    smp.init({"microbatches": 1})
    lin = smp.nn.DistributedLinear(input_features=8, output_features=4)
    assert lin(torch.randn(2, 8)).shape == (2, 4)
    ln = smp.nn.DistributedLayerNorm(normalized_shape=(8,), eps=1e-6)
    assert ln(torch.randn(2, 8)).shape == (2, 8) and tuple(ln.normalized_shape) == (8,) and ln.orig_last_dim == 8
    emb = smp.nn.DistributedEmbedding(16, 8, vocab_parallel=False, _skip_scatter_and_merge=True)
    assert emb(torch.randint(0, 16, (2, 3))).shape == (2, 3, 8)
    net = torch.nn.Linear(4, 2)
    Bit16_Optimizer(torch.optim.SGD(net.parameters(), lr=0.1), params_have_main_grad=False, bf16=True)
