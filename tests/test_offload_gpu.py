"""Activation checkpointing with offloading to pinned host memory on one GPU: same loss and gradients as plain
checkpointing, tensors really leave the device, nothing is left behind after the step (reference
test/torch/mpi_4ps/test_offloading.py:37-103)."""

import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(offload):
    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.models.configs import MODEL_CONFIGS

    smp.reset()
    smp.init({"microbatches": 2, "bf16": True, "offload_activations": offload, "activation_loading_horizon": 2})
    cfg = dict(MODEL_CONFIGS["gptj-test"])
    cfg.update(num_layers=4)
    torch.manual_seed(1)
    with smp.model_creation(dtype=torch.bfloat16):
        module = smp.nn.DistributedTransformerLMHead(**cfg)
    smp.set_activation_checkpointing(module.transformer.seq_layers, pack_args_as_tuple=True, strategy="each")
    model = smp.DistributedModel(module)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.0))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    torch.manual_seed(2)
    ids = torch.randint(0, cfg["vocab_size"], (4, 512), device="cuda")
    opt.zero_grad()
    loss = train_step(model, ids).reduce_mean().item()
    grads = {n: p.main_grad.clone() if hasattr(p, "main_grad") else p.grad.clone()
             for n, p in model.local_named_parameters()}
    stats = None
    if offload:
        offs = list(smp.state.offloaders.values())
        stats = {"bytes": sum(o.bytes_offloaded for o in offs), "left": sum(o.saved_tensors() for o in offs)}
    torch.cuda.synchronize()
    return loss, grads, stats


def test_offload_matches_plain_checkpointing():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    l0, g0, _ = _run(False)
    l1, g1, st = _run(True)
    assert abs(l0 - l1) < 1e-6, (l0, l1)
    for n in g0:
        assert torch.equal(g0[n], g1[n]), n
    # 4 layers x 2 microbatches x [2, 512, 768] bf16 inputs were copied out, and all were consumed
    assert st["bytes"] >= 4 * 2 * 2 * 512 * 768 * 2, st
    assert st["left"] == 0, st
