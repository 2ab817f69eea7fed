"""No device memory is left behind once the model, optimizer and step outputs are dropped and ``smp.reset()`` has
run (reference test/torch/mpi/test_leak.py:86-116 asserts ``memory_allocated() == 0`` the same way)."""

import gc

import pytest
import torch

pytestmark = pytest.mark.gpu


def _train_and_drop(ckpt):
    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.models.configs import MODEL_CONFIGS

    smp.reset()
    smp.init({"microbatches": 2, "bf16": True})
    cfg = dict(MODEL_CONFIGS["gptj-test"])
    cfg.update(num_layers=2)
    with smp.model_creation(dtype=torch.bfloat16):
        module = smp.nn.DistributedTransformerLMHead(**cfg)
    if ckpt:
        smp.set_activation_checkpointing(module.transformer.seq_layers, pack_args_as_tuple=True, strategy="each")
    model = smp.DistributedModel(module)
    opt = smp.DistributedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-3))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    ids = torch.randint(0, cfg["vocab_size"], (4, 256), device="cuda")
    for _ in range(2):
        opt.zero_grad()
        out = train_step(model, ids)
        opt.step()
    loss = out.reduce_mean().item()
    del out, ids, opt, model, module, train_step
    smp.reset()
    return loss


@pytest.mark.parametrize("ckpt", [False, True])
def test_no_device_memory_left_after_reset(ckpt):
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    gc.collect()
    torch.cuda.synchronize()
    base = torch.cuda.memory_allocated()
    loss = _train_and_drop(ckpt)
    assert loss == loss
    gc.collect()
    torch.cuda.synchronize()
    left = torch.cuda.memory_allocated() - base
    assert left <= 1 << 20, f"{left / 2 ** 20:.1f} MiB still allocated"
