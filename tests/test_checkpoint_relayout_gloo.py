"""A full (gathered) checkpoint written under one parallel layout, loaded under another: tensor-parallel degree 2 -> 4,
tensor parallel -> pipeline parallel, an uneven head split (5 heads over 3 ranks) -> degree 2, and the ``optimize:
memory`` parameter layout -> ``optimize: speed`` (reference torch/checkpoint.py:124-484 ``partial=False`` files are
layout-free; re-slicing on load: torch/utils.py:395-658).  The model that was saved and every model that loaded it must
give the same loss on the same batch."""

import os
import tempfile

import pytest
import torch

from dist_utils import run_distributed

CFG = dict(num_layers=2, num_attention_heads=5, attention_head_size=4, hidden_size=22, intermediate_size=34,
           vocab_size=64, num_positions=16, attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
           embedding_dropout_prob=0.0, activation="gelu", causal_mask_size=16, pre_layernorm=True,
           post_layernorm=False, final_layernorm=True, use_normal_initialization=True)


def _worker(rank, world, conf, ckpt_dir, save):
    import smdistributed_modelparallel_b200 as smp

    conf = dict(conf)
    conf.update({"ddp": True, "microbatches": 2, "ddp_dist_backend": "gloo"})
    smp.init(conf)
    tp = conf.get("tensor_parallel_degree", 1)
    torch.manual_seed(5 if save else 77)                      # a loader starts from other values
    with smp.model_creation(tensor_parallelism=tp > 1):
        module = smp.nn.DistributedTransformerLMHead(**CFG)
    model = smp.DistributedModel(module)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    @smp.step
    def eval_step(model, ids):
        return model((ids, None, None, None, ids))[0]

    ids = torch.randint(0, CFG["vocab_size"], (4, 16), generator=torch.Generator().manual_seed(9))
    if save:
        opt.zero_grad()
        train_step(model, ids)                                # move away from the initial values
        opt.step()
        smp.save_checkpoint(ckpt_dir, tag="full", partial=False, model=model)
    else:
        with torch.no_grad():
            eval_step(model, ids)                             # partitions the model (pipeline layouts need a step)
        smp.resume_from_checkpoint(ckpt_dir, tag="full", partial=False)
    with torch.no_grad():
        loss = eval_step(model, ids).reduce_mean().item()     # every rank sees the same batch
    return {"loss": loss}


LAYOUTS = {
    "tp2": (2, {"tensor_parallel_degree": 2}),
    "tp4": (4, {"tensor_parallel_degree": 4}),
    "tp3": (3, {"tensor_parallel_degree": 3}),
    "pp2": (2, {"pipeline_parallel_degree": 2}),
    "tp2pp2": (4, {"tensor_parallel_degree": 2, "pipeline_parallel_degree": 2}),
    "tp2-memory": (2, {"tensor_parallel_degree": 2, "optimize": "memory"}),
    "dp1": (1, {}),
}


@pytest.mark.parametrize("src,dsts", [("tp2", ["tp4", "pp2", "dp1"]), ("tp3", ["tp2", "tp2pp2"]),
                                      ("tp2-memory", ["tp2", "tp4"]), ("pp2", ["tp2-memory"])])
def test_full_checkpoint_loads_under_other_layouts(src, dsts):
    with tempfile.TemporaryDirectory() as d:
        n, conf = LAYOUTS[src]
        want = run_distributed(_worker, n, conf, d, True, timeout=300)[0]["loss"]
        assert os.path.exists(os.path.join(d, "full"))
        for dst in dsts:
            n2, conf2 = LAYOUTS[dst]
            for r in run_distributed(_worker, n2, conf2, d, False, timeout=300):
                assert abs(r["loss"] - want) < 2e-5, (src, dst, r["loss"], want)
