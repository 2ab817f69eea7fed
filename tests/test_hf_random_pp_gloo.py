"""Unmodified HuggingFace causal LMs (GPT-2, GPT-J, GPT-NeoX; random depth / width) split by the auto-partitioner over
two or three pipeline stages, with and without activation checkpointing of the blocks and a data-parallel replica:
loss and every local gradient equal the single-process model's (keyword arguments, ``ModelOutput`` results, tied
embeddings, rotary caches and masks created inside the modules all cross stage boundaries).  Reference tier:
test/torch/mpi_hybrid/test_gpt2.py.  ``SMP_RANDOM_HFPP_SEEDS=a-b`` widens the search."""

import copy
import os
import random

import pytest
import torch

transformers = pytest.importorskip("transformers")

from dist_utils import run_distributed
from test_hf_random_configs_gloo import make


def options(seed):
    rng = random.Random(seed)
    heads = rng.randint(2, 4)
    world, pp = rng.choice([(2, 2), (3, 3), (4, 2)])
    return {"family": rng.choice(["gpt2", "gptj", "gptneox"]), "heads": heads, "hidden": heads * 8,
            "layers": rng.randint(pp, 5), "rotary": 4, "window": 8, "vocab": rng.choice([50, 64]), "seq": 12,
            "world": world, "pp": pp, "ckpt": rng.random() < 0.4, "microbatches": rng.choice([2, 4]),
            "memory_weight": rng.choice([0.0, 0.8])}


def _worker(rank, world, seed):
    import smdistributed_modelparallel_b200 as smp

    o = options(seed)
    conf = {"pipeline_parallel_degree": o["pp"], "microbatches": o["microbatches"], "memory_weight": o["memory_weight"],
            "ddp": world > o["pp"]}
    if conf["ddp"]:
        conf["ddp_dist_backend"] = "gloo"
    smp.init(conf)
    cfg, Cls = make(o)
    torch.manual_seed(seed)
    net = Cls(cfg)
    oracle = copy.deepcopy(net)
    if o["ckpt"]:
        blocks = net.transformer.h if hasattr(net, "transformer") else net.gpt_neox.layers
        smp.set_activation_checkpointing(blocks, strategy="each")
    model = smp.DistributedModel(net)

    @smp.step
    def step(model, ids):
        out = model(input_ids=ids, labels=ids)
        model.backward(out.loss)
        return out.loss

    dp = world // o["pp"]
    per = 2 * o["microbatches"]
    all_ids = torch.randint(0, o["vocab"], (dp * per, o["seq"]), generator=torch.Generator().manual_seed(seed + 5))
    ids = all_ids[smp.dp_rank() * per:(smp.dp_rank() + 1) * per]
    loss = step(model, ids).reduce_mean().item()
    mbs = o["microbatches"]
    n_mb = dp * mbs
    ol = sum(oracle(input_ids=all_ids[i:i + 2], labels=all_ids[i:i + 2]).loss / n_mb for i in range(0, dp * per, 2))
    ol.backward()
    own = sum(oracle(input_ids=ids[i:i + 2], labels=ids[i:i + 2]).loss.item() / mbs for i in range(0, per, 2))
    ref = dict(oracle.named_parameters())
    worst = max((p.grad - ref[n].grad).abs().max().item() for n, p in model.local_named_parameters())
    return {"loss": loss, "own": own, "worst": worst,
            "stages": sorted(set(smp.state.module_manager.partition_info().values()))}


def _seeds():
    spec = os.environ.get("SMP_RANDOM_HFPP_SEEDS", "")
    if spec:
        a, b = spec.split("-")
        return list(range(int(a), int(b) + 1))
    return [1, 2, 4, 5, 8, 11]


@pytest.mark.parametrize("seed", _seeds())
def test_random_stock_huggingface_models_under_pipeline_parallelism(seed):
    o = options(seed)
    for r in run_distributed(_worker, o["world"], seed, timeout=300):
        assert r["stages"] == list(range(o["pp"])), (seed, o, r["stages"])
        assert abs(r["loss"] - r["own"]) < 5e-5 and r["worst"] < 5e-5, (seed, o, r)
