"""Partial (per-rank) checkpoints in random layouts on four gloo ranks - pipeline x tensor x data parallel, fp32 / bf16 /
fp16 with a dynamic loss scale, optimizer-state sharding, sharded data parallelism, delayed initialisation: train, save,
train on (the reference run), then rebuild everything from other initial values, resume, and train on again - the
continuation has to repeat the reference run step by step (model, optimizer state, fp32 masters and loss scaler all
restored; reference test/torch/mpi_hybrid/test_checkpoint_api.py, test_gpt_checkpoint.py, test_zero_checkpoint.py).
``SMP_RANDOM_RESUME_SEEDS=a-b`` widens the search."""

import os
import random
import tempfile

import pytest
import torch

from dist_utils import run_distributed

CFG = dict(num_layers=4, num_attention_heads=4, attention_head_size=4, hidden_size=16, intermediate_size=32,
           vocab_size=48, num_positions=12, attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
           embedding_dropout_prob=0.0, activation="gelu", causal_mask_size=12, pre_layernorm=True, post_layernorm=False,
           final_layernorm=True, use_normal_initialization=True)


def options(seed):
    rng = random.Random(seed)
    kind = rng.choice(["pp_tp", "pp_tp", "sdp"])
    o = {"kind": kind, "precision": rng.choice(["fp32", "bf16", "fp16"]), "microbatches": rng.choice([2, 4]),
         "delayed": rng.random() < 0.3}
    if kind == "sdp":
        o.update(pp=1, tp=1, sdp=rng.choice([2, 4]), shard=False)
    else:
        pp, tp = rng.choice([(2, 1), (2, 2), (1, 2), (4, 1), (1, 1), (1, 4)])
        o.update(pp=pp, tp=tp, sdp=1, shard=(4 // (pp * tp) > 1 and rng.random() < 0.5))
    return o


def _worker(rank, world, seed, ckpt_dir):
    import smdistributed_modelparallel_b200 as smp

    o = options(seed)
    conf = {"pipeline_parallel_degree": o["pp"], "tensor_parallel_degree": o["tp"], "microbatches": o["microbatches"],
            "ddp": True, "ddp_dist_backend": "gloo", "shard_optimizer_state": o["shard"],
            "delayed_parameter_initialization": o["delayed"], "bf16": o["precision"] == "bf16",
            "fp16": o["precision"] == "fp16"}
    if o["sdp"] > 1:
        conf.update(sharded_data_parallel_degree=o["sdp"], sdp_param_persistence_threshold=100)
        os.environ["SMP_SDP_UNIT_ELEMENTS"] = "3000"
    half = {"bf16": torch.bfloat16, "fp16": torch.float16}.get(o["precision"])

    def build(init_seed):
        smp.init(conf)
        torch.manual_seed(init_seed)
        with smp.model_creation(tensor_parallelism=o["tp"] > 1, dtype=half):
            module = smp.nn.DistributedTransformerLMHead(**CFG)
        model = smp.DistributedModel(module)
        kw = {"dynamic_loss_scale": True, "dynamic_loss_args": {"init_scale": 2.0 ** 10, "scale_window": 2}} \
            if o["precision"] == "fp16" else {}
        opt = smp.DistributedOptimizer(torch.optim.AdamW(model.parameters(), lr=5e-3, weight_decay=0.01), **kw)
        return model, opt

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    def run(model, opt, n, data_seed):
        g = torch.Generator().manual_seed(data_seed + smp.dp_rank())
        out = []
        for _ in range(n):
            ids = torch.randint(0, CFG["vocab_size"], (2 * o["microbatches"], 12), generator=g)
            opt.zero_grad()
            out.append(train_step(model, ids).reduce_mean().item())
            opt.step()
        return out

    model, opt = build(1)
    run(model, opt, 3, 10)
    smp.save_checkpoint(ckpt_dir, tag="r", partial=True, model=model, optimizer=opt, user_content={"seed": seed})
    want = run(model, opt, 3, 20)
    smp.reset()
    model2, opt2 = build(99)
    uc = smp.resume_from_checkpoint(ckpt_dir, tag="r", partial=True)
    got = run(model2, opt2, 3, 20)
    return {"want": want, "got": got, "uc": uc}


def _seeds():
    spec = os.environ.get("SMP_RANDOM_RESUME_SEEDS", "")
    if spec:
        a, b = spec.split("-")
        return list(range(int(a), int(b) + 1))
    return [1, 2, 3, 5, 8, 13]


@pytest.mark.parametrize("seed", _seeds())
def test_resume_repeats_the_uninterrupted_run(seed):
    with tempfile.TemporaryDirectory() as d:
        res = run_distributed(_worker, 4, seed, d, timeout=300)
    for r in res:
        assert r["uc"] == {"seed": seed}
        assert r["got"] == pytest.approx(r["want"], abs=1e-6), (seed, options(seed), r["want"], r["got"])
