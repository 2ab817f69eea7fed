"""Random combinations of engine options on four gloo ranks - pipeline degree 2 / 4, tensor-parallel degree 1 / 2,
microbatch count and window, ``simple`` / ``interleaved`` schedule, fast mode, static mode, activation checkpointing
strategy, prescaled batch, schedule tuning, optimizer-state sharding, reducer overlap - each trained for three SGD steps and compared, loss by loss, with the
dense model on the union batch.  Options are not independent in the engine (fast-mode chains change who talks to whom,
static mode replays receptions, checkpointing re-runs forwards inside backward ...), so they are exercised together.
``SMP_RANDOM_ENGINE_SEEDS=a-b`` widens the search."""

import os
import random

import pytest
import torch

from dist_utils import run_distributed

CFG = dict(num_layers=4, num_attention_heads=4, attention_head_size=4, hidden_size=16, intermediate_size=32,
           vocab_size=48, num_positions=12, attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
           embedding_dropout_prob=0.0, activation="gelu", causal_mask_size=12, pre_layernorm=True, post_layernorm=False,
           final_layernorm=True, use_normal_initialization=True)


def options(seed):
    rng = random.Random(seed)
    pp, tp = rng.choice([(2, 1), (2, 2), (4, 1), (1, 2), (1, 4), (2, 2)])
    mb = rng.choice([2, 4])
    o = {"pipeline_parallel_degree": pp, "tensor_parallel_degree": tp, "microbatches": mb,
         "pipeline": rng.choice(["simple", "interleaved"]), "fast_mode": rng.random() < 0.4,
         "static_mode": rng.random() < 0.4, "prescaled_batch": tp > 1 and rng.random() < 0.3}
    if o["pipeline"] == "interleaved":
        o["active_microbatches"] = rng.choice([mb, max(1, mb // 2), min(mb, pp + 1)])
    extra = {"ckpt": rng.choice([None, None, "each", "contiguous", "group_2"]),
             "record_steps": rng.choice([0, 0, 2]), "delayed": rng.random() < 0.3,
             "overlap": rng.random() < 0.7}
    if 4 // (pp * tp) > 1 and rng.random() < 0.5:
        o["shard_optimizer_state"] = True
    if extra["delayed"]:
        o["delayed_parameter_initialization"] = True
    return o, extra


def _worker(rank, world, seed):
    import smdistributed_modelparallel_b200 as smp

    o, extra = options(seed)
    os.environ["SMP_SCHEDULE_RECORD_STEPS"] = str(extra["record_steps"])
    conf = dict(o)
    conf.update({"ddp": True, "ddp_dist_backend": "gloo"})
    smp.init(conf)
    tp, pp = o["tensor_parallel_degree"], o["pipeline_parallel_degree"]
    torch.manual_seed(31)
    with smp.model_creation(tensor_parallelism=tp > 1):
        module = smp.nn.DistributedTransformerLMHead(**CFG)
    if extra["ckpt"]:
        smp.set_activation_checkpointing(module.transformer.seq_layers, pack_args_as_tuple=True, strategy=extra["ckpt"])
    model = smp.DistributedModel(module, overlapping_allreduce=extra["overlap"])
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.2))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    @smp.step
    def eval_step(model, ids):
        return model((ids, None, None, None, ids))[0]

    dp = world // (tp * pp)
    groups = dp * (1 if o["prescaled_batch"] else tp)
    per = 4 * o["microbatches"]                     # four sequences per microbatch: divisible by every tp degree used
    all_ids = torch.randint(0, CFG["vocab_size"], (3, groups * per, 12), generator=torch.Generator().manual_seed(seed))
    mine = smp.rdp_rank() if o["prescaled_batch"] else smp.dp_rank()
    with torch.no_grad():
        eval_step(model, all_ids[0, mine * per:(mine + 1) * per])      # partitions / materialises the model
    start = {k: v.clone() for k, v in model.state_dict().items()}      # gathered over tp and pp: the dense twin's start
    losses = []
    for it in range(3):
        opt.zero_grad()
        out = train_step(model, all_ids[it, mine * per:(mine + 1) * per])
        opt.step()
        vals = smp.state.comm.allgather(out.reduce_mean().item(), smp.CommGroup.WORLD)
        losses.append(sum(vals) / len(vals))
    return {"losses": losses, "all_ids": all_ids, "start": start}


def _seeds():
    spec = os.environ.get("SMP_RANDOM_ENGINE_SEEDS", "")
    if spec:
        a, b = spec.split("-")
        return list(range(int(a), int(b) + 1))
    return [2, 7, 9, 14, 20, 33]


@pytest.mark.parametrize("seed", _seeds())
def test_random_engine_option_combinations_track_dense(seed):
    import smdistributed_modelparallel_b200  # noqa: F401
    from smdistributed_modelparallel_b200.nn.transformer import DistributedTransformerLMHead

    res = run_distributed(_worker, 4, seed, timeout=300)
    dense = DistributedTransformerLMHead(**CFG)
    missing = dense.load_state_dict(res[0]["start"], strict=False)
    assert not [k for k in missing.missing_keys if "lm_head" not in k], missing
    opt = torch.optim.SGD(dense.parameters(), lr=0.2)
    want = []
    for it in range(3):
        ids = res[0]["all_ids"][it].long()             # tensors come back from the workers as float arrays
        opt.zero_grad()
        loss, _ = dense((ids, None, None, None, ids))
        loss.backward()
        opt.step()
        want.append(loss.item())
    for r in res:
        assert r["losses"] == pytest.approx(want, abs=1e-4), (seed, options(seed), r["losses"], want)
