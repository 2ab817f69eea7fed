"""BASELINE config 5 in miniature on CPU/gloo, 4 processes: BERT-style (post-LayerNorm, bidirectional) model,
tensor_parallel_degree=2 x data parallel 2, optimizer-state sharding, fp16 with dynamic loss scaling.  Training
must run (no overflow skips after the scale settles), the loss must fall, and every tensor-parallel replica pair
must end with identical parameters (reference test tier: test/torch/mpi_hybrid/test_opt_sharding.py)."""

import torch

from dist_utils import run_distributed

CFG = dict(num_layers=2, num_attention_heads=4, attention_head_size=8, hidden_size=32, intermediate_size=64,
           vocab_size=64, num_positions=16, attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
           embedding_dropout_prob=0.0, activation="gelu", layernorm_epsilon=1e-12, causal_mask_size=None,
           pre_layernorm=False, post_layernorm=True, num_token_types=2, use_normal_initialization=True,
           initializer_range=0.02, add_lm_head=True)


def _worker(rank, world):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"tensor_parallel_degree": 2, "ddp": True, "microbatches": 2, "shard_optimizer_state": True, "fp16": True,
              "ddp_dist_backend": "gloo"})
    torch.manual_seed(3)
    with smp.model_creation(tensor_parallelism=True, dtype=torch.float16):
        module = smp.nn.DistributedTransformerLMHead(**CFG)
    model = smp.DistributedModel(module)
    opt = smp.DistributedOptimizer(torch.optim.AdamW(model.parameters(), lr=5e-3), dynamic_loss_scale=True,
                                   dynamic_loss_args={"init_scale": 2.0 ** 10, "scale_window": 4})

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    torch.manual_seed(100 + rank)
    ids = torch.randint(0, CFG["vocab_size"], (4, 16))
    losses, scales = [], []
    for _ in range(12):
        opt.zero_grad()
        out = train_step(model, ids)
        opt.step()
        losses.append(out.reduce_mean().item())
        scales.append(float(opt.loss_scale))
    sd = {k: v.float().clone() for k, v in model.state_dict().items()}      # gathered over tp
    return {"losses": losses, "scales": scales, "sd": sd, "tp_rank": smp.tp_rank(), "rdp_rank": smp.rdp_rank()}


def test_bert_tp2_dp2_optshard_fp16():
    res = run_distributed(_worker, 4, timeout=420)
    for r in res:
        assert all(l == l for l in r["losses"]), r["losses"]
        assert r["losses"][-1] < r["losses"][0] - 0.3, r["losses"]
        assert r["scales"][-1] >= 2.0 ** 10, r["scales"]            # the scale grew or held: no persistent overflow
    # data-parallel replicas stay in lock-step: the gathered state dict is the same on every rank
    ref = res[0]["sd"]
    for r in res[1:]:
        for k, v in ref.items():
            assert torch.equal(v, r["sd"][k]), k
