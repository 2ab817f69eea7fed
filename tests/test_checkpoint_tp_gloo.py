"""Checkpoints under tensor parallelism on 2 CPU/gloo ranks (reference test tier: mpi_hybrid checkpoint round trips):
partial checkpoint (per-rank shards + optimizer state) restores training exactly; a full checkpoint written under
tp=2 loads into a tp=1 model in a single process; a full checkpoint loads back under tp=2 with re-slicing."""

import os
import tempfile

import torch

from dist_utils import run_distributed

CFG = dict(num_layers=2, num_attention_heads=4, attention_head_size=8, hidden_size=32, intermediate_size=64,
           vocab_size=64, num_positions=16, attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
           embedding_dropout_prob=0.0, activation="gelu_new", causal_mask_size=16, pre_layernorm=True,
           post_layernorm=False, final_layernorm=True, use_normal_initialization=True)


def _worker(rank, world, ckpt_dir):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"tensor_parallel_degree": 2, "ddp": True, "microbatches": 1, "ddp_dist_backend": "gloo"})

    def build(seed):
        torch.manual_seed(seed)
        with smp.tensor_parallelism(enabled=True):
            module = smp.nn.DistributedTransformerLMHead(**CFG)
        model = smp.DistributedModel(module)
        opt = smp.DistributedOptimizer(torch.optim.Adam(model.parameters(), lr=1e-2))
        return model, opt

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    def run(model, opt, n, seed):
        g = torch.Generator().manual_seed(seed + rank)
        out = []
        for _ in range(n):
            ids = torch.randint(0, CFG["vocab_size"], (2, 16), generator=g)
            opt.zero_grad()
            out.append(train_step(model, ids).reduce_mean().item())
            opt.step()
        return out

    model, opt = build(1)
    run(model, opt, 2, 10)
    smp.save_checkpoint(ckpt_dir, tag="p", partial=True, model=model, optimizer=opt, user_content={"step": 2})
    smp.save_checkpoint(ckpt_dir, tag="f", partial=False, model=model)
    want = run(model, opt, 2, 20)                        # the continuation every restored copy must reproduce
    files = sorted(os.listdir(os.path.join(ckpt_dir, "p_partial")))
    # ---- partial: fresh model with different init, same layout
    smp.reset()
    model2, opt2 = build(99)
    uc = smp.resume_from_checkpoint(ckpt_dir, tag="p", partial=True)
    got = run(model2, opt2, 2, 20)
    # ---- full checkpoint back into tp=2 (re-sliced on load); optimizer state is not part of full checkpoints
    smp.reset()
    model3, opt3 = build(7)
    smp.resume_from_checkpoint(ckpt_dir, tag="f", partial=False)
    sd3 = {k: v.clone() for k, v in model3.state_dict().items()}
    return {"want": want, "got": got, "uc": uc, "files": files, "sd3": sd3}


def test_tp2_checkpoints():
    with tempfile.TemporaryDirectory() as d:
        res = run_distributed(_worker, 2, d, timeout=300)
        full = torch.load(os.path.join(d, "f"), weights_only=False)
    for r in res:
        assert r["uc"] == {"step": 2}
        assert all(abs(a - b) < 1e-6 for a, b in zip(r["want"], r["got"])), (r["want"], r["got"])
    assert {"model_0_0.pt", "model_0_1.pt", "optimizer_states_0_0.pt", "optimizer_states_0_1.pt", "smp_config.pt",
            "user_content.pt"} <= set(res[0]["files"])
    # the full file is an ordinary un-sharded state dict: it loads into a tp=1 model built in this process
    import smdistributed_modelparallel_b200  # noqa: F401
    from smdistributed_modelparallel_b200.nn.transformer import DistributedTransformerLMHead

    dense = DistributedTransformerLMHead(**CFG)
    sd = {k: v for k, v in full.items() if not k.startswith("_smp")}
    missing = dense.load_state_dict(sd, strict=False)
    assert not [k for k in missing.missing_keys if "lm_head" not in k], missing
    assert sd["transformer.seq_layers.0.attention.query.weight"].shape == (32, 32)
    # and what tp=2 re-loaded from the full file gathers back to the same tensors
    for k, v in sd.items():
        if k in res[0]["sd3"]:
            assert torch.equal(res[0]["sd3"][k], v), k


def _optshard_worker(rank, world, ckpt_dir):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"ddp": True, "microbatches": 1, "shard_optimizer_state": True, "bf16": True, "ddp_dist_backend": "gloo"})

    def build(seed):
        torch.manual_seed(seed)
        with smp.model_creation(dtype=torch.bfloat16):
            module = smp.nn.DistributedTransformerLMHead(**CFG)
        model = smp.DistributedModel(module)
        opt = smp.DistributedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-2))
        return model, opt

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    def run(model, opt, n, seed):
        g = torch.Generator().manual_seed(seed + rank)
        out = []
        for _ in range(n):
            ids = torch.randint(0, CFG["vocab_size"], (2, 16), generator=g)
            opt.zero_grad()
            out.append(train_step(model, ids).reduce_mean().item())
            opt.step()
        return out

    model, opt = build(1)
    run(model, opt, 2, 10)
    smp.save_checkpoint(ckpt_dir, tag="s", partial=True, model=model, optimizer=opt)
    want = run(model, opt, 2, 20)
    files = sorted(os.listdir(os.path.join(ckpt_dir, "s_partial")))
    smp.reset()
    model2, opt2 = build(99)
    smp.resume_from_checkpoint(ckpt_dir, tag="s", partial=True)
    got = run(model2, opt2, 2, 20)
    return {"want": want, "got": got, "files": files}


def test_optimizer_state_sharding_checkpoint():
    """shard_optimizer_state: every data-parallel rank writes its own optimizer / 16-bit master slice
    (``optimizer_states_{pp}_{tp}_{rdp}.pt``, the v3 naming) and training continues exactly after a resume."""
    with tempfile.TemporaryDirectory() as d:
        res = run_distributed(_optshard_worker, 2, d, timeout=300)
    for r in res:
        assert all(abs(a - b) < 1e-6 for a, b in zip(r["want"], r["got"])), (r["want"], r["got"])
    files = set(res[0]["files"])
    assert {"optimizer_states_0_0_0.pt", "optimizer_states_0_0_1.pt", "model_0_0.pt"} <= files, files
