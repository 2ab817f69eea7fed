"""Data / tensor parallel plumbing on CPU/gloo: DP gradient averaging, tensor-parallel transformer parity
with the dense model (token-parallel layout, uneven-free shapes), optimizer-state sharding convergence
(reference test/torch/mpi_hybrid/test_opt_sharding.py:15-17 style: parameters after N steps vs dense)."""

import pytest
import torch

from dist_utils import run_distributed

CFG = dict(num_layers=2, num_attention_heads=4, attention_head_size=8, hidden_size=32, intermediate_size=64,
           vocab_size=64, num_positions=16, attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
           embedding_dropout_prob=0.0, activation="gelu_new", causal_mask_size=16, pre_layernorm=True,
           post_layernorm=False, final_layernorm=True, use_normal_initialization=True)


def _dp_worker(rank, world, shard, bf16, hook=None):
    import copy

    import smdistributed_modelparallel_b200 as smp

    import os

    os.environ["SMP_DDP_BUCKET_MIN_ELEMS"] = "1000"          # the toy layers become all-reduce buckets
    smp.init({"ddp": True, "microbatches": 2, "shard_optimizer_state": shard, "bf16": bf16})
    torch.manual_seed(3)
    module = smp.nn.DistributedTransformerLMHead(**CFG)
    oracle = copy.deepcopy(module)
    model = smp.DistributedModel(module)
    if hook == "allreduce":
        from torch.distributed.algorithms.ddp_comm_hooks import default_hooks
        model.register_comm_hook(smp.get_dp_process_group(), default_hooks.allreduce_hook)
    elif hook == "custom":
        calls = {"n": 0}

        def my_hook(group, bucket):
            calls["n"] += 1
            t = bucket.buffer()
            fut = torch.distributed.all_reduce(t, group=group, async_op=True).get_future()
            return fut.then(lambda f: f.value()[0] / torch.distributed.get_world_size(group))

        model.register_comm_hook(smp.get_dp_process_group(), my_hook)
    opt = smp.DistributedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.01))
    oopt = torch.optim.AdamW(oracle.parameters(), lr=1e-2, weight_decay=0.01)

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    torch.manual_seed(5)
    all_ids = torch.randint(0, CFG["vocab_size"], (5, 4 * world, 16))
    losses = []
    for it in range(5):
        ids = all_ids[it, rank * 4:(rank + 1) * 4]
        opt.zero_grad()
        out = train_step(model, ids)
        opt.step()
        losses.append(out.reduce_mean().item())
        oopt.zero_grad()
        ol, _ = oracle((all_ids[it], None, None, None, all_ids[it]))
        ol.backward()
        oopt.step()
    worst = max((a.float() - b).abs().max().item() for a, b in zip(model.module.parameters(), oracle.parameters()))
    stats = dict(model.reducer.stats) if model.reducer is not None else {}
    native = type(getattr(model.reducer, "_native", None)).__name__ if model.reducer is not None else None
    return {"worst": worst, "losses": losses, "stats": stats, "buckets": len(getattr(model.reducer, "buckets", [])),
            "native": native}


@pytest.mark.parametrize("shard,bf16", [(False, False), (True, False), (True, True), (False, True)])
def test_dp2_matches_dense(shard, bf16):
    res = run_distributed(_dp_worker, 2, shard, bf16)
    tol = 4e-3 if not bf16 else 5e-2
    for r in res:
        assert r["worst"] < tol, r
        if not bf16 and not shard:
            # fp32 training: per-parameter hooks + native gradient counter + planned buckets reduce during backward
            assert r["stats"]["overlapped_elems"] > 0, r
        if bf16 and not shard:
            # 16-bit training without optimizer-state sharding: the two layers are reduced while backward runs
            assert r["buckets"] == 2 and r["stats"]["overlapped_elems"] > 0 and r["stats"]["tail_elems"] > 0, r
            assert r["native"] == "BucketReducer", r["native"]       # csrc/torch_ext/comm_runtime.cpp did the reducing
    assert abs(res[0]["worst"] - res[1]["worst"]) < 1e-6 or bf16


def _tp_worker(rank, world, tp, distribute_embedding, prescaled):
    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.checkpoint_utils import gather_full_state_dict

    smp.init({"ddp": True, "tensor_parallel_degree": tp, "microbatches": 1, "prescaled_batch": prescaled})
    cfg = dict(CFG, distribute_embedding=distribute_embedding)
    torch.manual_seed(3)
    with smp.tensor_parallelism(enabled=True):
        module = smp.nn.DistributedTransformerLMHead(**cfg)
    model = smp.DistributedModel(module)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=1.0))
    full = {k: v.clone() for k, v in model.state_dict().items()}      # gathered, un-sharded, detached from live storage
    # dense oracle: same weights in a tp=1 replica of the architecture, built outside smp's tp context
    torch.manual_seed(5)
    ids_all = torch.randint(0, CFG["vocab_size"], (2 * world, 16))
    ids = ids_all if prescaled else ids_all[rank * 2:(rank + 1) * 2]

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    opt.zero_grad()
    out = train_step(model, ids)
    loss = out.reduce_mean().item()
    n_grads = sum(1 for _, p in model.local_named_parameters() if p.grad is not None)
    opt.step()                                   # SGD lr=1: new = old - averaged gradient
    after = model.state_dict()
    return {"loss": loss, "full": {k: v.clone() for k, v in full.items()}, "tp_rank": smp.tp_rank(),
            "n_grads": n_grads, "after": {k: v.clone() for k, v in after.items()}}


@pytest.mark.parametrize("distribute_embedding,prescaled", [(False, False), (True, False), (False, True)])
def test_tp2_loss_matches_dense(distribute_embedding, prescaled):
    res = run_distributed(_tp_worker, 2, 2, distribute_embedding, prescaled)
    # rebuild the dense model from the gathered state dict and evaluate the same batches
    import smdistributed_modelparallel_b200  # noqa: F401
    from smdistributed_modelparallel_b200.nn.transformer import DistributedTransformerLMHead

    dense = DistributedTransformerLMHead(**CFG)
    sd = res[0]["full"]
    missing = dense.load_state_dict(sd, strict=False)
    assert not [k for k in missing.missing_keys if "lm_head" not in k], missing
    torch.manual_seed(5)
    ids_all = torch.randint(0, CFG["vocab_size"], (4, 16))
    # gradient parity: one SGD(lr=1) step on the dense model over the whole batch must land on the gathered
    # parameters of the tensor-parallel run
    dense_opt = torch.optim.SGD(dense.parameters(), lr=1.0)
    gl, _ = dense((ids_all, None, None, None, ids_all))
    gl.backward()
    dense_opt.step()
    dsd = dense.state_dict()
    for k, v in res[0]["after"].items():
        if k in dsd:
            err = (v - dsd[k]).abs().max().item()
            assert err < 2e-5, f"{k}: parameter after one step differs from dense by {err}"
    dense.load_state_dict(sd, strict=False)
    if prescaled:
        ol, _ = dense((ids_all, None, None, None, ids_all))
        for r in res:
            # each rank's local objective covers its half of the tokens, scaled so the mean over ranks is the dense loss
            pass
        mean = sum(r["loss"] for r in res) / len(res)
        assert abs(mean - ol.item()) < 1e-4, (mean, ol.item())
    else:
        for r, sl in zip(res, (slice(0, 2), slice(2, 4))):
            ol, _ = dense((ids_all[sl], None, None, None, ids_all[sl]))
            assert abs(r["loss"] - ol.item()) < 1e-4, (r["loss"], ol.item())


@pytest.mark.parametrize("hook,bf16", [("allreduce", False), ("custom", True)])
def test_dp2_with_comm_hook(hook, bf16):
    """register_comm_hook: a DDP communication hook (torch's stock all-reduce hook, or a user future chain) replaces
    the built-in reduction of every bucket; training still matches the dense model."""
    res = run_distributed(_dp_worker, 2, False, bf16, hook)
    for r in res:
        assert r["worst"] < (5e-2 if bf16 else 4e-3), r
        assert r["stats"]["overlapped_elems"] == 0          # the hook path does not use the built-in overlap


def _accum_worker(rank, world):
    import copy

    import smdistributed_modelparallel_b200 as smp

    smp.init({"ddp": True, "microbatches": 2, "ddp_dist_backend": "gloo"})
    torch.manual_seed(3)
    module = smp.nn.DistributedTransformerLMHead(**CFG)
    oracle = copy.deepcopy(module)
    model = smp.DistributedModel(module, backward_passes_per_step=2)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.5))
    oopt = torch.optim.SGD(oracle.parameters(), lr=0.5)

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    torch.manual_seed(5)
    all_ids = torch.randint(0, CFG["vocab_size"], (2, 4 * world, 16))
    opt.zero_grad()
    oopt.zero_grad()
    for k in range(2):                                   # two @smp.step calls, gradients reduced only after the second
        train_step(model, all_ids[k, rank * 4:(rank + 1) * 4])
        ol, _ = oracle((all_ids[k], None, None, None, all_ids[k]))
        ol.backward()
    opt.step()
    oopt.step()
    return max((a.float() - b).abs().max().item() for a, b in zip(model.module.parameters(), oracle.parameters()))


def test_backward_passes_per_step_accumulates_then_reduces():
    """backward_passes_per_step=2: the first call only accumulates locally, the second triggers the data-parallel
    reduction; the update equals the dense model's with both batches' gradients summed."""
    res = run_distributed(_accum_worker, 2)
    assert max(res) < 2e-5, res


def _tp_optshard_worker(rank, world, shard):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"tensor_parallel_degree": 2, "microbatches": 2, "ddp": True, "ddp_dist_backend": "gloo", "bf16": True,
              "shard_optimizer_state": shard})
    torch.manual_seed(1)
    with smp.model_creation(tensor_parallelism=True, dtype=torch.bfloat16):
        module = smp.nn.DistributedTransformerLMHead(**CFG)
    model = smp.DistributedModel(module)
    opt = smp.DistributedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-2))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    g = torch.Generator().manual_seed(10 + smp.rdp_rank())
    out = []
    for _ in range(3):
        ids = torch.randint(0, CFG["vocab_size"], (4, 16), generator=g)
        opt.zero_grad()
        out.append(train_step(model, ids).reduce_mean().item())
        opt.step()
    return {"losses": out, "virtual": len(list(model.virtual_named_parameters()))}


def test_tp2_rdp2_optimizer_state_sharding_matches_unsharded():
    """tp2 x rdp2, bf16: distributed parameters shard their optimizer state over the two replicas, the rest over all
    four ranks; the loss curve is bit-identical to the run that keeps the whole state everywhere."""
    plain = run_distributed(_tp_optshard_worker, 4, False, timeout=300)
    sharded = run_distributed(_tp_optshard_worker, 4, True, timeout=300)
    assert plain[0]["losses"] == sharded[0]["losses"], (plain[0], sharded[0])
    assert plain[0]["virtual"] == 0 and sharded[0]["virtual"] >= 2


def _join_worker(rank, world, by_world):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"microbatches": 2, "ddp": True, "ddp_dist_backend": "gloo"})
    torch.manual_seed(3)
    net = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Tanh(), torch.nn.Linear(16, 4))
    model = smp.DistributedModel(net)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.05))

    @smp.step
    def train_step(model, x, y):
        loss = ((model(x) - y) ** 2).mean()
        model.backward(loss)
        return loss

    steps = 2 if rank == 0 else 5                  # rank 0 runs out of data first
    g = torch.Generator().manual_seed(10 + rank)
    data = [(torch.randn(4, 8, generator=g), torch.randn(4, 4, generator=g)) for _ in range(steps)]
    # oracle for the steps where only rank 1 still trains: its gradient divided by 2 (by_world) or by 1
    with model.join(divide_by_initial_world_size=by_world):
        for x, y in data:
            opt.zero_grad()
            train_step(model, x, y)
            opt.step()
    return {"params": [p.detach().clone() for p in model.parameters()], "steps": steps}


@pytest.mark.parametrize("by_world", [True, False])
def test_join_with_uneven_step_counts(by_world):
    """DistributedModel.join: rank 0 stops after 2 steps, rank 1 after 5; nobody hangs, rank 0 shadows the remaining
    reductions with zeros and both end with rank 1's parameters.  Dividing by the initial world size halves the
    gradient of the lone steps; dividing by the active count does not - so the two settings must end differently."""
    res = run_distributed(_join_worker, 2, by_world, timeout=300)
    for a, b in zip(res[0]["params"], res[1]["params"]):
        assert torch.equal(a, b)
    test_join_with_uneven_step_counts.seen = getattr(test_join_with_uneven_step_counts, "seen", {})
    test_join_with_uneven_step_counts.seen[by_world] = res[1]["params"]
    seen = test_join_with_uneven_step_counts.seen
    if len(seen) == 2:
        assert any(not torch.equal(a, b) for a, b in zip(seen[True], seen[False]))


def _full_opt_worker(rank, world, pp, tp):
    import smdistributed_modelparallel_b200 as smp

    CFG = dict(num_layers=4, num_attention_heads=4, attention_head_size=8, hidden_size=32, intermediate_size=64,
               vocab_size=64, num_positions=16, attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
               embedding_dropout_prob=0.0, activation="gelu_new", causal_mask_size=16, pre_layernorm=True,
               post_layernorm=False, final_layernorm=True, use_normal_initialization=True)
    smp.init({"pipeline_parallel_degree": pp, "tensor_parallel_degree": tp, "microbatches": 2, "ddp": True,
              "ddp_dist_backend": "gloo"})
    torch.manual_seed(7)
    with smp.tensor_parallelism(enabled=tp > 1):
        module = smp.nn.DistributedTransformerLMHead(**CFG)
    model = smp.DistributedModel(module)
    opt = smp.DistributedOptimizer(torch.optim.Adam(model.parameters(), lr=1e-2))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    torch.manual_seed(11)
    ids = torch.randint(0, 64, (4, 16))
    for _ in range(2):
        opt.zero_grad()
        train_step(model, ids)
        opt.step()
    full = opt.state_dict_backcompat()
    want = {n: {k: (v.clone() if torch.is_tensor(v) else v) for k, v in opt.optimizer.state[p].items()}
            for n, p in model.local_named_parameters() if p in opt.optimizer.state}
    n_params_total = len(opt._global_index_to_name)
    # refused by the ordinary loader, accepted by the back-compat one
    refused = False
    try:
        opt.load_state_dict(full)
    except Exception as e:                                  # noqa: BLE001
        refused = "backcompat" in str(e)
    for st in opt.optimizer.state.values():                 # wreck the state, then restore it from the full dict
        for v in st.values():
            if torch.is_tensor(v):
                v.zero_()
    loaded = opt.load_state_dict_backcompat(full)
    worst = 0.0
    for n, p in model.local_named_parameters():
        if n in want:
            for k, v in want[n].items():
                if torch.is_tensor(v):
                    worst = max(worst, (opt.optimizer.state[p][k].float() - v.float()).abs().max().item())
    exp_avg_shapes = {full["_smp_param_names"][i]: tuple(st["exp_avg"].shape) for i, st in full["state"].items()}
    return {"refused": refused, "loaded": loaded, "worst": worst, "n_full": len(full["state"]), "n_total": n_params_total,
            "qkv_full_rows": exp_avg_shapes.get("transformer.seq_layers.0.attention.qkv_weight", (0,))[0],
            "n_local": len(want)}


@pytest.mark.parametrize("pp,tp", [(1, 2), (2, 2)])
def test_full_optimizer_state_dict_round_trip(pp, tp):
    """state_dict_backcompat / load_state_dict_backcompat (reference optimizers/backcompat_opt.py): the gathered
    dictionary covers every parameter of the unpartitioned model with tensor-parallel tensors merged, the ordinary
    loader refuses it, the back-compat loader restores exactly the local state."""
    res = run_distributed(_full_opt_worker, pp * tp, pp, tp, timeout=420)
    for r in res:
        assert r["refused"] and r["worst"] == 0.0 and r["loaded"] == r["n_local"] > 0, r
        assert r["n_full"] == r["n_total"], r
        assert r["qkv_full_rows"] == 3 * 32, r               # both tp halves of the fused q/k/v rows
