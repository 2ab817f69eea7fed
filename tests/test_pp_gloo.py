"""Pipeline parallelism on CPU/gloo, world size 2: GPT-2-style model, pp=2, 4 microbatches.  Loss and every
parameter gradient must match the dense single-process model (the reference's oracle design,
reference test/torch/smp_test_base.py:603-639; BASELINE config #1)."""

import pytest
import torch

from dist_utils import run_distributed

CFG = dict(num_layers=4, num_attention_heads=4, attention_head_size=8, hidden_size=32, intermediate_size=64,
           vocab_size=64, num_positions=16, attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
           embedding_dropout_prob=0.0, activation="gelu_new", causal_mask_size=16, pre_layernorm=True,
           post_layernorm=False, final_layernorm=True, use_normal_initialization=True)


def _worker(rank, world, pipeline, microbatches, trace=False):
    import copy

    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "microbatches": microbatches, "pipeline": pipeline, "ddp": False})
    torch.manual_seed(7)
    module = smp.nn.DistributedTransformerLMHead(**CFG)
    oracle = copy.deepcopy(module)
    model = smp.DistributedModel(module, trace_execution_times=trace, trace_memory_usage=trace)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    torch.manual_seed(11)
    ids = torch.randint(0, CFG["vocab_size"], (8, 16))
    opt.zero_grad()
    out = train_step(model, ids)
    loss = out.reduce_mean().item()
    # oracle: same batch in one piece
    ol, _ = oracle((ids, None, None, None, ids))
    ol.backward()
    ref = {n: p.grad for n, p in oracle.named_parameters()}
    worst = 0.0
    n_local = 0
    for n, p in model.local_named_parameters():
        assert p.grad is not None, f"{n} has no grad on pp_rank {smp.pp_rank()}"
        worst = max(worst, (p.grad - ref[n]).abs().max().item())
        n_local += 1
    part = smp.state.module_manager.partition_info()
    opt.step()
    tr = smp.state.module_manager.trace
    pool = smp.state.core.get_and_reset_memory_stats()
    allocs = smp.state.core.get_and_reset_alloc_stats()
    eng = smp.state.exec_server.p2p._engine
    native = dict(eng.stats()) if eng is not None else None
    return {"native_p2p": native, "pool_mb": pool["d2d_peak_allocated_mb"], "pool_reserved_mb": pool["d2d_peak_reserved_mb"],
            "pool_allocs": allocs["success"], "pool_failures": allocs["failure"], "loss": loss, "oracle": ol.item(), "worst_grad_err": worst, "n_local": n_local,
            "ranks_used": sorted(set(part.values())), "hops": smp.state.module_manager.hops,
            "traced": len(tr.order), "trace_time": sum(tr.times.values()), "trace_bytes": sum(tr.memory.values())}


@pytest.mark.parametrize("pipeline", ["interleaved", "simple"])
def test_pp2_matches_dense(pipeline):
    res = run_distributed(_worker, 2, pipeline, 4)
    for r in res:
        assert abs(r["loss"] - r["oracle"]) < 5e-5, r      # fp32 noise: one batch vs microbatches (other GEMM shapes)
        assert r["worst_grad_err"] < 3e-5, r
        assert r["ranks_used"] == [0, 1]
        assert r["n_local"] > 0
        # every stage received tensors from the other one: the transport's receive buffers show up in the native
        # core's pool metrics (4 microbatches x [2, 16, hidden] fp32 activations at least)
        assert r["pool_allocs"] >= 4 and r["pool_failures"] == 0, r
        assert r["pool_mb"] * 2 ** 20 >= 4 * 2 * 16 * CFG["hidden_size"] * 4, r
        assert r["pool_reserved_mb"] >= r["pool_mb"]
        # ... and were moved by the native runtime's handle engine (csrc/torch_ext/comm_runtime.cpp)
        assert r["native_p2p"]["sent"] >= 4 and r["native_p2p"]["received"] >= 4 and r["native_p2p"]["pending"] == 0, r
    assert sum(r["n_local"] for r in res) > 0


def test_pp2_with_first_step_tracing():
    """trace_execution_times / trace_memory_usage: the leader runs one traced microbatch before partitioning and
    feeds times and activation sizes to the partitioner; training is unaffected."""
    res = run_distributed(_worker, 2, "interleaved", 4, True)
    for r in res:
        assert abs(r["loss"] - r["oracle"]) < 5e-5, r      # fp32 noise: one batch vs microbatches (other GEMM shapes)
        assert r["worst_grad_err"] < 3e-5, r
        assert r["ranks_used"] == [0, 1]
    leader = res[0]
    assert leader["traced"] > 10 and leader["trace_time"] > 0 and leader["trace_bytes"] > 0, leader
    assert res[1]["traced"] == 0


def _pp_tp_worker(rank, world, window, prescaled=False):
    """pp=2 x tp=2 on four processes.  Exercises the deterministic leader schedule: all tp ranks of a stage must
    issue their collectives in the same order."""
    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "tensor_parallel_degree": 2, "microbatches": 4, "ddp": True,
              "active_microbatches": window, "ddp_dist_backend": "gloo", "prescaled_batch": prescaled})
    torch.manual_seed(7)
    with smp.tensor_parallelism(enabled=True):
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

    torch.manual_seed(11)
    all_ids = torch.randint(0, CFG["vocab_size"], (3, 16, 16))
    with torch.no_grad():
        eval_step(model, all_ids[0, :8])          # partitions the model
    full = {k: v.clone() for k, v in model.state_dict().items()}
    losses = []
    for it in range(3):
        # prescaled_batch: every tp rank is handed the whole group batch; otherwise each brings its own half
        ids = all_ids[it] if prescaled else all_ids[it, smp.tp_rank() * 8:(smp.tp_rank() + 1) * 8]
        opt.zero_grad()
        out = train_step(model, ids)
        opt.step()
        losses.append(out.reduce_mean().item())
    return {"losses": losses, "full": full, "tp_rank": smp.tp_rank(), "pp_rank": smp.pp_rank(), "rank": rank}


@pytest.mark.parametrize("window,prescaled", [(2, False), (4, True)])
def test_pp2_tp2_tracks_dense(window, prescaled):
    """Several optimizer steps must track the dense model (built here, tp=1) trained on the union of the
    tensor-parallel ranks' batches - every tp rank brings its own sequences."""
    res = run_distributed(_pp_tp_worker, 4, window, prescaled, timeout=420)
    import smdistributed_modelparallel_b200  # noqa: F401
    from smdistributed_modelparallel_b200.nn.transformer import DistributedTransformerLMHead

    dense = DistributedTransformerLMHead(**CFG)
    r0 = next(r for r in res if r["rank"] == 0)
    missing = dense.load_state_dict(r0["full"], strict=False)
    assert not [k for k in missing.missing_keys if "lm_head" not in k], missing
    oopt = torch.optim.SGD(dense.parameters(), lr=0.1)
    torch.manual_seed(11)
    all_ids = torch.randint(0, CFG["vocab_size"], (3, 16, 16))
    for it in range(3):
        oopt.zero_grad()
        ol, _ = dense((all_ids[it], None, None, None, all_ids[it]))
        ol.backward()
        oopt.step()
        mean = sum(r["losses"][it] for r in res if r["pp_rank"] == 0) / 2
        assert abs(mean - ol.item()) < 3e-5, (it, mean, ol.item())


def _pp4_worker(rank, world, fast_mode=False, static_mode=False):
    import copy

    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 4, "microbatches": 4, "ddp": False, "active_microbatches": 3,
              "fast_mode": fast_mode, "static_mode": static_mode})
    torch.manual_seed(7)
    cfg = dict(CFG, num_layers=8)
    module = smp.nn.DistributedTransformerLMHead(**cfg)
    oracle = copy.deepcopy(module)
    model = smp.DistributedModel(module)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    torch.manual_seed(11)
    ids = torch.randint(0, cfg["vocab_size"], (8, 16))
    opt.zero_grad()
    loss = train_step(model, ids).reduce_mean().item()
    ol, _ = oracle((ids, None, None, None, ids))
    ol.backward()
    ref = {n: p.grad for n, p in oracle.named_parameters()}
    worst, n_local = 0.0, 0
    for n, p in model.local_named_parameters():
        worst = max(worst, (p.grad - ref[n]).abs().max().item())
        n_local += 1
    layers = [smp.state.module_manager.get_partition(l) for l in model.module.transformer.seq_layers]
    opt.step()
    loss2 = train_step(model, ids).reduce_mean().item()
    return {"loss": loss, "oracle": ol.item(), "worst": worst, "n_local": n_local, "layers": layers,
            "loss2": loss2, "hop_bytes": smp.state.module_manager.comm_volume_bytes,
            "static": dict(smp.state.exec_server.p2p.stats)}


@pytest.mark.parametrize("fast_mode,static_mode", [(False, False), (True, False), (False, True)])
def test_pp4_matches_dense(fast_mode, static_mode):
    """Four pipeline stages (reference test/torch/mpi_4ps): the layer stack is cut into four contiguous chunks
    served in sequence; loss and every local gradient match the dense model.  In fast mode the chunks hand
    their outputs to one another directly (and the gradients walk back the same way), so the leader moves
    far fewer bytes."""
    res = run_distributed(_pp4_worker, 4, fast_mode, static_mode, timeout=420)
    for r in res:
        assert abs(r["loss"] - r["oracle"]) < 5e-5, r
        assert r["worst"] < 3e-5, r
        assert r["n_local"] > 0
        assert r["loss2"] < r["loss"]
    layers = res[0]["layers"]
    assert sorted(set(layers)) == [0, 1, 2, 3] and layers == sorted(layers), layers
    if static_mode:                      # the second step replayed the record of the first on every stage
        assert all(r["static"]["replayed_steps"] == 1 and r["static"]["preposted"] > 0 for r in res), res
        return
    test_pp4_matches_dense.leader_bytes[fast_mode] = res[0]["hop_bytes"]
    if len(test_pp4_matches_dense.leader_bytes) == 2:
        assert test_pp4_matches_dense.leader_bytes[True] < 0.6 * test_pp4_matches_dense.leader_bytes[False]


test_pp4_matches_dense.leader_bytes = {}


def _pp_dp_worker(rank, world):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 2, "ddp": True, "bf16": True, "ddp_dist_backend": "gloo"})
    torch.manual_seed(7)
    with smp.model_creation(dtype=torch.bfloat16):
        module = smp.nn.DistributedTransformerLMHead(**CFG)
    smp.set_activation_checkpointing(module.transformer.seq_layers, pack_args_as_tuple=True, strategy="group_2")
    model = smp.DistributedModel(module)
    opt = smp.DistributedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-2))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    @smp.step
    def eval_step(model, ids):
        return model((ids, None, None, None, ids))[0]

    torch.manual_seed(11)
    all_ids = torch.randint(0, CFG["vocab_size"], (4, 8, 16))
    with torch.no_grad():
        eval_step(model, all_ids[0, :4])
    full = {k: v.float().clone() for k, v in model.state_dict().items()}
    losses = []
    for it in range(4):
        ids = all_ids[it, smp.dp_rank() * 4:(smp.dp_rank() + 1) * 4]
        opt.zero_grad()
        losses.append(train_step(model, ids).reduce_mean().item())
        opt.step()
    after = {k: v.float().clone() for k, v in model.state_dict().items()}
    return {"losses": losses, "full": full, "after": after, "pp_rank": smp.pp_rank(), "dp_rank": smp.dp_rank(), "rank": rank}


def test_pp2_dp2_bf16_with_activation_checkpointing():
    """Two pipeline stages x two data-parallel replicas, bf16 with fp32 masters, group_2 activation checkpointing:
    parameters after four AdamW steps track the dense fp32 model trained on both replicas' batches."""
    res = run_distributed(_pp_dp_worker, 4, timeout=420)
    import smdistributed_modelparallel_b200  # noqa: F401
    from smdistributed_modelparallel_b200.nn.transformer import DistributedTransformerLMHead

    dense = DistributedTransformerLMHead(**CFG)
    r0 = next(r for r in res if r["rank"] == 0)
    dense.load_state_dict(r0["full"], strict=False)
    oopt = torch.optim.AdamW(dense.parameters(), lr=1e-2)
    torch.manual_seed(11)
    all_ids = torch.randint(0, CFG["vocab_size"], (4, 8, 16))
    for it in range(4):
        oopt.zero_grad()
        ol, _ = dense((all_ids[it], None, None, None, all_ids[it]))
        ol.backward()
        oopt.step()
        leaders = [r for r in res if r["pp_rank"] == 0]
        mean = sum(r["losses"][it] for r in leaders) / len(leaders)
        assert abs(mean - ol.item()) < 5e-2, (it, mean, ol.item())
    dsd = dense.state_dict()
    worst = max((r0["after"][k] - dsd[k]).abs().max().item() for k in dsd if k in r0["after"])
    assert worst < 5e-2, worst
    # replicas hold identical parameters
    other = next(r for r in res if r["pp_rank"] == 0 and r["dp_rank"] == 1)
    for k, v in r0["after"].items():
        assert torch.equal(v, other["after"][k]), k


def _delayed_worker(rank, world, trace=False, tp=1, trace_device="gpu", break_trace=False):
    import smdistributed_modelparallel_b200 as smp

    conf = {"pipeline_parallel_degree": 2, "microbatches": 2, "ddp": tp > 1, "delayed_parameter_initialization": True,
            "memory_weight": 0.0 if trace else 0.8}
    if tp > 1:
        conf.update(tensor_parallel_degree=tp, ddp_dist_backend="gloo")
    smp.init(conf)
    with smp.model_creation(tensor_parallelism=tp > 1):
        module = smp.nn.DistributedTransformerLMHead(**CFG)
    assert all(p.is_meta for p in module.parameters()), "nothing may be materialised before partitioning"
    model = smp.DistributedModel(module, trace_execution_times=trace, trace_memory_usage=trace,
                                 trace_device=trace_device)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))
    if break_trace:
        def boom(mod, args):
            if smp.is_tracing():
                raise RuntimeError("this module cannot run in the traced pass")
        module.transformer.seq_layers[0].register_forward_pre_hook(boom)

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    torch.manual_seed(11)
    ids = torch.randint(0, CFG["vocab_size"], (8, 16))
    losses = []
    for _ in range(3):
        opt.zero_grad()
        losses.append(train_step(model, ids).reduce_mean().item())
        opt.step()
    local = list(model.local_parameters())
    everything = list(model.module.parameters())
    tr = smp.state.module_manager.trace
    return {"losses": losses, "local_elems": sum(p.numel() for p in local),
            "held_elems": sum(p.numel() for p in everything if not p.is_meta),
            "meta_left": sum(1 for p in local if p.is_meta), "traced": len(tr.order),
            "trace_time": sum(tr.times.values()), "pp_rank": smp.pp_rank(),
            "partition": dict(smp.state.module_manager.partition_info())}


def test_delayed_parameter_initialisation_under_pp():
    """delayed_parameter_initialization: the model is built on the meta device; after partitioning each pipeline rank
    materialises (and initialises) only the parameters it owns, and training runs."""
    res = run_distributed(_delayed_worker, 2, timeout=300)
    total = 0
    for r in res:
        assert r["meta_left"] == 0
        assert r["held_elems"] == r["local_elems"], r          # nothing beyond the local partition holds storage
        assert all(l == l for l in r["losses"]) and r["losses"][-1] < r["losses"][0]
        assert r["losses"][0] < 4.6, r["losses"]      # ~ln(64): the constructor's initialisers were replayed
        total += r["local_elems"]
    assert res[0]["local_elems"] > 0 and res[1]["local_elems"] > 0 and total > 0


@pytest.mark.parametrize("tp", [1, 2])
def test_tracing_under_delayed_initialisation(tp):
    """With delayed initialisation the parameters have no storage when the first step arrives: the leaders borrow
    storage for the traced passes, hand it back, and every rank - whatever times its own leader measured - uses rank
    0's partition.  Afterwards only the local partition is materialised, as without tracing."""
    res = run_distributed(_delayed_worker, 2 * tp, True, tp, timeout=420)
    for r in res:
        assert r["meta_left"] == 0 and r["held_elems"] == r["local_elems"], r
        assert all(l == l for l in r["losses"]) and r["losses"][-1] < r["losses"][0]
        assert r["partition"] == res[0]["partition"]
        assert (r["traced"] > 10 and r["trace_time"] > 0) if r["pp_rank"] == 0 else r["traced"] == 0, r
    assert set(res[0]["partition"].values()) == {0, 1}


def test_tracing_on_the_host_and_a_trace_that_fails():
    """``trace_device="cpu"`` runs the traced passes on borrowed host storage; a traced pass that raises does not end
    the job - the model is partitioned from the untraced estimates and training goes on."""
    host = run_distributed(_delayed_worker, 2, True, 1, "cpu", timeout=420)
    broken = run_distributed(_delayed_worker, 2, True, 1, "gpu", True, timeout=420)
    for r in host:
        assert r["meta_left"] == 0 and r["losses"][-1] < r["losses"][0]
        assert (r["traced"] > 10 and r["trace_time"] > 0) if r["pp_rank"] == 0 else r["traced"] == 0, r
    for r in broken:
        assert r["meta_left"] == 0 and r["losses"][-1] < r["losses"][0]
        assert r["trace_time"] == 0                      # nothing of the failed trace is used
        assert r["partition"] == broken[0]["partition"] and set(r["partition"].values()) == {0, 1}


def _pp_optshard_worker(rank, world):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 2, "ddp": True, "fp16": True, "shard_optimizer_state": True,
              "ddp_dist_backend": "gloo"})
    torch.manual_seed(7)
    with smp.model_creation(dtype=torch.float16):
        module = smp.nn.DistributedTransformerLMHead(**CFG)
    model = smp.DistributedModel(module)
    opt = smp.DistributedOptimizer(torch.optim.AdamW(model.parameters(), lr=5e-3), dynamic_loss_scale=True,
                                   dynamic_loss_args={"init_scale": 2.0 ** 12, "scale_window": 3})

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    torch.manual_seed(100 + smp.dp_rank())
    ids = torch.randint(0, CFG["vocab_size"], (4, 16))
    losses, scales = [], []
    for _ in range(10):
        opt.zero_grad()
        out = train_step(model, ids)
        opt.step()
        losses.append(out.reduce_mean().item())
        scales.append(float(opt.loss_scale))
    sd = {k: v.float().clone() for k, v in model.state_dict().items()}
    return {"losses": losses, "scales": scales, "sd": sd, "pp_rank": smp.pp_rank(), "dp_rank": smp.dp_rank()}


def test_pp2_dp2_fp16_dynamic_scale_with_optimizer_state_sharding():
    """Pipeline x data parallel with fp16 dynamic loss scaling and optimizer-state sharding: overflow decisions are
    shared by every rank (same scale history), the loss falls, replicas stay identical."""
    res = run_distributed(_pp_optshard_worker, 4, timeout=420)
    for r in res:
        assert r["scales"] == res[0]["scales"], (r["scales"], res[0]["scales"])
        if r["pp_rank"] == 0:
            assert all(l == l for l in r["losses"]) and r["losses"][-1] < r["losses"][0] - 0.2, r["losses"]
    for r in res[1:]:
        for k, v in res[0]["sd"].items():
            assert torch.equal(v, r["sd"][k]), k


def _failing_step_worker(rank, world):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 2, "ddp": False})
    torch.manual_seed(7)
    model = smp.DistributedModel(smp.nn.DistributedTransformerLMHead(**CFG))

    @smp.step
    def bad_step(model, ids, fail):
        loss, _ = model((ids, None, None, None, ids))
        if fail:
            raise ValueError("deliberate failure in the step function")
        model.backward(loss)
        return loss

    ids = torch.randint(0, CFG["vocab_size"], (4, 16), generator=torch.Generator().manual_seed(3))
    ok = bad_step(model, ids, False).reduce_mean().item()
    try:
        bad_step(model, ids, True)
    except Exception as e:   # noqa: BLE001
        return {"ok": ok, "error": f"{type(e).__name__}: {e}"}
    return {"ok": ok, "error": None}


def test_step_failure_reaches_every_pipeline_rank():
    """An exception raised by the step function on the leader ends the step on the other stage too (it is told
    why), instead of leaving it waiting for messages that will never come."""
    res = run_distributed(_failing_step_worker, 2, timeout=180)
    assert "deliberate failure" in res[0]["error"] and res[0]["error"].startswith("ValueError")
    assert "deliberate failure" in res[1]["error"], res[1]


def _static_worker(rank, world, static_mode):
    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 4, "ddp": False, "static_mode": static_mode})
    torch.manual_seed(7)
    model = smp.DistributedModel(smp.nn.DistributedTransformerLMHead(**CFG))
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    @smp.step
    def eval_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        return loss

    g = torch.Generator().manual_seed(3)
    losses = []
    for i in range(4):
        ids = torch.randint(0, CFG["vocab_size"], (8, 16), generator=g)
        opt.zero_grad()
        losses.append(train_step(model, ids).reduce_mean().item())
        opt.step()
        if i % 2 == 1:
            with torch.no_grad():
                losses.append(eval_step(model, ids).reduce_mean().item())
    stats = dict(smp.state.exec_server.p2p.stats)
    return {"losses": losses, "stats": stats, "host": dict(smp.state.exec_server.p2p.host_stats)}


def test_static_mode_preposts_receives_and_changes_nothing():
    """``static_mode``: the first step of each step function records the receptions per source, later steps post
    them ahead of the announcing message; results are identical to the dynamic protocol, the train and eval step
    functions keep separate records."""
    plain = run_distributed(_static_worker, 2, False, timeout=240)
    static = run_distributed(_static_worker, 2, True, timeout=240)
    assert plain[0]["losses"] == static[0]["losses"], (plain[0]["losses"], static[0]["losses"])
    for r in static:
        assert r["stats"]["recorded_steps"] == 2 and r["stats"]["replayed_steps"] == 4, r["stats"]
        assert r["stats"]["preposted"] > 0
    assert all(r["stats"]["preposted"] == 0 for r in plain)


def test_static_mode_refuses_a_step_that_differs_from_its_record():
    from types import SimpleNamespace

    from smdistributed_modelparallel_b200.backend.exceptions import SMPRuntimeError
    from smdistributed_modelparallel_b200.runtime.server import PipelineP2P

    st = SimpleNamespace(core=SimpleNamespace(rank=lambda: 0, pp_rank=lambda: 0), cfg=SimpleNamespace(static_mode=True, fast_mode=False))
    p2p = PipelineP2P(st)
    posted = []
    p2p._post = lambda metas, src, device: (posted.append((tuple(metas), src)) or ([None] * len(metas), []))
    a, b = [((2, 4), torch.float32)], [((2, 8), torch.float32), ((2,), torch.int64)]
    p2p.begin_step("k")
    p2p.recv(a, 1, "cpu")
    p2p.recv(b, 1, "cpu")
    p2p.recv(a, 1, "cpu")
    p2p.drain()
    assert p2p.stats["recorded_steps"] == 1 and len(posted) == 3
    p2p.begin_step("k")                              # replay: the first two receptions are posted up front
    assert len(posted) == 5 and posted[3:] == [(tuple(a), 1), (tuple(b), 1)]
    p2p.recv(a, 1, "cpu")
    assert len(posted) == 6                          # ... and the window slides as they are consumed
    p2p.recv(b, 1, "cpu")
    with pytest.raises(SMPRuntimeError, match="static_mode"):
        p2p.recv(b, 1, "cpu")                        # the record says `a` comes third
    p2p._open = False
    p2p.begin_step("k")
    p2p.recv(a, 1, "cpu")
    with pytest.raises(SMPRuntimeError, match="identical steps"):
        p2p.drain()                                  # fewer receptions than recorded


def test_host_bus_fallback_for_pipeline_tensors(monkeypatch):
    """``SMP_DISABLE_D2D=1``: pipeline tensors travel over the host message bus instead of the p2p process groups
    (the reference's MPI fallback of its D2D transport); same numbers, with and without static mode."""
    plain = run_distributed(_static_worker, 2, False, timeout=240)
    monkeypatch.setenv("SMP_DISABLE_D2D", "1")
    host = run_distributed(_static_worker, 2, False, timeout=240)
    host_static = run_distributed(_static_worker, 2, True, timeout=240)
    assert plain[0]["losses"] == host[0]["losses"] == host_static[0]["losses"]
    assert all(r["host"]["sent"] == 0 for r in plain)
    for r in host + host_static:
        assert r["host"]["sent"] > 0 and r["host"]["received"] > 0 and r["host"]["bytes"] > 0
    assert host_static[0]["stats"]["replayed_steps"] == 4


def _seq_ckpt_worker(rank, world, configured):
    import copy

    import torch.nn as nn

    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 2, "ddp": False, "auto_partition": False,
              "default_partition": 0})
    torch.manual_seed(1)
    calls = {"n": 0}

    class Block(nn.Module):
        def __init__(self):
            super().__init__()
            self.lin = nn.Linear(16, 16)

        def forward(self, x):
            calls["n"] += 1
            return torch.tanh(self.lin(x))

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.inp = nn.Linear(8, 16)
            self.seq = nn.Sequential(*[Block() for _ in range(4)])
            self.out = nn.Linear(16, 2)

        def forward(self, x):
            h = self.inp(x)
            h = self.seq(h) if configured else smp.checkpoint_sequential(self.seq, h, strategy="group_2")
            return self.out(h)

    net = Net()
    oracle = copy.deepcopy(net)
    for i, layer in enumerate(net.seq):
        smp.set_partition(layer, 0 if i < 2 else 1)
    if configured:
        smp.set_activation_checkpointing(net.seq, strategy="each")
    model = smp.DistributedModel(net)

    @smp.step
    def step(model, x):
        loss = model(x).pow(2).mean()
        model.backward(loss)
        return loss

    x = torch.randn(4, 8, generator=torch.Generator().manual_seed(2))
    loss = step(model, x).reduce_mean().item()
    mine = calls["n"]
    ol = oracle(x).pow(2).mean() if configured else None
    if ol is None:
        h = oracle.inp(x)
        for blk in oracle.seq:
            h = blk(h)
        ol = oracle.out(h).pow(2).mean()
    ol.backward()
    ref = dict(oracle.named_parameters())
    worst = max((p.grad - ref[n].grad).abs().max().item() for n, p in model.local_named_parameters())
    return {"loss": loss, "oracle": ol.item(), "worst": worst, "calls": mine}


@pytest.mark.parametrize("configured", [False, True])
def test_checkpointed_sequential_spanning_two_stages(configured):
    """An ``nn.Sequential`` cut across two pipeline ranks, checkpointed through ``smp.checkpoint_sequential`` or
    ``smp.set_activation_checkpointing``: each rank checkpoints its own run of layers (every local layer runs twice
    per microbatch: forward and recomputation), results match the plain model."""
    res = run_distributed(_seq_ckpt_worker, 2, configured, timeout=240)
    for r in res:
        assert abs(r["loss"] - r["oracle"]) < 5e-6 and r["worst"] < 5e-6, r
        assert r["calls"] == 2 * 2 * 2, r          # 2 local layers x 2 microbatches x (forward + recompute)


def _amp_worker(rank, world):
    import torch.nn as nn

    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 2, "ddp": False})
    torch.manual_seed(1)
    net = nn.Sequential(nn.Linear(8, 16), nn.Tanh(), nn.Linear(16, 16), nn.Tanh(), nn.Linear(16, 2))
    model = smp.DistributedModel(net)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))
    scaler = smp.amp.GradScaler(init_scale=2.0 ** 8, growth_interval=2)
    seen = []
    for m in net:
        if isinstance(m, nn.Linear):
            m.register_forward_hook(lambda mod, a, out: seen.append(out.dtype))

    @smp.step
    def step(model, x):
        with torch.autocast("cpu", dtype=torch.bfloat16):
            loss = model(x).float().pow(2).mean()
        model.backward(scaler.scale(loss))
        return loss

    out = []
    for i in range(3):
        x = torch.randn(4, 8, generator=torch.Generator().manual_seed(2 + i))
        opt.zero_grad()
        out.append(step(model, x).reduce_mean().item())
        scaler.step(opt)
        scaler.update()
    return {"losses": out, "scale": scaler.get_scale(), "dtypes": sorted({str(d) for d in seen}),
            "n_local": len(list(model.local_parameters()))}


def test_autocast_and_grad_scaler_across_pipeline_stages():
    """The caller's autocast region covers modules executed on other stages; ``smp.amp.GradScaler`` works on stages
    that never ran the step function (they start from the same scale) and all stages agree on the scale."""
    res = run_distributed(_amp_worker, 2, timeout=240)
    assert res[0]["losses"] == res[1]["losses"] and res[0]["losses"][-1] < res[0]["losses"][0]
    assert res[0]["scale"] == res[1]["scale"] == 2.0 ** 9          # grew once after two clean steps
    for r in res:
        assert r["n_local"] > 0 and r["dtypes"] == ["torch.bfloat16"], r


def _failing_stage_worker(rank, world):
    import torch.nn as nn

    import smdistributed_modelparallel_b200 as smp

    class Boom(nn.Module):
        def __init__(self):
            super().__init__()
            self.lin = nn.Linear(8, 8)
            self.fail = False

        def forward(self, x):
            if self.fail:
                raise ValueError("deliberate failure on the second stage")
            return self.lin(x)

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 2, "ddp": False, "auto_partition": False,
              "default_partition": 0})
    net = nn.Sequential(nn.Linear(8, 8), Boom(), nn.Linear(8, 2))
    smp.set_partition(net[1], 1)
    model = smp.DistributedModel(net)

    @smp.step
    def step(model, x):
        loss = model(x).pow(2).mean()
        model.backward(loss)
        return loss

    x = torch.randn(4, 8)
    step(model, x)
    net[1].fail = True
    try:
        step(model, x)
    except Exception as e:   # noqa: BLE001
        return f"{type(e).__name__}: {e}"
    return None


def test_failure_inside_a_remote_module_fails_both_stages():
    """A module raising on the second stage: that stage re-raises its own exception, the leader gets it back as
    the failure of the remote call (with the remote traceback); nobody hangs."""
    res = run_distributed(_failing_stage_worker, 2, timeout=180)
    assert res[0].startswith("SMPRuntimeError") and "deliberate failure on the second stage" in res[0]
    assert res[1].startswith("ValueError: deliberate failure")


def _reinit_worker(rank, world):
    import smdistributed_modelparallel_b200 as smp

    res = []
    for conf in ({"tensor_parallel_degree": 2}, {"pipeline_parallel_degree": 2},
                 {"tensor_parallel_degree": 2, "pipeline_parallel_degree": 2}):
        smp.init(dict(conf, microbatches=2, ddp=True, ddp_dist_backend="gloo"))
        torch.manual_seed(1)
        with smp.model_creation(tensor_parallelism=smp.tp_size() > 1):
            module = smp.nn.DistributedTransformerLMHead(**CFG)
        model = smp.DistributedModel(module)
        opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))

        @smp.step
        def step(model, ids):
            loss, _ = model((ids, None, None, None, ids))
            model.backward(loss)
            return loss

        ids = torch.randint(0, CFG["vocab_size"], (4, 16), generator=torch.Generator().manual_seed(3))
        opt.zero_grad()
        first = step(model, ids).reduce_mean().item()
        opt.step()
        second = step(model, ids).reduce_mean().item()
        res.append((smp.pp_size(), smp.tp_size(), smp.dp_size(), first, second))
    return res


def test_reinitialising_with_other_degrees_in_one_process():
    """``smp.init`` again with another layout (tp2 -> pp2 -> pp2 x tp2 on the same four processes): groups, engine
    and partition are rebuilt each time and training works in each."""
    res = run_distributed(_reinit_worker, 4, timeout=300)
    for r in res:
        assert [x[:3] for x in r] == [(1, 2, 4), (2, 1, 2), (2, 2, 2)]
        assert all(second < first for *_, first, second in r)
    assert res[0][0][3] == res[0][2][3]            # the same tensor-parallel model whether or not it is pipelined


def _tuned_worker(rank, world, record_steps):
    import os

    import smdistributed_modelparallel_b200 as smp

    os.environ["SMP_SCHEDULE_RECORD_STEPS"] = str(record_steps)
    smp.init({"pipeline_parallel_degree": 2, "tensor_parallel_degree": 2, "microbatches": 4, "ddp": True,
              "active_microbatches": 4, "ddp_dist_backend": "gloo"})
    torch.manual_seed(7)
    with smp.tensor_parallelism(enabled=True):
        module = smp.nn.DistributedTransformerLMHead(**CFG)
    model = smp.DistributedModel(module)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    torch.manual_seed(11)
    all_ids = torch.randint(0, CFG["vocab_size"], (7, 16, 16))
    losses, warms = [], []
    for it in range(7):
        ids = all_ids[it, smp.tp_rank() * 8:(smp.tp_rank() + 1) * 8]
        opt.zero_grad()
        out = train_step(model, ids)
        opt.step()
        losses.append(out.reduce_mean().item())
        tuners = list(smp.state.exec_server._tuners.values())
        warms.append(tuners[0].choice if tuners else None)
    t = tuners[0] if tuners else None
    return {"losses": losses, "warms": warms, "pp_rank": smp.pp_rank(),
            "timed": {c: len(v) for c, v in t.times.items()} if t else None,
            "cands": t.candidates if t else None}


def test_schedule_record_and_pick_fastest():
    """SMP_SCHEDULE_RECORD_STEPS=4 on pp2 x tp2: steps 1-4 are timed under the candidate warm-up depths, at step 5
    rank 0's fastest candidate becomes the schedule of both leaders, and the numbers are those of an untuned run
    (the schedule changes the order of work, not the work)."""
    tuned = run_distributed(_tuned_worker, 4, 4, timeout=420)
    plain = run_distributed(_tuned_worker, 4, 0, timeout=420)
    leaders = [r for r in tuned if r["pp_rank"] == 0]
    assert len(leaders) == 2
    for r in leaders:
        assert r["cands"] == [2, 3, 1, 4]
        assert r["warms"][:5] == [None] * 5 and r["warms"][5] in r["cands"]      # decided when step 5 starts
        assert r["warms"][5] == r["warms"][6] == leaders[0]["warms"][5]           # the same on every leader
        assert r["timed"] == {2: 1, 3: 1, 1: 1, 4: 1}                             # step 0 untimed, one step each
    for r in tuned:
        if r["pp_rank"] != 0:
            assert r["cands"] is None                                             # only leaders schedule
    for a, b in zip(tuned, plain):
        assert a["losses"] == pytest.approx(b["losses"], abs=2e-5)
    assert all(r["cands"] is None for r in plain)


def _saved_partition_worker(rank, world, tmp):
    import os

    import torch.nn as nn
    import torch.nn.functional as F

    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.backend.exceptions import StepFunctionCalledError

    class Net(nn.Module):
        def __init__(self, dim=16):
            super().__init__()
            self.lin0_0, self.lin0_1, self.lin0_2 = nn.Linear(dim, dim), nn.Linear(dim, dim), nn.Linear(dim, dim)
            self.lin1, self.lin2, self.lin3 = nn.Linear(dim, dim), nn.Linear(dim, dim), nn.Linear(dim, 10)

        def forward(self, x):
            for m in (self.lin0_0, self.lin0_1, self.lin0_2, self.lin1, self.lin2):
                x = m(x)
            return self.lin3(x)

    @smp.step
    def train_step(model, x, y):
        loss = F.cross_entropy(model(x), y)
        model.backward(loss)
        return loss

    want = {"lin0_0": 0, "lin0_1": 0, "lin0_2": 0, "lin1": 1, "lin2": 2, "lin3": 3}

    def placed(model):
        mm = smp.state.module_manager
        return {n: mm.get_partition(getattr(model.module, n)) for n in want}

    x, y = torch.randn(8, 16), torch.randint(0, 10, (8,))
    out = {}
    base = {"pipeline_parallel_degree": 4, "microbatches": 2, "ddp": False}
    pfile = os.path.join(tmp, "partition.data")
    # 1. manual partition; the assignment is also written to partition_file
    smp.init({**base, "auto_partition": False, "default_partition": 0, "partition_file": pfile})
    model = smp.DistributedModel(Net())
    for n, r in want.items():
        smp.set_partition(getattr(model.module, n), r)
    train_step(model, x, y)
    out["manual"] = placed(model)
    info = smp.state.module_manager.get_model_partition_info()
    smp.barrier()
    out["file_written"] = os.path.exists(pfile)
    # 2. the saved assignment wins over later set_partition calls; loading after a step is refused
    smp.reset()
    smp.init({**base, "auto_partition": False, "default_partition": 0})
    model = smp.DistributedModel(Net())
    model.load_saved_partition(info)
    for n, r in {"lin0_0": 0, "lin0_1": 0, "lin0_2": 1, "lin1": 2, "lin2": 3, "lin3": 3}.items():
        smp.set_partition(getattr(model.module, n), r)
    train_step(model, x, y)
    out["loaded_over_manual"] = placed(model)
    try:
        model.load_saved_partition(info)
        out["refused_after_step"] = False
    except StepFunctionCalledError:
        out["refused_after_step"] = True
    # 3. ... and over the auto-partitioner
    smp.reset()
    smp.init({**base, "auto_partition": True})
    model = smp.DistributedModel(Net())
    model.load_saved_partition(info)
    train_step(model, x, y)
    out["loaded_over_auto"] = placed(model)
    # 4. config keys: load_partition + partition_file
    smp.reset()
    smp.init({**base, "auto_partition": True, "load_partition": True, "partition_file": pfile})
    model = smp.DistributedModel(Net())
    loss = train_step(model, x, y)
    out["loaded_from_file"] = placed(model)
    out["loss"] = loss.reduce_mean().item()
    return out


def test_saved_partition_is_reloaded(tmp_path):
    """model.load_saved_partition(state.module_manager.get_model_partition_info()) and the load_partition /
    partition_file config keys (reference test/torch/mpi_4ps/test_load_partition.py:60-136)."""
    want = {"lin0_0": 0, "lin0_1": 0, "lin0_2": 0, "lin1": 1, "lin2": 2, "lin3": 3}
    for r in run_distributed(_saved_partition_worker, 4, str(tmp_path), timeout=420):
        assert r["manual"] == want and r["file_written"]
        assert r["loaded_over_manual"] == want and r["refused_after_step"]
        assert r["loaded_over_auto"] == want
        assert r["loaded_from_file"] == want
        assert r["loss"] == r["loss"]
