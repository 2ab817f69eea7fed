"""Loss scalers, fused optimizers (CPU restatements of the multi-tensor kernels vs torch / reference formulas),
the TP-consistent RNG manager, the NCCL throttler and the ZeRO-2D config mapping (reference tiers: fp16 /
optimizer unit tests, test/backend/test_config.py)."""

import math

import pytest
import torch

import smdistributed_modelparallel_b200 as smp
from smdistributed_modelparallel_b200.backend.config import ModelParallelConfig
from smdistributed_modelparallel_b200.backend.zero_config import construct_zero2d_config_dict
from smdistributed_modelparallel_b200.precision.loss_scaler import DynamicLossScaler, LossScaler
from smdistributed_modelparallel_b200.utils.random import RngManager
from smdistributed_modelparallel_b200.utils.throttler import NCCLThrottler


def test_static_and_dynamic_loss_scalers():
    s = LossScaler(128.0)
    s.update_scale(True)
    assert s.loss_scale == 128.0
    d = DynamicLossScaler(init_scale=2.0 ** 8, scale_factor=2.0, scale_window=3, min_scale=1.0)
    d.update_scale(True)                                   # overflow halves at once (delayed_shift 1)
    assert d.loss_scale == 2.0 ** 7
    for _ in range(3):
        d.update_scale(False)                              # three clean steps after the overflow -> doubles
    assert d.loss_scale == 2.0 ** 8
    for _ in range(20):
        d.update_scale(True)
    assert d.loss_scale == 1.0                             # clamped at min_scale
    sd = d.state_dict()
    d2 = DynamicLossScaler()
    d2.load_state_dict(sd)
    assert d2.loss_scale == 1.0 and d2.cur_iter == d.cur_iter
    h = DynamicLossScaler(init_scale=16.0, delayed_shift=2)
    h.update_scale(True)
    assert h.loss_scale == 16.0                            # first overflow is absorbed by the hysteresis
    h.update_scale(True)
    assert h.loss_scale == 8.0
    p = torch.nn.Parameter(torch.zeros(3))
    p.grad = torch.tensor([1.0, float("inf"), 0.0])
    assert d.has_overflow_serial([p])
    p.grad = torch.tensor([1.0, 2.0, 0.0])
    assert not d.has_overflow_serial([p])


def _params(seed):
    torch.manual_seed(seed)
    ps = [torch.nn.Parameter(torch.randn(n)) for n in (7, 64, 1025)]
    gs = [torch.randn_like(p) for p in ps]
    return ps, gs


def _run(opt_cls, ref_cls, kw, ref_kw, steps=3):
    ps, gs = _params(0)
    rs = [torch.nn.Parameter(p.detach().clone()) for p in ps]
    opt, ref = opt_cls(ps, **kw), ref_cls(rs, **ref_kw)
    for _ in range(steps):
        for p, r, g in zip(ps, rs, gs):
            p.grad, r.grad = g.clone(), g.clone()
        opt.step()
        ref.step()
    return max((p - r).abs().max().item() for p, r in zip(ps, rs))


def test_fused_adam_matches_torch():
    from smdistributed_modelparallel_b200.optimizers import FusedAdam

    assert _run(FusedAdam, torch.optim.AdamW, dict(lr=1e-2, weight_decay=0.1), dict(lr=1e-2, weight_decay=0.1)) < 1e-6
    assert _run(FusedAdam, torch.optim.Adam, dict(lr=1e-2, weight_decay=0.1, adam_w_mode=False),
                dict(lr=1e-2, weight_decay=0.1)) < 1e-6
    with pytest.raises(RuntimeError):
        FusedAdam([torch.nn.Parameter(torch.zeros(1))], amsgrad=True)


def test_fused_sgd_and_adagrad_match_torch():
    from smdistributed_modelparallel_b200.optimizers.fused import FusedAdagrad, FusedSGD

    assert _run(FusedSGD, torch.optim.SGD, dict(lr=0.1, momentum=0.9, weight_decay=0.01),
                dict(lr=0.1, momentum=0.9, weight_decay=0.01)) < 1e-6
    assert _run(FusedSGD, torch.optim.SGD, dict(lr=0.1, momentum=0.9, nesterov=True),
                dict(lr=0.1, momentum=0.9, nesterov=True)) < 1e-6
    assert _run(FusedAdagrad, torch.optim.Adagrad, dict(lr=0.05, eps=1e-10), dict(lr=0.05, eps=1e-10)) < 1e-6


def test_fused_lamb_trust_ratio():
    """One LAMB step by hand: Adam direction + weight decay, scaled per tensor by |w| / |update|."""
    from smdistributed_modelparallel_b200.optimizers import FusedLAMB

    torch.manual_seed(1)
    w = torch.nn.Parameter(torch.randn(50))
    g = torch.randn(50) * 0.1
    w0 = w.detach().clone()
    opt = FusedLAMB([w], lr=0.01, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.01, max_grad_norm=1e9)
    w.grad = g.clone()
    opt.step()
    m = 0.1 * g
    v = 0.001 * g * g
    upd = (m / (1 - 0.9)) / ((v / (1 - 0.999)).sqrt() + 1e-6) + 0.01 * w0
    ratio = w0.norm() / upd.norm()
    expect = w0 - 0.01 * ratio * upd
    assert torch.allclose(w.detach(), expect, atol=1e-5), (w.detach() - expect).abs().max()


def test_fused_novograd_runs_and_decreases_loss():
    from smdistributed_modelparallel_b200.optimizers import FusedNovoGrad

    torch.manual_seed(2)
    w = torch.nn.Parameter(torch.randn(20))
    opt = FusedNovoGrad([w], lr=0.05)
    first = None
    for _ in range(30):
        opt.zero_grad()
        loss = (w ** 2).sum()
        loss.backward()
        opt.step()
        first = first if first is not None else loss.item()
    assert (w ** 2).sum().item() < first * 0.95          # per-tensor normalised steps: slow but steady


def test_rng_manager_is_a_separate_reproducible_stream():
    r = RngManager()
    r.set_state(42)
    torch.manual_seed(7)
    before = torch.get_rng_state().clone()
    with r.consistent_rng_state():
        a = torch.rand(4)
    assert torch.equal(torch.get_rng_state(), before)       # the default generator was not advanced
    with r.consistent_rng_state():
        b = torch.rand(4)
    assert not torch.equal(a, b)                            # the managed stream advances between uses
    r2 = RngManager()
    r2.set_state(42)
    with r2.consistent_rng_state():
        a2 = torch.rand(4)
    assert torch.equal(a, a2)                               # same seed -> same stream on every tp rank
    st = r.get_states()
    with r.consistent_rng_state():
        c = torch.rand(4)
    r.set_states(st)
    with r.consistent_rng_state():
        c2 = torch.rand(4)
    assert torch.equal(c, c2)


def test_throttler_runs_everything_in_order():
    t = NCCLThrottler()
    t.set_throttle_limit(2)
    out = []
    for i in range(7):
        t.run(out.append, i)
    t.reset()
    assert out == list(range(7))


def test_zero2d_config_mapping():
    cfg = ModelParallelConfig({"sharded_data_parallel_degree": 4, "ddp": True, "bf16": True,
                               "sdp_reduce_bucket_size": int(2e8), "sdp_gradient_clipping": 0.5,
                               "sdp_param_persistence_threshold": 12345})

    class Core:
        _size = 8

    d = construct_zero2d_config_dict(cfg, Core())
    z = d["zero_optimization"]
    assert z["stage"] == 3 and z["zero2d_shard_size"] == 4
    assert z["reduce_bucket_size"] == int(2e8) and z["stage3_param_persistence_threshold"] == 12345
    assert d["gradient_clipping"] == 0.5 and d["bf16"]["enabled"] and not d["fp16"]["enabled"]
    Core._size = 6
    with pytest.raises(Exception):
        construct_zero2d_config_dict(cfg, Core())           # 4 does not divide 6
    assert construct_zero2d_config_dict(ModelParallelConfig({}), Core()) == {}


def test_amp_grad_scaler_single_process():
    smp.reset()
    smp.init({"microbatches": 1})
    scaler = smp.amp.GradScaler(init_scale=4.0, enabled=False)
    w = torch.nn.Parameter(torch.ones(3))
    opt = torch.optim.SGD([w], lr=0.1)
    loss = (w * 2).sum()
    scaler.scale(loss).backward()
    scaler.step(opt)
    scaler.update()
    assert torch.allclose(w.detach(), torch.full((3,), 0.8))
    assert math.isclose(scaler.get_scale(), 1.0)


def test_fp16_dynamic_loss_scale_skips_overflowing_steps():
    """fp16 with dynamic loss scaling: a step whose scaled gradients overflow is skipped (parameters and optimizer
    state untouched, scale halved); once the scale fits, training proceeds."""
    import torch.nn as nn

    smp.reset()
    smp.init({"microbatches": 1, "fp16": True})
    torch.manual_seed(0)
    with smp.model_creation(dtype=torch.float16):
        net = nn.Sequential(nn.Linear(8, 16), nn.Tanh(), nn.Linear(16, 4))
    model = smp.DistributedModel(net)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1), dynamic_loss_scale=True,
                                   dynamic_loss_args={"init_scale": 2.0 ** 30, "scale_window": 1000})

    @smp.step
    def train_step(model, x, y):
        loss = ((model(x).float() - y) ** 2).mean()
        model.backward(loss)
        return loss

    x, y = torch.randn(4, 8).half(), torch.randn(4, 4)
    before = [p.detach().clone() for p in model.module.parameters()]
    scales, changed = [], []
    for _ in range(24):
        opt.zero_grad()
        train_step(model, x, y)
        opt.step()
        scales.append(float(opt.loss_scale))
        changed.append(any(not torch.equal(a, b.detach()) for a, b in zip(before, model.module.parameters())))
    assert scales[0] == 2.0 ** 29 and not changed[0]          # the first step overflowed and was skipped
    first_ok = changed.index(True)
    assert first_ok > 0 and scales[first_ok] == scales[first_ok - 1]   # the scale stops shrinking once a step fits
    assert all(s == scales[-1] for s in scales[first_ok:])
    assert all(torch.isfinite(p).all() for p in model.module.parameters())


def test_bit16_master_inspection():
    """fp32 masters / master gradients of the 16-bit optimizer can be inspected (unscaled), the masters follow a
    change of the model weights made from outside, and the Megatron-style getters of the state answer."""
    import torch.nn as nn

    smp.reset()
    smp.init({"microbatches": 1, "bf16": True})
    torch.manual_seed(0)
    with smp.model_creation(dtype=torch.bfloat16):
        net = nn.Sequential(nn.Linear(8, 16), nn.Tanh(), nn.Linear(16, 4))
    model = smp.DistributedModel(net)
    opt = smp.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1))

    @smp.step
    def train_step(model, x, y):
        loss = ((model(x).float() - y) ** 2).mean()
        model.backward(loss)
        return loss

    x, y = torch.randn(4, 8).bfloat16(), torch.randn(4, 4)
    opt.zero_grad()
    train_step(model, x, y)
    inner = opt
    masters = inner.get_master_params()
    assert sum(len(g) for g in masters) == 4 and all(m.dtype == torch.float32 for g in masters for m in g)
    grads = inner.inspect_master_grad_data()
    flat = [g for grp in grads for g in grp]
    assert all(g is not None and g.dtype == torch.float32 and torch.isfinite(g).all() for g in flat)
    assert any(g.abs().sum() > 0 for g in flat)
    scaled = [g for grp in inner.get_master_grads() for g in grp]
    ratio = float(inner.loss_scale)
    assert all(torch.allclose(a, b * ratio) for a, b in zip(scaled, flat))
    assert set(inner.state_dict_for_save_checkpoint()) == set(opt.fp16_optimizer.state_dict())
    with torch.no_grad():
        net[0].weight.fill_(0.25)
    inner.reload_model_params()
    assert all(torch.equal(m, torch.full_like(m, 0.25)) for g in inner.get_master_params() for m in g
               if m.shape == net[0].weight.shape)
    st = smp.state
    assert st.get_data_parallel_world_size() == 1 and st.get_model_parallel_world_size() == 1
    assert st.get_data_parallel_rank() == 0 and st.get_model_parallel_rank() == 0 and not st.is_tracing()
