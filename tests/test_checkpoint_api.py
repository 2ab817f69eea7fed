"""smp.save_checkpoint / smp.resume_from_checkpoint / smp.save / smp.load on one process (file layout,
``newest`` bookkeeping, retention, user content, lazy load before the model exists, config verification):
reference test/torch/mpi_hybrid/test_checkpoint_api.py tier."""

import os
import tempfile

import pytest
import torch
import torch.nn as nn

import smdistributed_modelparallel_b200 as smp
from smdistributed_modelparallel_b200.backend.exceptions import SMPValidationError


def _setup(cfg=None, seed=0):
    smp.reset()
    smp.init(dict({"microbatches": 2}, **(cfg or {})))
    torch.manual_seed(seed)
    net = nn.Sequential(nn.Linear(8, 16), nn.Tanh(), nn.Linear(16, 4))
    model = smp.DistributedModel(net)
    opt = smp.DistributedOptimizer(torch.optim.Adam(model.parameters(), lr=1e-2))

    @smp.step
    def train_step(model, x, y):
        loss = ((model(x) - y) ** 2).mean()
        model.backward(loss)
        return loss

    return model, opt, train_step


def _train(model, opt, step, n=2, seed=1):
    torch.manual_seed(seed)
    for _ in range(n):
        x, y = torch.randn(4, 8), torch.randn(4, 4)
        opt.zero_grad()
        step(model, x, y)
        opt.step()


def test_partial_round_trip_and_layout():
    with tempfile.TemporaryDirectory() as d:
        model, opt, step = _setup()
        _train(model, opt, step)
        smp.save_checkpoint(d, tag="step2", partial=True, model=model, optimizer=opt, user_content={"epoch": 3})
        files = set(os.listdir(os.path.join(d, "step2_partial")))
        assert {"model_0_0.pt", "optimizer_states_0_0.pt", "smp_config.pt", "user_content.pt"} <= files
        assert open(os.path.join(d, "newest")).read() == "step2_partial\n"
        want = {k: v.clone() for k, v in model.state_dict().items()}
        want_opt = opt.state_dict()["optimizer_state_dict"]
        want_opt = {"state": {k: {n: (t.clone() if torch.is_tensor(t) else t) for n, t in st.items()}
                              for k, st in want_opt["state"].items()}}
        _train(model, opt, step, seed=9)              # diverge
        assert any(not torch.equal(want[k], v) for k, v in model.state_dict().items())
        uc = smp.resume_from_checkpoint(d)            # newest tag
        assert uc == {"epoch": 3}
        for k, v in model.state_dict().items():
            assert torch.equal(want[k], v), k
        got_opt = opt.state_dict()["optimizer_state_dict"]
        for pid, st in want_opt["state"].items():
            for name, val in st.items():
                if torch.is_tensor(val):
                    assert torch.allclose(val.float(), got_opt["state"][pid][name].float()), (pid, name)


def test_retention_keeps_the_last_n():
    with tempfile.TemporaryDirectory() as d:
        model, opt, step = _setup()
        for i in range(4):
            smp.save_checkpoint(d, tag=f"t{i}", model=model, optimizer=opt, num_kept_partial_checkpoints=2)
        dirs = sorted(x for x in os.listdir(d) if x.endswith("_partial"))
        assert dirs == ["t2_partial", "t3_partial"]
        assert open(os.path.join(d, "newest")).read().split() == ["t2_partial", "t3_partial"]


def test_full_checkpoint_loads_into_plain_module():
    with tempfile.TemporaryDirectory() as d:
        model, opt, step = _setup()
        _train(model, opt, step)
        smp.save_checkpoint(d, tag="final", partial=False, model=model, user_content=[1, 2])
        assert os.path.isfile(os.path.join(d, "final")) and os.path.isfile(os.path.join(d, "user_content_final"))
        sd = torch.load(os.path.join(d, "final"), weights_only=False)
        plain = nn.Sequential(nn.Linear(8, 16), nn.Tanh(), nn.Linear(16, 4))
        plain.load_state_dict({k: v for k, v in sd.items() if not k.startswith("_smp")})
        x = torch.randn(3, 8)
        assert torch.allclose(plain(x), model.module(x))
        with pytest.raises(SMPValidationError):
            smp.resume_from_checkpoint(d, partial=False)       # a full checkpoint needs an explicit tag
        model2, opt2, _ = _setup(seed=5)
        assert smp.resume_from_checkpoint(d, tag="final", partial=False) == [1, 2]
        assert torch.allclose(model2.module(x), plain(x))


def test_resume_before_model_creation_is_deferred():
    with tempfile.TemporaryDirectory() as d:
        model, opt, step = _setup()
        _train(model, opt, step)
        smp.save_checkpoint(d, tag="a", model=model, optimizer=opt)
        want = {k: v.clone() for k, v in model.state_dict().items()}
        smp.reset()
        smp.init({"microbatches": 2})
        smp.resume_from_checkpoint(d, tag="a")
        torch.manual_seed(77)
        model2 = smp.DistributedModel(nn.Sequential(nn.Linear(8, 16), nn.Tanh(), nn.Linear(16, 4)))
        opt2 = smp.DistributedOptimizer(torch.optim.Adam(model2.parameters(), lr=1e-2))

        @smp.step
        def fwd(model, x):
            return model(x)

        fwd(model2, torch.randn(4, 8))                # first step applies the deferred state
        for k, v in model2.state_dict().items():
            assert torch.equal(want[k], v), k
        assert opt2 is not None


def test_errors():
    with tempfile.TemporaryDirectory() as d:
        model, opt, step = _setup()
        with pytest.raises(SMPValidationError):
            smp.resume_from_checkpoint(d)              # no `newest`
        with pytest.raises(SMPValidationError):
            smp.resume_from_checkpoint(d, tag="missing")
        with pytest.raises(Exception):
            smp.save_checkpoint(d, tag=123, model=model)
        smp.save_checkpoint(d, tag="x", model=model, optimizer=opt)
        # a different parallel layout is refused when the saved config is verified
        smp.reset()
        smp.init({"microbatches": 4})
        nn_model = smp.DistributedModel(nn.Sequential(nn.Linear(8, 16), nn.Tanh(), nn.Linear(16, 4)))
        assert nn_model is not None
        smp.resume_from_checkpoint(d, tag="x", load_optimizer=False)   # microbatches may differ


def test_save_load_partial_names():
    with tempfile.TemporaryDirectory() as d:
        _setup()
        smp.save({"a": 1}, os.path.join(d, "obj"), partial=True)
        assert os.path.isfile(os.path.join(d, "obj_0_0.pt"))
        assert smp.load(os.path.join(d, "obj"), partial=True) == {"a": 1}
        smp.save({"b": 2}, os.path.join(d, "whole.pt"), partial=False)
        assert smp.load(os.path.join(d, "whole.pt"), partial=False) == {"b": 2}


def test_bf16_masters_survive_a_resume():
    """16-bit training keeps fp32 master weights in the optimizer wrapper; a partial checkpoint stores them
    (fp16_states) and the run continues bit-for-bit after a resume."""
    with tempfile.TemporaryDirectory() as d:
        def build(seed):
            smp.reset()
            smp.init({"microbatches": 2, "bf16": True})
            torch.manual_seed(seed)
            with smp.model_creation(dtype=torch.bfloat16):
                net = nn.Sequential(nn.Linear(8, 16), nn.Tanh(), nn.Linear(16, 4))
            model = smp.DistributedModel(net)
            opt = smp.DistributedOptimizer(torch.optim.Adam(model.parameters(), lr=1e-2))

            @smp.step
            def train_step(model, x, y):
                loss = ((model(x).float() - y) ** 2).mean()
                model.backward(loss)
                return loss

            return model, opt, train_step

        def run(model, opt, step, n, seed):
            g = torch.Generator().manual_seed(seed)
            out = []
            for _ in range(n):
                x, y = torch.randn(4, 8, generator=g).bfloat16(), torch.randn(4, 4, generator=g)
                opt.zero_grad()
                out.append(step(model, x, y).reduce_mean().item())
                opt.step()
            return out

        model, opt, step = build(0)
        run(model, opt, step, 3, 1)
        smp.save_checkpoint(d, tag="b", model=model, optimizer=opt)
        assert "fp16_states_0_0.pt" in os.listdir(os.path.join(d, "b_partial"))
        want = run(model, opt, step, 3, 2)
        model2, opt2, step2 = build(42)
        smp.resume_from_checkpoint(d, tag="b")
        got = run(model2, opt2, step2, 3, 2)
        assert want == got, (want, got)


def test_reference_shaped_fp16_states():
    """fp16_states files carry the loss scaler as a pickled OBJECT whose class resolves under
    smdistributed.modelparallel.torch.fp16.loss_scaler (reference torch/optimizers/optimizer.py:185-190,270-271).
    A file shaped like the reference's loads here, and what this library writes has the same shape."""
    import pickle

    from smdistributed.modelparallel.torch.fp16.loss_scaler import DynamicLossScaler as RefPathScaler

    with tempfile.TemporaryDirectory() as d:
        smp.reset()
        smp.init({"microbatches": 1, "fp16": True})
        torch.manual_seed(0)
        with smp.model_creation(dtype=torch.float16):
            net = nn.Sequential(nn.Linear(8, 8), nn.Tanh(), nn.Linear(8, 4))
        model = smp.DistributedModel(net)
        opt = smp.DistributedOptimizer(torch.optim.Adam(model.parameters(), lr=1e-3), dynamic_loss_scale=True)

        @smp.step
        def train_step(model, x):
            loss = model(x).float().pow(2).mean()
            model.backward(loss)
            return loss

        opt.zero_grad()
        train_step(model, torch.randn(2, 8).half())
        opt.step()

        # (1) what this library writes: an object at the reference's class path with the reference's fields
        mine = opt.local_fp16_state_dict()
        ls = mine["loss_scaler"]
        assert type(ls).__module__ == "smdistributed.modelparallel.torch.fp16.loss_scaler"
        assert type(ls).__name__ == "DynamicLossScaler" and ls.model is None
        for field in ("cur_scale", "cur_iter", "last_overflow_iter", "scale_factor", "scale_window", "min_scale",
                      "delayed_shift", "cur_hysteresis", "consecutive_hysteresis"):
            assert hasattr(ls, field), field
        assert set(mine) == {"loss_scaler", "dynamic_loss_scale", "overflow", "first_closure_call_this_step",
                             "fp32_from_fp16"}
        pickle.loads(pickle.dumps(mine))

        # (2) a file shaped like the reference's: the object is built field by field the way the reference's
        # constructor leaves it (no `dynamic` attribute, a `model` slot), saved with torch.save
        ref_obj = RefPathScaler.__new__(RefPathScaler)
        ref_obj.__dict__.update(cur_scale=2.0 ** 14, cur_iter=37, last_overflow_iter=30, scale_factor=2.0,
                                scale_window=1000, min_scale=1, delayed_shift=2, cur_hysteresis=1,
                                consecutive_hysteresis=False, model=None)
        masters = [[m.detach().cpu().clone() + 0.25 for m in grp] for grp in opt.fp16_optimizer.fp32_from_fp16_groups]
        path = os.path.join(d, "fp16_states_0_0.pt")
        torch.save({"loss_scaler": ref_obj, "dynamic_loss_scale": True, "overflow": False,
                    "first_closure_call_this_step": True, "fp32_from_fp16": masters}, path)
        loaded = torch.load(path, weights_only=False)
        opt.load_optimizer_checkpoint_state({"optimizer_state_dict": opt.optimizer.state_dict(), "fp16": loaded})
        b = opt.fp16_optimizer
        assert b.loss_scaler.cur_scale == 2.0 ** 14 and b.loss_scaler.cur_iter == 37
        assert b.loss_scaler.cur_hysteresis == 1 and b.loss_scaler.delayed_shift == 2
        assert b.loss_scaler.dynamic is True                      # own field survives a reference-shaped load
        for cur, saved in zip(b.fp32_from_fp16_groups, masters):
            for c, s in zip(cur, saved):
                assert torch.equal(c.detach().cpu(), s)
        # the older dict form still loads
        b.loss_scaler.load_state_dict({"cur_scale": 8.0})
        assert b.loss_scaler.cur_scale == 8.0
