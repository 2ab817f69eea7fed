"""Hybrid tier: one HuggingFace GPT-2 trained densely in fp32 and the same weights trained under smp in a matrix of
layouts x precisions x options on four ranks; after every step the loss, and after the last step every parameter (gathered
and translated back to HuggingFace names), must track the dense run - with SGD a parameter after a step is the sliced
gradient, so this is the gradient comparison of the reference's tier (reference test/torch/mpi_hybrid/smp_test_base.py:175-971,
test_gpt_grad.py:88-193, test_opt_sharding.py:15-17, test_zero.py:40-76).

The matrix runs on CPU / gloo in the default tier (four processes) and, unchanged, on four GPUs over NCCL under
``-m gpu`` with ``SMP_RUN_HYBRID_GPU=1`` - there the 16-bit cases go through the tcgen05 GEMMs, the flash attention
kernels, the fused tensor-parallel collectives and the peer-memory ZeRO-3 kernels.
"""

import pytest
import torch

transformers = pytest.importorskip("transformers")

from dist_utils import run_distributed

VOCAB, SEQ, STEPS = 128, 32, 3


def _hf(seed=0):
    from transformers import GPT2Config, GPT2LMHeadModel

    cfg = GPT2Config(n_layer=4, n_head=4, n_embd=64, n_positions=SEQ, vocab_size=VOCAB, resid_pdrop=0.0, embd_pdrop=0.0,
                     attn_pdrop=0.0)
    torch.manual_seed(seed)
    return GPT2LMHeadModel(cfg)


# name -> (smp config, options)
SCENARIOS = {
    "tp4-fp32": ({"tensor_parallel_degree": 4}, {}),
    "tp4-bf16-ckpt": ({"tensor_parallel_degree": 4, "bf16": True}, {"ckpt": True}),
    "tp4-fp32-memory": ({"tensor_parallel_degree": 4, "optimize": "memory"}, {}),
    "tp2pp2-fp32-delayed": ({"tensor_parallel_degree": 2, "pipeline_parallel_degree": 2,
                             "delayed_parameter_initialization": True}, {}),
    "tp2pp2-bf16-prescaled-ckpt": ({"tensor_parallel_degree": 2, "pipeline_parallel_degree": 2, "bf16": True,
                                    "prescaled_batch": True}, {"ckpt": True}),
    "tp2pp2-fp16": ({"tensor_parallel_degree": 2, "pipeline_parallel_degree": 2, "fp16": True}, {"loss_scale": 64.0}),
    "tp2rdp2-fp32-optshard": ({"tensor_parallel_degree": 2, "shard_optimizer_state": True}, {"adam": True}),
    "pp2dp2-bf16": ({"pipeline_parallel_degree": 2, "bf16": True}, {}),
    "sdp4-fp32": ({"sharded_data_parallel_degree": 4, "sdp_param_persistence_threshold": 100}, {"adam": True}),
    "sdp4-bf16-ckpt": ({"sharded_data_parallel_degree": 4, "sdp_param_persistence_threshold": 100, "bf16": True},
                       {"adam": True, "ckpt": True}),
}
GPU_ONLY = {
    "sdp4-bf16-ckpt-offload": ({"sharded_data_parallel_degree": 4, "sdp_param_persistence_threshold": 100, "bf16": True,
                                "offload_activations": True}, {"adam": True, "ckpt": True}),
    "tp4-bf16-ckpt-offload": ({"tensor_parallel_degree": 4, "bf16": True, "offload_activations": True}, {"ckpt": True}),
}


def _worker(rank, world, name, cuda):
    import os

    import smdistributed_modelparallel_b200 as smp

    conf, opts = dict({**SCENARIOS, **GPU_ONLY}[name][0]), {**SCENARIOS, **GPU_ONLY}[name][1]
    os.environ["SMP_SDP_UNIT_ELEMENTS"] = "20000"
    conf.update({"microbatches": 2, "ddp": True})
    if not cuda:
        conf["ddp_dist_backend"] = "gloo"
    smp.init(conf)
    tp = conf.get("tensor_parallel_degree", 1)
    half = torch.bfloat16 if conf.get("bf16") else (torch.float16 if conf.get("fp16") else None)
    dense = _hf()
    hf_sd = {k: v.clone() for k, v in dense.state_dict().items()}
    with smp.model_creation(tensor_parallelism=tp > 1, dtype=half):   # delayed initialisation comes from the config
        net = _hf(seed=1)                                  # other values: everything must come from load_state_dict
    if opts.get("ckpt"):
        smp.set_activation_checkpointing(net.transformer.h if hasattr(net.transformer, "h") else net.transformer,
                                         strategy="each")
    model = smp.DistributedModel(net)
    lr = 1e-2 if opts.get("adam") else 0.2
    make = (lambda p: torch.optim.Adam(p, lr=lr)) if opts.get("adam") else (lambda p: torch.optim.SGD(p, lr=lr))
    okw = {"static_loss_scale": opts["loss_scale"]} if "loss_scale" in opts else {}
    opt = smp.DistributedOptimizer(make(model.parameters()), **okw)
    model.load_state_dict(hf_sd)
    dopt = make(dense.parameters())

    @smp.step
    def train_step(model, ids):
        loss = model(input_ids=ids, labels=ids).loss
        model.backward(loss)
        return loss

    g = torch.Generator().manual_seed(17)
    dp = world // (tp * conf.get("pipeline_parallel_degree", 1))
    groups = dp * (1 if conf.get("prescaled_batch") else tp)      # independent batches per step
    per = 4
    # one batch, repeated: the loss has to fall from step to step
    all_ids = torch.randint(0, VOCAB, (1, groups * per, SEQ), generator=g).expand(STEPS, -1, -1)
    mine = smp.dp_rank() if not conf.get("prescaled_batch") else smp.rdp_rank()
    dev = torch.device("cuda", smp.local_rank()) if cuda else torch.device("cpu")
    losses, dlosses = [], []
    for it in range(STEPS):
        ids = all_ids[it, mine * per:(mine + 1) * per].to(dev)
        opt.zero_grad()
        out = train_step(model, ids)
        opt.step()
        # this rank's loss is the mean over its own sequences; the mean over the data-parallel ranks is the
        # dense model's loss on the union batch
        own = torch.tensor([out.reduce_mean().item()], dtype=torch.float64)
        gathered = smp.state.comm.allgather(own.item(), smp.CommGroup.WORLD)
        losses.append(sum(gathered) / len(gathered))
        dopt.zero_grad()
        dl = dense(input_ids=all_ids[it], labels=all_ids[it]).loss
        dl.backward()
        dopt.step()
        dlosses.append(dl.item())
    to_hf = smp.state.tp_registry.active_translate_functions[0][0] if smp.state.tp_registry.active_translate_functions \
        else (lambda sd: sd)
    back = to_hf(model.state_dict())
    dsd = dense.state_dict()
    num = den_a = den_b = 0.0
    worst = 0.0
    missing = [k for k in dsd if k not in back and not k.endswith(".attn.bias") and not k.endswith("masked_bias")]
    for k, v in dsd.items():
        if k not in back or not v.is_floating_point():
            continue
        a = back[k].float().cpu() - hf_sd[k]                      # the update smp applied
        b = v - hf_sd[k]                                          # the update the dense run applied
        worst = max(worst, (a - b).abs().max().item())
        num += (a * b).sum().item()
        den_a += a.pow(2).sum().item()
        den_b += b.pow(2).sum().item()
    cos = num / max((den_a * den_b) ** 0.5, 1e-30)
    return {"losses": losses, "dense": dlosses, "worst": worst, "cos": cos, "missing": missing,
            "replaced": type(model.module).__name__, "scale": (den_a / max(den_b, 1e-30)) ** 0.5}


def _check(name, res):
    conf = {**SCENARIOS, **GPU_ONLY}[name][0]
    sixteen = bool(conf.get("bf16") or conf.get("fp16"))
    r = res[0]
    assert not r["missing"], r["missing"][:5]
    if conf.get("tensor_parallel_degree", 1) > 1:
        assert r["replaced"] == "DistributedTransformerLMHead"
    for a, b in zip(r["losses"], r["dense"]):
        assert abs(a - b) < (6e-2 if sixteen else 2e-4), (name, r["losses"], r["dense"])
    assert r["dense"][-1] < r["dense"][0]
    # the update applied to every parameter: exact in fp32, same direction and size in 16-bit
    if sixteen:
        assert r["cos"] > 0.97 and 0.9 < r["scale"] < 1.1, (name, r["cos"], r["scale"])
    elif {**SCENARIOS, **GPU_ONLY}[name][1].get("adam"):
        # Adam divides by sqrt(v) + eps: elements whose gradient is at rounding level move by +-lr on noise, so
        # the element-wise bound is loose here; the losses of the later steps (above) pin the update
        assert r["cos"] > 0.995 and 0.97 < r["scale"] < 1.03, (name, r["worst"], r["cos"], r["scale"])
    else:
        assert r["worst"] < 2e-4 and r["cos"] > 0.9999, (name, r["worst"], r["cos"])
    for other in res[1:]:
        assert other["losses"] == pytest.approx(r["losses"], abs=1e-9)


@pytest.mark.parametrize("name", list(SCENARIOS))
def test_hybrid_layouts_track_dense_on_gloo(name):
    _check(name, run_distributed(_worker, 4, name, False, timeout=600))


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(SCENARIOS) + list(GPU_ONLY))
def test_hybrid_layouts_track_dense_on_gpus(name):
    if torch.cuda.device_count() < 4:
        pytest.skip("needs 4 GPUs")
    import os
    if os.environ.get("SMP_RUN_HYBRID_GPU", "0") != "1":
        # the matrix has run on gloo only so far (this round's multi-GPU budget was spent before it existed); until a
        # 4-GPU log of it is under profiles/, a -m gpu run on a larger box does not meet it unannounced
        pytest.skip("set SMP_RUN_HYBRID_GPU=1 to run the hybrid matrix on GPUs")
    _check(name, run_distributed(_worker, 4, name, True, timeout=900, use_cuda=True))
