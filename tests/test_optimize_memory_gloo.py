"""optimize='memory' on 2 CPU/gloo ranks: activations enter a stack of transformer layers through an all-to-all
([local batch, S, H] -> [full batch, S, H/t]), every GEMM multiplies this rank's H/t (or heads / intermediate) slice and
is reduce-scattered over its output features, LayerNorms are hidden-sharded, and the last layer converts back
(reference torch/nn/transformer.py:1429-1454 for the layer-level all-to-alls, :1243-1262 / :1869-1905 for the
reduce-scatter projections).  Outputs, input gradients and the gathered state dict are compared with the same layers
built densely."""

import pytest
import torch

from dist_utils import run_distributed

KW = dict(num_attention_heads=5, attention_head_size=4, hidden_size=22, intermediate_size=30,
          attention_dropout_prob=0.0, hidden_dropout_prob=0.0, activation="gelu", pre_layernorm=True,
          post_layernorm=True, causal_mask_size=8, use_normal_initialization=True)
XKW = dict(KW, add_cross_attention=True, num_attention_heads=4, hidden_size=16, intermediate_size=24)


def _worker(rank, world):
    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.nn.transformer import DistributedTransformer

    smp.init({"tensor_parallel_degree": world, "ddp": True, "ddp_dist_backend": "gloo", "optimize": "memory"})
    out = {}
    data = torch.Generator().manual_seed(77)
    torch.manual_seed(3)
    with smp.tensor_parallelism(enabled=True):
        stack = DistributedTransformer(num_layers=2, **KW)
    first, last = stack.seq_layers[0], stack.seq_layers[-1]
    out["flags"] = (first.input_layer, first.output_layer, last.input_layer, last.output_layer)
    out["shapes"] = {n: tuple(p.shape) for n, p in first.named_parameters()}
    h_all = torch.randn(4, 8, 22, generator=data)
    h = h_all[rank * 2:(rank + 1) * 2].clone().requires_grad_(True)
    y = stack((h, None))[0]
    y.square().sum().backward()
    out["y"], out["dx"], out["h_all"] = y.detach(), h.grad.clone(), h_all
    model = smp.DistributedModel(stack)
    out["sd"] = {k: v.clone() for k, v in model.state_dict().items()}
    # ---- a decoder layer with cross attention, entry and exit in one layer
    smp.reset()
    smp.init({"tensor_parallel_degree": world, "ddp": True, "ddp_dist_backend": "gloo", "optimize": "memory"})
    torch.manual_seed(4)
    with smp.tensor_parallelism(enabled=True):
        xl = smp.nn.DistributedTransformerLayer(**XKW)
    d_all = torch.randn(4, 8, 16, generator=data)
    e_all = torch.randn(4, 6, 16, generator=data)
    dh = d_all[rank * 2:(rank + 1) * 2].clone().requires_grad_(True)
    enc = e_all[rank * 2:(rank + 1) * 2].clone().requires_grad_(True)
    yx = xl((dh, None, enc, None))[0]
    yx.square().sum().backward()
    out["x_y"], out["x_dh"], out["x_denc"], out["d_all"], out["e_all"] = yx.detach(), dh.grad.clone(), enc.grad.clone(), \
        d_all, e_all
    xm = smp.DistributedModel(xl)
    out["x_sd"] = {k: v.clone() for k, v in xm.state_dict().items()}
    # ---- GPT-J style blocks are refused
    try:
        with smp.tensor_parallelism(enabled=True):
            smp.nn.DistributedTransformerLayer(**dict(KW, parallel_attn_output=True, post_layernorm=False))
        out["refused"] = False
    except Exception as e:      # noqa: BLE001
        out["refused"] = type(e).__name__
    return out


def test_memory_layout_matches_dense():
    import smdistributed_modelparallel_b200  # noqa: F401
    from smdistributed_modelparallel_b200.nn.transformer import DistributedTransformer, DistributedTransformerLayer

    res = run_distributed(_worker, 2, timeout=300)
    r0, r1 = res
    assert r0["flags"] == (True, False, False, True)
    # rank 0 holds 11 of the 22 hidden features, 3 of the 5 heads, 15 of 30 intermediate features
    sh0, sh1 = r0["shapes"], r1["shapes"]
    assert sh0["attention.qkv_weight"] == (3 * 20, 11) and sh0["attention.dense_weight"] == (22, 12)
    assert sh1["attention.dense_weight"] == (22, 8) and sh0["attention.qkv_bias"] == (36,)
    assert sh0["output.dense1_weight"] == (30, 11) and sh0["output.dense2_weight"] == (22, 15)
    assert sh0["output.dense2_bias"] == (11,) and sh0["attention.dense_bias"] == (11,)
    dense = DistributedTransformer(num_layers=2, **KW)
    missing = dense.load_state_dict(r0["sd"], strict=False)
    assert not missing.missing_keys and not missing.unexpected_keys, missing
    ha = r0["h_all"].clone().requires_grad_(True)
    yd = dense((ha, None))[0]
    yd.square().sum().backward()
    for r, rr in enumerate(res):
        assert rr["y"].shape == (2, 8, 22)
        assert torch.allclose(rr["y"], yd[r * 2:(r + 1) * 2].detach(), atol=3e-5), (rr["y"] - yd[r * 2:(r + 1) * 2]).abs().max()
        assert torch.allclose(rr["dx"], ha.grad[r * 2:(r + 1) * 2], atol=3e-5)
    xdense = DistributedTransformerLayer(**XKW)
    missing = xdense.load_state_dict(r0["x_sd"], strict=False)
    assert not missing.missing_keys and not missing.unexpected_keys, missing
    da = r0["d_all"].clone().requires_grad_(True)
    ea = r0["e_all"].clone().requires_grad_(True)
    yd = xdense((da, None, ea, None))[0]
    yd.square().sum().backward()
    for r, rr in enumerate(res):
        assert torch.allclose(rr["x_y"], yd[r * 2:(r + 1) * 2].detach(), atol=3e-5)
        assert torch.allclose(rr["x_dh"], da.grad[r * 2:(r + 1) * 2], atol=3e-5)
        assert torch.allclose(rr["x_denc"], ea.grad[r * 2:(r + 1) * 2], atol=3e-5)
    assert r0["refused"] == "DistTransformerConfigError"
