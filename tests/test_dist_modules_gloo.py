"""Tensor-parallel building blocks on 2 CPU/gloo ranks against dense PyTorch (reference
test/torch/mpi_4ps/test_dist_modules.py:81-130): DistributedLinear (input-partitioned, uneven split), vocab- and
hidden-parallel DistributedEmbedding, hidden-sharded DistributedLayerNorm, and a transformer layer whose head
count does not divide the tensor-parallel degree - outputs and input gradients."""

import torch

from dist_utils import run_distributed


def _worker(rank, world):
    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.checkpoint_utils import gather_full_state_dict  # noqa: F401

    smp.init({"tensor_parallel_degree": world, "ddp": True, "ddp_dist_backend": "gloo"})
    out = {}
    torch.manual_seed(0)
    data = torch.Generator().manual_seed(123)     # inputs come from their own stream: parameter shapes differ per rank
    # ---- DistributedLinear, 10 input features over 2 ranks x batch 3 per rank
    lin = smp.nn.DistributedLinear(11, 6)                   # 11 = 6 + 5: uneven split
    x_all = torch.randn(2 * 3, 11, generator=data)
    x = x_all[rank * 3:(rank + 1) * 3].clone().requires_grad_(True)
    y = lin(x)
    y.sum().backward()
    out["lin_w"], out["lin_b"] = lin.weight.detach().clone(), (lin.bias.detach().clone() if lin.bias is not None else None)
    out["lin_y"], out["lin_dx"], out["x_all"] = y.detach(), x.grad.clone(), x_all
    # ---- vocab-parallel embedding
    emb = smp.nn.DistributedEmbedding(13, 8, vocab_parallel=True)
    ids_all = torch.randint(0, 13, (4, 5), generator=data)
    ids = ids_all[rank * 2:(rank + 1) * 2]
    e = emb(ids)
    e.sum().backward()
    out["emb_w"], out["emb_range"] = emb.weight.detach().clone(), (emb.vocab_start_idx, emb.vocab_end_idx)
    out["emb_y"], out["ids_all"], out["emb_gw"] = e.detach(), ids_all, emb.weight.grad.clone()
    # ---- hidden-parallel embedding
    emb2 = smp.nn.DistributedEmbedding(13, 8, vocab_parallel=False)
    e2 = emb2(ids)
    out["emb2_w"], out["emb2_y"] = emb2.weight.detach().clone(), e2.detach()
    # ---- hidden-sharded LayerNorm: every rank holds H/t channels of the same rows
    ln = smp.nn.DistributedLayerNorm(10)
    xl_full = torch.randn(4, 10, generator=data)
    lo = 5 * rank
    xl = xl_full[:, lo:lo + 5].clone().requires_grad_(True)
    yl = ln(xl)
    (yl * torch.arange(5.0)).sum().backward()
    out["ln_y"], out["ln_dx"], out["ln_full"] = yl.detach(), xl.grad.clone(), xl_full
    # ---- transformer layer with 5 heads on 2 ranks (3 + 2)
    torch.manual_seed(1)
    with smp.tensor_parallelism(enabled=True):
        layer = smp.nn.DistributedTransformerLayer(num_attention_heads=5, attention_head_size=4, hidden_size=20,
                                                   intermediate_size=30, attention_dropout_prob=0.0,
                                                   hidden_dropout_prob=0.0, activation="gelu", pre_layernorm=True,
                                                   post_layernorm=False, causal_mask_size=8,
                                                   use_normal_initialization=True)
    h_all = torch.randn(4, 8, 20, generator=data)
    h = h_all[rank * 2:(rank + 1) * 2].clone().requires_grad_(True)
    yo, _ = layer((h, None))
    yo.square().sum().backward()
    out["layer_y"], out["layer_dx"], out["h_all"] = yo.detach(), h.grad.clone(), h_all
    out["layer_heads"] = layer.attention.local_attention_heads
    model = smp.DistributedModel(layer)
    out["layer_sd"] = {k: v.clone() for k, v in model.state_dict().items()}     # gathered over tp
    # ---- decoder layer with cross attention over encoder states of a different length, post-LayerNorm
    smp.reset()
    torch.manual_seed(2)
    with smp.tensor_parallelism(enabled=True):
        xl = smp.nn.DistributedTransformerLayer(num_attention_heads=4, attention_head_size=4, hidden_size=16,
                                                intermediate_size=24, attention_dropout_prob=0.0,
                                                hidden_dropout_prob=0.0, activation="relu", pre_layernorm=False,
                                                post_layernorm=True, causal_mask_size=8, add_cross_attention=True,
                                                use_normal_initialization=True)
    d_all = torch.randn(4, 8, 16, generator=data)
    enc_all = torch.randn(4, 6, 16, generator=data)
    dh = d_all[rank * 2:(rank + 1) * 2].clone().requires_grad_(True)
    enc = enc_all[rank * 2:(rank + 1) * 2].clone().requires_grad_(True)
    yx = xl((dh, None, enc, None))[0]
    yx.square().sum().backward()
    out["x_y"], out["x_dh"], out["x_denc"] = yx.detach(), dh.grad.clone(), enc.grad.clone()
    out["d_all"], out["enc_all"] = d_all, enc_all
    xm = smp.DistributedModel(xl)
    out["x_sd"] = {k: v.clone() for k, v in xm.state_dict().items()}
    return out


def test_tp2_modules_match_dense():
    import torch.nn.functional as F

    res = run_distributed(_worker, 2, timeout=300)
    r0, r1 = res
    r0["ids_all"] = r0["ids_all"].long()
    # DistributedLinear: W = [W0 | W1] over input features, bias lives on tp rank 0
    W = torch.cat([r0["lin_w"], r1["lin_w"]], dim=1)
    assert W.shape == (6, 11) and r0["lin_w"].shape[1] == 6 and r1["lin_w"].shape[1] == 5
    assert r0["lin_b"] is not None and r1["lin_b"] is None
    xa = r0["x_all"].clone().requires_grad_(True)
    ya = xa @ W.t() + r0["lin_b"]
    ya.sum().backward()
    for r, rr in enumerate(res):
        assert torch.allclose(rr["lin_y"], ya[r * 3:(r + 1) * 3].detach(), atol=1e-5)
        assert torch.allclose(rr["lin_dx"], xa.grad[r * 3:(r + 1) * 3], atol=1e-5)
    # vocab-parallel embedding: rows 0..6 on rank 0, 7..12 on rank 1
    assert r0["emb_range"] == (0, 7) and r1["emb_range"] == (7, 13)
    E = torch.cat([r0["emb_w"], r1["emb_w"]], dim=0)
    for r, rr in enumerate(res):
        assert torch.allclose(rr["emb_y"], E[r0["ids_all"][r * 2:(r + 1) * 2]], atol=1e-6)
    counts = torch.bincount(r0["ids_all"].reshape(-1), minlength=13).float()
    assert torch.allclose(r0["emb_gw"][:, 0], counts[:7]) and torch.allclose(r1["emb_gw"][:, 0], counts[7:])
    # hidden-parallel embedding: columns split
    E2 = torch.cat([r0["emb2_w"], r1["emb2_w"]], dim=1)
    for r, rr in enumerate(res):
        assert torch.allclose(rr["emb2_y"], E2[r0["ids_all"][r * 2:(r + 1) * 2]], atol=1e-6)
    # hidden-sharded LayerNorm == LayerNorm over the concatenated channels
    xf = r0["ln_full"].clone().requires_grad_(True)
    yf = F.layer_norm(xf, (10,))
    (yf * torch.arange(5.0).repeat(2)).sum().backward()
    for r, rr in enumerate(res):
        assert torch.allclose(rr["ln_y"], yf[:, 5 * r:5 * r + 5].detach(), atol=1e-5)
        assert torch.allclose(rr["ln_dx"], xf.grad[:, 5 * r:5 * r + 5], atol=1e-5)
    # transformer layer, 5 heads split 3 + 2: compare with the same layer built densely from the gathered weights
    assert (r0["layer_heads"], r1["layer_heads"]) == (3, 2)
    import smdistributed_modelparallel_b200  # noqa: F401
    from smdistributed_modelparallel_b200.nn.transformer import DistributedTransformerLayer

    dense = DistributedTransformerLayer(num_attention_heads=5, attention_head_size=4, hidden_size=20,
                                        intermediate_size=30, attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
                                        activation="gelu", pre_layernorm=True, post_layernorm=False,
                                        causal_mask_size=8, use_normal_initialization=True)
    missing = dense.load_state_dict(r0["layer_sd"], strict=False)
    assert not missing.missing_keys, missing
    ha = r0["h_all"].clone().requires_grad_(True)
    yd, _ = dense((ha, None))
    yd.square().sum().backward()
    for r, rr in enumerate(res):
        assert torch.allclose(rr["layer_y"], yd[r * 2:(r + 1) * 2].detach(), atol=2e-5)
        assert torch.allclose(rr["layer_dx"], ha.grad[r * 2:(r + 1) * 2], atol=2e-5)
    # cross-attention decoder layer
    xdense = DistributedTransformerLayer(num_attention_heads=4, attention_head_size=4, hidden_size=16,
                                         intermediate_size=24, attention_dropout_prob=0.0, hidden_dropout_prob=0.0,
                                         activation="relu", pre_layernorm=False, post_layernorm=True, causal_mask_size=8,
                                         add_cross_attention=True, use_normal_initialization=True)
    missing = xdense.load_state_dict(r0["x_sd"], strict=False)
    assert not missing.missing_keys, missing
    da = r0["d_all"].clone().requires_grad_(True)
    ea = r0["enc_all"].clone().requires_grad_(True)
    yd = xdense((da, None, ea, None))[0]
    yd.square().sum().backward()
    for r, rr in enumerate(res):
        assert torch.allclose(rr["x_y"], yd[r * 2:(r + 1) * 2].detach(), atol=2e-5)
        assert torch.allclose(rr["x_dh"], da.grad[r * 2:(r + 1) * 2], atol=2e-5)
        assert torch.allclose(rr["x_denc"], ea.grad[r * 2:(r + 1) * 2], atol=2e-5)
