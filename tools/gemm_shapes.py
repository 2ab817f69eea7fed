"""Per-shape GEMM timing for the GPT-J 6B step (micro-batch 4 x 2048 tokens): tcgen05 kernel vs cuBLAS."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from smdistributed_modelparallel_b200.ops.gemm import gemm

dev = "cuda"
T, H, I, V, S, d, BH = 8192, 4096, 16384, 50400, 2048, 256, 64
bf = torch.bfloat16


def timeit(fn, iters=8, label="timed loop"):
    """ms per call over a ~0.25 s loop, with the clock record of that loop (SM clock, throttle reasons) printed."""
    from tools._clocks import calibrated_iters, clocks
    for _ in range(3):
        fn()
    iters = calibrated_iters(fn, floor=iters)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    with clocks(label):
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def case(name, M, N, K, batch=1, layout="nt", out_fp32_acc=False, count=1):
    shp = (lambda r, c: (batch, r, c) if batch > 1 else (r, c))
    A = torch.randn(shp(M, K), device=dev, dtype=bf) if layout != "tn" else torch.randn(shp(K, M), device=dev, dtype=bf)
    B = torch.randn(shp(N, K), device=dev, dtype=bf) if layout == "nt" else torch.randn(shp(K, N), device=dev, dtype=bf)
    flops = 2.0 * M * N * K * batch
    if out_fp32_acc:
        out = torch.zeros(shp(M, N), device=dev, dtype=torch.float32)
        mine = lambda: gemm(A, B, a_mn_major=(layout == "tn"), b_mn_major=(layout != "nt"), out=out, residual=out, beta=1.0)
        ref_out = torch.zeros(shp(M, N), device=dev, dtype=torch.float32)
        def ref():
            ref_out.add_(torch.matmul(A.transpose(-1, -2), B))
    else:
        out = torch.empty(shp(M, N), device=dev, dtype=bf)
        mine = lambda: gemm(A, B, a_mn_major=(layout == "tn"), b_mn_major=(layout != "nt"), out=out)
        if layout == "nt":
            ref = lambda: torch.matmul(A, B.transpose(-1, -2))
        elif layout == "nn":
            ref = lambda: torch.matmul(A, B)
        else:
            ref = lambda: torch.matmul(A.transpose(-1, -2), B)
    tm, tr = timeit(mine), timeit(ref)
    print(f"{name:28s} {layout} M={M:6d} N={N:6d} K={K:6d} b={batch:3d}  mine {tm:7.3f} ms {flops/tm/1e9:7.0f} TF | "
          f"cublas {tr:7.3f} ms {flops/tr/1e9:7.0f} TF | ratio {tr/tm:5.2f} | x{count}/layer -> {tm*count:6.2f} ms", flush=True)
    return tm * count, tr * count


tot_m = tot_r = 0.0
for args in [
    ("qkv fwd", T, 3 * H, H, 1, "nt", False, 1), ("attn-dense fwd", T, H, H, 1, "nt", False, 1),
    ("mlp1 fwd", T, I, H, 1, "nt", False, 1), ("mlp2 fwd", T, H, I, 1, "nt", False, 1),
    ("qkv dgrad", T, H, 3 * H, 1, "nn", False, 1), ("attn-dense dgrad", T, H, H, 1, "nn", False, 1),
    ("mlp1 dgrad", T, H, I, 1, "nn", False, 1), ("mlp2 dgrad", T, I, H, 1, "nn", False, 1),
    ("qkv wgrad", 3 * H, H, T, 1, "tn", True, 1), ("attn-dense wgrad", H, H, T, 1, "tn", True, 1),
    ("mlp1 wgrad", I, H, T, 1, "tn", True, 1), ("mlp2 wgrad", H, I, T, 1, "tn", True, 1),
    ("QK^T", S, S, d, BH, "nt", False, 1), ("PV", S, d, S, BH, "nn", False, 1),
    ("dP = dO V^T", S, S, d, BH, "nt", False, 1), ("dV = P^T dO", S, d, S, BH, "tn", False, 1),
    ("dQ = dS K", S, d, S, BH, "nn", False, 1), ("dK = dS^T Q", S, d, S, BH, "tn", False, 1),
]:
    a, b = case(*args)
    tot_m += a
    tot_r += b
print(f"per-layer GEMM total: mine {tot_m:.2f} ms, cublas {tot_r:.2f} ms")
case("lm head fwd", T, V, H, 1, "nt")
case("lm head dgrad", T, H, V, 1, "nn")
case("lm head wgrad", V, H, T, 1, "tn", True)
