"""Quick GEMM bring-up probe: small problems first, error maps on mismatch, then a timing sweep."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from smdistributed_modelparallel_b200.ops.gemm import gemm, stats

torch.manual_seed(0)
dev = "cuda"


def check(M, N, K, layout, dtype=torch.bfloat16):
    A = torch.randn(M, K, device=dev, dtype=dtype)
    B = torch.randn(N, K, device=dev, dtype=dtype)
    ref = A.float() @ B.float().t()
    if layout == "nt":
        out = gemm(A, B)
    elif layout == "nn":
        out = gemm(A, B.t().contiguous(), b_mn_major=True)
    else:
        out = gemm(A.t().contiguous(), B.t().contiguous(), a_mn_major=True, b_mn_major=True)
    torch.cuda.synchronize()
    err = ((out.float() - ref).norm() / ref.norm()).item()
    print(f"gemm {layout} {M}x{N}x{K} {dtype}: rel {err:.3e} native={stats['native']} fb={stats['fallback']}", flush=True)
    if err > 1e-2:
        d = (out.float() - ref).abs()
        bm, bn = min(M, 256) // 32, min(N, 256) // 32
        tile = d[: bm * 32, : bn * 32].reshape(bm, 32, bn, 32).amax(dim=(1, 3))
        print("max-abs error per 32x32 block (first 256x256):")
        print((tile > 0.5).int())
        print("out[0,:8]", out[0, :8].float().tolist())
        print("ref[0,:8]", ref[0, :8].tolist())
    return err


for lay in ("nt", "nn", "tn"):
    for shp in ((128, 64, 64), (128, 256, 64), (128, 256, 128), (256, 512, 256), (384, 640, 192)):
        check(*shp, lay)

# timing
print("---- timing (bf16, nt) ----")
for (M, N, K) in ((4096, 4096, 4096), (8192, 8192, 8192), (16384, 4096, 4096), (16384, 16384, 4096), (16384, 4096, 16384)):
    A = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
    B = torch.randn(N, K, device=dev, dtype=torch.bfloat16)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    for fn, name in ((lambda: gemm(A, B, out=out), "tcgen05"), (lambda: torch.matmul(A, B.t(), out=out), "cublas")):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"{name} {M}x{N}x{K}: {ms:.3f} ms  {2*M*N*K/ms/1e9:.1f} TFLOP/s", flush=True)
