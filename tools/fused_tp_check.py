"""2..8-GPU check of the fused tensor-parallel kernels: numerics vs NCCL + fp32 matmul, then timing.
Launch: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/fused_tp_check.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import smdistributed_modelparallel_b200 as smp
from smdistributed_modelparallel_b200.parallel import fused_tp, tp_ops
from smdistributed_modelparallel_b200.ops.gemm import gemm

n = int(os.environ["WORLD_SIZE"])
os.environ.setdefault("SMP_LOG_LEVEL", "warning")
smp.init({"tensor_parallel_degree": n, "ddp": True})
r = smp.tp_rank()
dev = torch.device("cuda", smp.local_rank())
g = smp.get_tp_process_group()
m = fused_tp.manager()
bf = torch.bfloat16


def log(*a):
    if r == 0:
        print(*a, flush=True)


def rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


torch.manual_seed(100 + r)
ok = True
for (T_l, K, N) in [(128, 64, 64), (256, 512, 384), (1024, 1024, 2048), (2048, 4096, 3072)]:
    for it in range(3):          # repeated: exercises buffer reuse / epochs
        x = torch.randn(T_l, K, device=dev, dtype=bf)
        w = torch.randn(N, K, device=dev, dtype=bf) * 0.05
        X = tp_ops.gather_tokens(x, g)
        ref = X.float() @ w.float().t()
        y, full = m.ag_gemm(x, w)
        e1 = rel(y, ref)
        e1b = (full - X).abs().max().item()
        # reduce-scatter: every rank has a full-T input with its own K shard
        xs = torch.randn(n * T_l, K, device=dev, dtype=bf)
        part = (xs.float() @ w.float().t())
        tot = part.clone()
        dist.all_reduce(tot, group=g)
        ref_rs = tot[r * T_l:(r + 1) * T_l]
        y2 = m.gemm_rs(xs, w)
        e2 = rel(y2, ref_rs)
        ag = m.all_gather(x).clone()
        e3 = (ag - X).abs().max().item()
        torch.cuda.synchronize()
        bad = e1 > 2e-2 or e2 > 2e-2 or e1b != 0 or e3 != 0
        ok = ok and not bad
        log(f"T_l={T_l} K={K} N={N} it={it}: ag_gemm rel {e1:.2e} (gather exact {e1b == 0}) | gemm_rs rel {e2:.2e} | "
            f"all_gather exact {e3 == 0} {'BAD' if bad else 'ok'}")
flag = torch.tensor([0.0 if ok else 1.0], device=dev)
dist.all_reduce(flag)
log("NUMERICS", "PASS" if flag.item() == 0 else "FAIL")


def timeit(fn, iters=10, label="timed loop"):
    """ms per call, max over ranks; the loop is long enough (~0.25 s, same count on every rank) for the clock
    sampler, whose record (SM clock, throttle reasons during the loop) rank 0 prints next to the number."""
    from tools._clocks import calibrated_iters, clocks
    for _ in range(3):
        fn()
    n_it = torch.tensor([calibrated_iters(fn, floor=iters)], device=dev)
    dist.all_reduce(n_it, op=dist.ReduceOp.MIN)
    iters = int(n_it.item())
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    import contextlib
    with (clocks(label) if dist.get_rank() == 0 else contextlib.nullcontext()):
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


# GPT-J 6B shapes at micro-batch 4 x 2048 tokens per tp group
T, H = int(os.environ.get("TP_CHECK_T", 8192)), int(os.environ.get("TP_CHECK_H", 4096))
T_l = T // n
log(f"shapes: T={T} H={H} tp={n}")


def roof(t_gemm, nbytes):
    """Roofline of a fused GEMM+collective: the slower of the GEMM alone and its bytes over NVLink at 900 GB/s."""
    return max(t_gemm, nbytes / 900e9 * 1e3)


for name, K, N in [("qkv+mlp1 (AG->GEMM)", H, (3 * H + 4 * H) // n), ("qkv (AG->GEMM)", H, 3 * H // n)]:
    x = torch.randn(T_l, K, device=dev, dtype=bf); w = torch.randn(N, K, device=dev, dtype=bf) * 0.02
    t_f = timeit(lambda: m.ag_gemm(x, w))
    t_n = timeit(lambda: gemm(tp_ops.gather_tokens(x, g), w))
    t_g = timeit(lambda: gemm(torch.empty(T, K, device=dev, dtype=bf), w))
    t_c = timeit(lambda: tp_ops.gather_tokens(x, g))
    fl = 2.0 * T * N * K
    rf = roof(t_g, T * K * 2 * (n - 1) / n)
    log(f"{name}: fused {t_f:.3f} ms ({fl/t_f/1e9:.0f} TF) | nccl AG + gemm {t_n:.3f} ms | gemm alone {t_g:.3f} | AG alone {t_c:.3f} | roofline {rf:.3f} ms -> {100 * rf / t_f:.0f} %")
for name, K, N in [("mlp2 (GEMM->RS)", 4 * H // n, H), ("attn dense (GEMM->RS)", H // n, H)]:
    x = torch.randn(T, K, device=dev, dtype=bf); w = torch.randn(N, K, device=dev, dtype=bf) * 0.02
    t_f = timeit(lambda: m.gemm_rs(x, w))
    t_n = timeit(lambda: tp_ops.scatter_sum_tokens(gemm(x, w), g))
    t_g = timeit(lambda: gemm(x, w))
    y = gemm(x, w)
    t_c = timeit(lambda: tp_ops.scatter_sum_tokens(y, g))
    fl = 2.0 * T * N * K
    rf = roof(t_g, T * N * 2 * (n - 1) / n)
    log(f"{name}: fused {t_f:.3f} ms ({fl/t_f/1e9:.0f} TF) | gemm + nccl RS {t_n:.3f} ms | gemm alone {t_g:.3f} | RS alone {t_c:.3f} | roofline {rf:.3f} ms -> {100 * rf / t_f:.0f} %")
dist.barrier()
