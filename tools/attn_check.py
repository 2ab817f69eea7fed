"""Numerics + timing check of the flash attention kernels against an fp32 PyTorch restatement (run on a B200).

    python tools/attn_check.py [--bwd] [--time]
"""
import argparse
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from smdistributed_modelparallel_b200.ops import _native as N  # noqa: E402
from tools._clocks import clocks  # noqa: E402


def ref_attention(q, k, v, scale, causal, window, mask):
    qf, kf, vf = q.float(), k.float(), v.float()
    s = torch.einsum("bhqd,bhkd->bhqk", qf, kf) * scale
    Sq, Sk = q.shape[2], k.shape[2]
    i = torch.arange(Sq, device=q.device)[:, None] + (Sk - Sq)
    j = torch.arange(Sk, device=q.device)[None, :]
    dead = torch.zeros(Sq, Sk, dtype=torch.bool, device=q.device)
    if causal:
        dead |= j > i
    if window:
        dead |= j <= i - window
    s = s.masked_fill(dead[None, None], float("-inf"))
    if mask is not None:
        s = s.masked_fill(mask.bool(), float("-inf"))
    p = torch.softmax(s, dim=-1)
    p = torch.nan_to_num(p, nan=0.0)
    return torch.einsum("bhqk,bhkd->bhqd", p, vf), torch.logsumexp(s, dim=-1)


def run_fwd(q, k, v, scale, causal=True, window=0, mask=None, p_drop=0.0, seed=0, offset=0, out_layout="bhsd"):
    B, H, Sq, d = q.shape
    Sk = k.shape[2]
    if out_layout == "bhsd":
        o = torch.empty(B, H, Sq, d, dtype=q.dtype, device=q.device)
        osb, osh, oss = o.stride(0), o.stride(1), o.stride(2)
    else:
        o = torch.empty(B, Sq, H, d, dtype=q.dtype, device=q.device)
        osb, osh, oss = o.stride(0), o.stride(2), o.stride(1)
    lse = torch.empty(B * H, Sq, dtype=torch.float32, device=q.device)
    if mask is not None:
        msb, msh, msq = mask.stride(0), (mask.stride(1) if mask.shape[1] > 1 else 0), (mask.stride(2) if mask.shape[2] > 1 else 0)
    else:
        msb = msh = msq = 0
    N.call("smpk_attn_fwd", N.stream_ptr(q.device), N.dtype_code(q.dtype), B, H, Sq, Sk, d,
           q.data_ptr(), q.stride(0), q.stride(1), q.stride(2), k.data_ptr(), k.stride(0), k.stride(1), k.stride(2),
           v.data_ptr(), v.stride(0), v.stride(1), v.stride(2), o.data_ptr(), osb, osh, oss, lse.data_ptr(),
           float(scale), int(causal), int(window), N.ptr(mask), msb, msh, msq, float(p_drop), seed, offset)
    return o, lse


def run_bwd(q, k, v, o, do, lse, scale, causal=True, window=0, mask=None, p_drop=0.0, seed=0, offset=0):
    B, H, Sq, d = q.shape
    Sk = k.shape[2]
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    delta = torch.empty(B * H, Sq, dtype=torch.float32, device=q.device)
    if mask is not None:
        msb, msh, msq = mask.stride(0), (mask.stride(1) if mask.shape[1] > 1 else 0), (mask.stride(2) if mask.shape[2] > 1 else 0)
    else:
        msb = msh = msq = 0
    st = lambda t: (t.stride(0), t.stride(1), t.stride(2))
    N.call("smpk_attn_bwd", N.stream_ptr(q.device), N.dtype_code(q.dtype), B, H, Sq, Sk, d,
           q.data_ptr(), *st(q), k.data_ptr(), *st(k), v.data_ptr(), *st(v), o.data_ptr(), *st(o), do.data_ptr(), *st(do),
           lse.data_ptr(), delta.data_ptr(), dq.data_ptr(), *st(dq), dk.data_ptr(), *st(dk), dv.data_ptr(), *st(dv),
           float(scale), int(causal), int(window), N.ptr(mask), msb, msh, msq, float(p_drop), seed, offset)
    return dq, dk, dv


def ref_grads(q, k, v, do, scale, causal, window, mask, keep=None, p_drop=0.0):
    qf, kf, vf = (t.float().detach().requires_grad_(True) for t in (q, k, v))
    s = torch.einsum("bhqd,bhkd->bhqk", qf, kf) * scale
    Sq, Sk = q.shape[2], k.shape[2]
    i = torch.arange(Sq, device=q.device)[:, None] + (Sk - Sq)
    j = torch.arange(Sk, device=q.device)[None, :]
    dead = torch.zeros(Sq, Sk, dtype=torch.bool, device=q.device)
    if causal:
        dead |= j > i
    if window:
        dead |= j <= i - window
    s = s.masked_fill(dead[None, None], float("-inf"))
    if mask is not None:
        s = s.masked_fill(mask.bool(), float("-inf"))
    p = torch.nan_to_num(torch.softmax(s, dim=-1), nan=0.0)
    if keep is not None:
        p = p * keep / (1.0 - p_drop)
    out = torch.einsum("bhqk,bhkd->bhqd", p, vf)
    out.backward(do.float())
    return out.detach(), qf.grad, kf.grad, vf.grad


def check_bwd(dev):
    ok = True
    cases = [
        (1, 1, 128, 128, 64, False, 0, None, torch.bfloat16),
        (1, 1, 128, 128, 64, True, 0, None, torch.bfloat16),
        (2, 3, 256, 256, 64, True, 0, None, torch.bfloat16),
        (2, 2, 512, 512, 128, True, 0, None, torch.bfloat16),
        (2, 2, 512, 512, 96, True, 0, None, torch.bfloat16),
        (1, 2, 512, 512, 256, True, 0, None, torch.bfloat16),
        (1, 2, 300, 300, 64, True, 0, None, torch.float16),
        (2, 2, 200, 333, 64, False, 0, "pad", torch.bfloat16),
        (1, 2, 384, 384, 128, False, 0, "full", torch.bfloat16),
        (1, 2, 640, 640, 128, True, 256, None, torch.bfloat16),
        (1, 2, 1024, 1024, 256, True, 0, None, torch.bfloat16),
        (1, 1, 192, 192, 192, True, 0, None, torch.bfloat16),
    ]
    for (B, H, Sq, Sk, d, causal, window, mt, dt) in cases:
        q = torch.randn(B, H, Sq, d, device=dev, dtype=dt)
        k = torch.randn(B, H, Sk, d, device=dev, dtype=dt)
        v = torch.randn(B, H, Sk, d, device=dev, dtype=dt)
        do = torch.randn(B, H, Sq, d, device=dev, dtype=dt)
        mask = None
        if mt == "pad":
            mask = torch.zeros(B, 1, 1, Sk, dtype=torch.uint8, device=dev)
            mask[0, :, :, Sk - 50:] = 1
            mask[1, :, :, Sk - 7:] = 1
        elif mt == "full":
            mask = (torch.rand(B, 1, Sq, Sk, device=dev) < 0.3).to(torch.uint8)
        scale = 1.0 / math.sqrt(d)
        o, lse = run_fwd(q, k, v, scale, causal, window, mask)
        dq, dk, dv = run_bwd(q, k, v, o, do, lse, scale, causal, window, mask)
        torch.cuda.synchronize()
        _, rq, rk, rv = ref_grads(q, k, v, do, scale, causal, window, mask)
        errs = []
        for name, a, r in (("dq", dq, rq), ("dk", dk, rk), ("dv", dv, rv)):
            rel = (a.float() - r).abs().max().item() / (r.abs().max().item() + 1e-6)
            errs.append((name, rel))
        good = all(e < 2e-2 for _, e in errs)
        ok &= good
        print(f"bwd B{B} H{H} Sq{Sq} Sk{Sk} d{d} causal={causal} win={window} mask={mt} {dt}: " +
              " ".join(f"{n} rel={e:.3e}" for n, e in errs) + (" OK" if good else " FAIL"), flush=True)
    # dropout: recover the keep mask through V = I (O row = dropped probabilities), then compare everything
    for d in (64, 256):
        B, H, S, pdrop, seed, offset = 1, 2, d, 0.2, 1234, 77
        dt = torch.bfloat16
        q = torch.randn(B, H, S, d, device=dev, dtype=dt)
        k = torch.randn(B, H, S, d, device=dev, dtype=dt)
        eye = torch.eye(S, device=dev, dtype=dt)[None, None].expand(B, H, S, d).contiguous()
        scale = 1.0 / math.sqrt(d)
        o_eye, _ = run_fwd(q, k, eye, scale, True, 0, None, pdrop, seed, offset)
        tri = torch.ones(S, S, device=dev, dtype=torch.bool).tril()
        keep = (o_eye != 0) | ~tri[None, None]
        frac = 1.0 - keep[:, :, tri].float().mean().item()
        v = torch.randn(B, H, S, d, device=dev, dtype=dt)
        do = torch.randn(B, H, S, d, device=dev, dtype=dt)
        o, lse = run_fwd(q, k, v, scale, True, 0, None, pdrop, seed, offset)
        o2, _ = run_fwd(q, k, v, scale, True, 0, None, pdrop, seed, offset)
        dq, dk, dv = run_bwd(q, k, v, o, do, lse, scale, True, 0, None, pdrop, seed, offset)
        torch.cuda.synchronize()
        ro, rq, rk, rv = ref_grads(q, k, v, do, scale, True, 0, None, keep.float(), pdrop)
        errs = [("o", (o.float() - ro).abs().max().item() / (ro.abs().max().item() + 1e-6))]
        for name, a, r in (("dq", dq, rq), ("dk", dk, rk), ("dv", dv, rv)):
            errs.append((name, (a.float() - r).abs().max().item() / (r.abs().max().item() + 1e-6)))
        good = all(e < 2e-2 for _, e in errs) and abs(frac - pdrop) < 0.03 and torch.equal(o, o2)
        ok &= good
        print(f"dropout d{d}: dropped fraction {frac:.3f} (p={pdrop}) deterministic={torch.equal(o, o2)} " +
              " ".join(f"{n} rel={e:.3e}" for n, e in errs) + (" OK" if good else " FAIL"), flush=True)
    return ok


def time_bwd(dev):
    for (name, B, H, S, d) in [("gptj", 4, 16, 2048, 256), ("neox20b", 2, 64, 2048, 96), ("gpt2xl", 4, 25, 1024, 64),
                               ("d128", 4, 32, 2048, 128)]:
        q = torch.randn(B, H, S, d, device=dev, dtype=torch.bfloat16)
        k, v, do = torch.randn_like(q), torch.randn_like(q), torch.randn_like(q)
        scale = 1.0 / math.sqrt(d)
        o, lse = run_fwd(q, k, v, scale, True)
        for _ in range(3):
            run_bwd(q, k, v, o, do, lse, scale, True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 400                       # ~0.4 s per shape: long enough for the clock sampler to see the load
        with clocks(f"bwd {name}"):
            e0.record()
            for _ in range(iters):
                run_bwd(q, k, v, o, do, lse, scale, True)
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        flops = 10.0 * B * H * S * S * d / 2
        print(f"time bwd {name} B{B} H{H} S{S} d{d} causal: {ms:.3f} ms  {flops / ms / 1e9:.0f} TFLOP/s (5 ideal causal products)", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--time", action="store_true")
    ap.add_argument("--bwd", action="store_true")
    args = ap.parse_args()
    torch.manual_seed(0)
    dev = torch.device("cuda")
    N.load()
    ok = True
    cases = [
        # B, H, Sq, Sk, d, causal, window, masktype, dtype
        (1, 1, 128, 128, 64, False, 0, None, torch.bfloat16),
        (1, 1, 128, 128, 64, True, 0, None, torch.bfloat16),
        (2, 3, 256, 256, 64, True, 0, None, torch.bfloat16),
        (2, 2, 512, 512, 128, True, 0, None, torch.bfloat16),
        (2, 2, 512, 512, 96, True, 0, None, torch.bfloat16),
        (1, 2, 1024, 1024, 256, True, 0, None, torch.bfloat16),
        (1, 2, 300, 300, 64, True, 0, None, torch.float16),
        (2, 2, 200, 333, 64, False, 0, "pad", torch.bfloat16),
        (1, 2, 384, 384, 128, False, 0, "full", torch.bfloat16),
        (1, 2, 640, 640, 128, True, 256, None, torch.bfloat16),
        (1, 4, 2048, 2048, 256, True, 0, None, torch.bfloat16),
    ]
    for (B, H, Sq, Sk, d, causal, window, mt, dt) in cases:
        q = torch.randn(B, H, Sq, d, device=dev, dtype=dt)
        k = torch.randn(B, H, Sk, d, device=dev, dtype=dt)
        v = torch.randn(B, H, Sk, d, device=dev, dtype=dt)
        mask = None
        if mt == "pad":
            mask = torch.zeros(B, 1, 1, Sk, dtype=torch.uint8, device=dev)
            mask[0, :, :, Sk - 50:] = 1
            mask[1, :, :, Sk - 7:] = 1
        elif mt == "full":
            mask = (torch.rand(B, 1, Sq, Sk, device=dev) < 0.3).to(torch.uint8)
        scale = 1.0 / math.sqrt(d)
        o, lse = run_fwd(q, k, v, scale, causal, window, mask)
        torch.cuda.synchronize()
        ro, rlse = ref_attention(q, k, v, scale, causal, window, mask)
        err = (o.float() - ro).abs().max().item()
        lerr = (lse.reshape(B, H, Sq) * math.log(2.0) - rlse).abs().max().item()
        good = err < 2e-2 and lerr < 2e-3
        ok &= good
        print(f"fwd B{B} H{H} Sq{Sq} Sk{Sk} d{d} causal={causal} win={window} mask={mt} {dt}: max|o-ref|={err:.3e} max|lse-ref|={lerr:.3e} {'OK' if good else 'FAIL'}", flush=True)
        # token-major output layout
        o2, _ = run_fwd(q, k, v, scale, causal, window, mask, out_layout="bshd")
        torch.cuda.synchronize()
        e2 = (o2.transpose(1, 2).float() - o.float()).abs().max().item()
        if e2 != 0.0:
            ok = False
            print(f"   token-major layout differs: {e2}")
    if args.time:
        for (name, B, H, S, d) in [("gptj", 4, 16, 2048, 256), ("neox20b", 2, 64, 2048, 96), ("gpt2xl", 4, 25, 1024, 64),
                                   ("d128", 4, 32, 2048, 128), ("gptj-8k", 1, 16, 8192, 256)]:
            q = torch.randn(B, H, S, d, device=dev, dtype=torch.bfloat16)
            k = torch.randn_like(q)
            v = torch.randn_like(q)
            scale = 1.0 / math.sqrt(d)
            for _ in range(3):
                run_fwd(q, k, v, scale, True)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            iters = 1500                  # ~0.3-0.8 s per shape (operands of the large shapes exceed the 126 MB L2)
            with clocks(f"fwd {name}"):
                e0.record()
                for _ in range(iters):
                    run_fwd(q, k, v, scale, True)
                e1.record()
                torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / iters
            flops = 4.0 * B * H * S * S * d / 2
            print(f"time fwd {name} B{B} H{H} S{S} d{d} causal: {ms:.3f} ms  {flops / ms / 1e9:.0f} TFLOP/s (causal flops)", flush=True)
    if args.bwd:
        ok &= check_bwd(dev)
        if args.time:
            time_bwd(dev)
    print("ALL OK" if ok else "SOME FAILED")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
