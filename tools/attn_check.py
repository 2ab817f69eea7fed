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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--time", action="store_true")
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
            e0.record()
            for _ in range(10):
                run_fwd(q, k, v, scale, True)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            flops = 4.0 * B * H * S * S * d / 2
            print(f"time fwd {name} B{B} H{H} S{S} d{d} causal: {ms:.3f} ms  {flops / ms / 1e9:.0f} TFLOP/s (causal flops)", flush=True)
    print("ALL OK" if ok else "SOME FAILED")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
