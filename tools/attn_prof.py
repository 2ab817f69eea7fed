"""One forward + backward launch of the fused attention kernels at a model shape (target of an ncu capture).

    python tools/attn_prof.py [gptj|neox|d128|gpt2xl]
"""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.attn_check import run_bwd, run_fwd  # noqa: E402

SHAPES = {"gptj": (4, 16, 2048, 256), "neox": (2, 64, 2048, 96), "d128": (4, 32, 2048, 128), "gpt2xl": (4, 25, 1024, 64)}


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "gptj"
    B, H, S, d = SHAPES[name]
    dev = torch.device("cuda")
    torch.manual_seed(0)
    q = torch.randn(B, H, S, d, device=dev, dtype=torch.bfloat16)
    k, v, do = torch.randn_like(q), torch.randn_like(q), torch.randn_like(q)
    scale = 1.0 / math.sqrt(d)
    for _ in range(2):
        o, lse = run_fwd(q, k, v, scale, True)
        run_bwd(q, k, v, o, do, lse, scale, True)
    torch.cuda.synchronize()
    print("done", name)


if __name__ == "__main__":
    main()
