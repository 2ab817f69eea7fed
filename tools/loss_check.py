"""Does a model 'train' or merely 'run'?  For a few small configurations (BERT-style post-LN encoder and GPT-J style
decoder, fp16 with dynamic loss scaling and bf16) build the model in half precision through the public API (tcgen05
GEMMs, flash attention, fused cross-entropy, multi-tensor Adam on fp32 masters), copy its weights into an fp32 twin
(fp32 inputs are outside the 16-bit kernels: library GEMMs, library attention) and compare: the loss of the first forward pass (must agree to ~1e-2 and be close to ln(vocab) at random init),
and the loss after a handful of optimizer steps on one repeated batch (must fall on both).

Usage (one GPU):  python tools/loss_check.py  [--steps 6]
Prints one line per configuration and ``ALL OK`` / ``MISMATCH``; exit code 0 only when every configuration agrees.
"""

from __future__ import annotations

import argparse
import copy
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def _cfg(kind, vocab, seq):
    from smdistributed_modelparallel_b200.models.configs import MODEL_CONFIGS

    base = dict(MODEL_CONFIGS["bert-large" if kind.startswith("bert") else "gpt-j-6b"])
    if kind == "bert-wide":        # BERT-large's own widths and vocabulary (30,522 is not a multiple of 8), 2 layers
        base.update(num_layers=2)
        return base
    base.update(num_layers=2, hidden_size=256, num_attention_heads=4, attention_head_size=64, intermediate_size=1024,
                vocab_size=vocab, num_positions=seq)
    if kind == "gptj":
        base.update(rotary_dim=16, causal_mask_size=seq)
    return base


def run(kind, dtype, steps, fp16_scale):
    import smdistributed_modelparallel_b200 as smp

    vocab, seq, batch = (30522, 512, 4) if kind == "bert-wide" else (3072, 128, 8)
    cfg = _cfg(kind, vocab, seq)
    losses = {}
    sd = None
    cuda = torch.cuda.is_available()           # without a GPU this is a dry run of the script itself
    for where in ("gpu", "cpu"):               # "cpu" = the fp32 twin (on the same device)
        smp.reset()
        conf = {"microbatches": 2}
        if where == "gpu":
            conf["fp16" if dtype == torch.float16 else "bf16"] = True
        smp.init(conf)
        torch.manual_seed(0)
        if where == "gpu":
            with smp.model_creation(dtype=dtype):
                module = smp.nn.DistributedTransformerLMHead(**cfg)
        else:
            module = smp.nn.DistributedTransformerLMHead(**cfg)
            module.load_state_dict(sd)
            module = module.float()
        if where == "gpu":
            sd = {k: v.detach().float().cpu().clone() for k, v in copy.deepcopy(module).state_dict().items()}
        model = smp.DistributedModel(module)
        opt = smp.DistributedOptimizer(torch.optim.AdamW(model.parameters(), lr=3e-4, weight_decay=0.0),
                                       **({"dynamic_loss_scale": True} if (where == "gpu" and fp16_scale) else {}))

        @smp.step
        def train_step(model, ids):
            loss, _ = model((ids, None, None, None, ids))
            model.backward(loss)
            return loss

        g = torch.Generator().manual_seed(5)
        ids = torch.randint(0, vocab, (batch, seq), generator=g)
        if cuda:
            ids = ids.cuda()
        out = []
        applied = 0
        while applied < steps and len(out) < steps + 40:
            opt.zero_grad()
            o = train_step(model, ids)
            before = [p.detach().float().sum().item() for p in list(model.parameters())[:2]]
            opt.step()
            after = [p.detach().float().sum().item() for p in list(model.parameters())[:2]]
            val = float(o.reduce_mean().item())
            if before != after or where == "cpu":       # a step skipped by the loss scaler changes nothing
                applied += 1
                out.append(val)
        losses[where] = out
    return losses, math.log(vocab)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=6)
    args = ap.parse_args()
    ok = True
    cases = [("bert", torch.float16, True), ("bert", torch.bfloat16, False), ("gptj", torch.bfloat16, False),
             ("gptj", torch.float16, True)]
    if torch.cuda.is_available():
        cases += [("bert-wide", torch.float16, True), ("bert-wide", torch.bfloat16, False)]
    for kind, dtype, scale in cases:
        losses, lnv = run(kind, dtype, args.steps, scale)
        g, c = losses["gpu"], losses["cpu"]
        first_ok = abs(g[0] - c[0]) < 5e-2 and abs(g[0] - lnv) < 2.5
        falls = g[-1] < g[0] - 0.02 and c[-1] < c[0] - 0.02
        tracks = abs(g[-1] - c[-1]) < 0.15
        good = first_ok and falls and tracks
        ok &= good
        print(f"{kind} {str(dtype).replace('torch.', '')}: ln(V)={lnv:.3f} gpu {' '.join(f'{x:.3f}' for x in g)} | "
              f"fp32-twin {' '.join(f'{x:.3f}' for x in c)} -> {'OK' if good else 'MISMATCH'}", flush=True)
    print("ALL OK" if ok else "MISMATCH")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
