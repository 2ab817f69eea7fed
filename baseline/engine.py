"""Training step of the comparison arm: stock PyTorch + NCCL (see the package docstring)."""

from __future__ import annotations

import contextlib
import os
import re
import time

import torch
import torch.distributed as dist
import torch.nn as nn

from .model import RefGPT


def parse_parallelism(n, spec):
    if not spec:
        spec = {1: "dp1", 2: "tp2", 4: "tp4", 8: "pp2tp4"}.get(n, f"tp{n}")
    pp = int(m.group(1)) if (m := re.search(r"pp(\d+)", spec)) else 1
    tp = int(m.group(1)) if (m := re.search(r"tp(\d+)", spec)) else 1
    sdp = int(m.group(1)) if (m := re.search(r"sdp(\d+)", spec)) else 1
    return pp, tp, sdp, spec


class _Masters:
    """16-bit parameters, fp32 master copies, fused AdamW on the masters (the reference's Bit16_Optimizer +
    multi-tensor Adam, fp16/fp16.py:380-464,597-651)."""

    def __init__(self, params, lr, betas, wd):
        self.params = [p for p in params if p.requires_grad]
        self.masters = [p.detach().float().clone().requires_grad_(True) for p in self.params]
        self.opt = torch.optim.AdamW(self.masters, lr=lr, betas=betas, weight_decay=wd, fused=self.masters[0].is_cuda)

    def zero_grad(self):
        for p in self.params:
            p.grad = None

    def step(self):
        grads = [p.grad for p in self.params]
        for m, g in zip(self.masters, grads):
            if m.grad is None:
                m.grad = torch.empty_like(m)
        torch._foreach_copy_([m.grad for m in self.masters], grads)          # 16-bit grads -> fp32 master grads
        self.opt.step()
        with torch.no_grad():
            torch._foreach_copy_(self.params, self.masters)                 # masters -> 16-bit model weights


def _groups(n, pp, tp):
    """rank = pp_rank * (dp * tp) + dp_rank * tp + tp_rank; returns (pp_rank, dp_rank, tp_rank, tp_group, dp_group)."""
    rank = dist.get_rank() if n > 1 else 0
    dp = n // (pp * tp)
    pr, rem = divmod(rank, dp * tp)
    dr, tr = divmod(rem, tp)
    tp_group = dp_group = None
    if n > 1:
        for a in range(pp):
            for b in range(dp):
                ranks = [a * dp * tp + b * tp + c for c in range(tp)]
                g = dist.new_group(ranks) if tp > 1 else None
                if a == pr and b == dr:
                    tp_group = g
            for c in range(tp):
                ranks = [a * dp * tp + b * tp + c for b in range(dp)]
                g = dist.new_group(ranks) if dp > 1 else None
                if a == pr and c == tr:
                    dp_group = g
    # one 2-rank group per direction between neighbouring stages: activations going forward and gradients coming back
    # are ordered independently (on one communicator, "send a1" on one side and "send g0" on the other wait for each
    # other's next receive)
    fwd_groups, bwd_groups = {}, {}
    if n > 1 and pp > 1:
        stage = dp * tp
        for a in range(pp - 1):
            for i in range(stage):
                pair = [a * stage + i, (a + 1) * stage + i]
                f, b = dist.new_group(pair), dist.new_group(pair)
                if rank in pair:
                    fwd_groups[tuple(pair)] = f
                    bwd_groups[tuple(pair)] = b
    return pr, dr, tr, dp, tp_group, dp_group, fwd_groups, bwd_groups


class Runner:
    def __init__(self, args, mcfg, n, device=None, dtype=torch.bfloat16):
        self.args, self.cfg, self.n = args, mcfg, n
        pp, tp, sdp, self.label = parse_parallelism(n, args.parallelism)
        self.pp, self.tp, self.sdp = pp, tp, sdp
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.dtype = dtype
        if device is not None and torch.device(device).type == "cpu":      # plumbing tests (gloo)
            self.dev = torch.device("cpu")
            if n > 1 and not dist.is_initialized():
                dist.init_process_group("gloo")
        else:
            self.dev = torch.device("cuda", self.local)
            torch.cuda.set_device(self.dev)
            if n > 1 and not dist.is_initialized():
                dist.init_process_group("nccl", device_id=self.dev)
        self.rank = dist.get_rank() if n > 1 else 0
        self.seq = args.seq or mcfg["num_positions"]
        self.micro = args.micro_batch or 4
        self.nmb = args.microbatches or (2 if pp == 1 else 8)
        self.ckpt = bool(args.checkpoint if args.checkpoint >= 0 else (1 if args.model == "gpt-neox-20b" else 0))
        self.offload = bool(args.offload)
        torch.manual_seed(1234 + self.rank)
        if sdp > 1:
            self._build_fsdp()
        else:
            self._build_mp()

    # ---- pp x tp x dp
    def _build_mp(self):
        n, pp, tp = self.n, self.pp, self.tp
        self.pr, self.dr, self.tr, self.dp, self.tp_group, self.dp_group, self.fwd_g, self.bwd_g = _groups(n, pp, tp)
        L = self.cfg["num_layers"]
        per = (L + pp - 1) // pp
        lo, hi = self.pr * per, min(L, (self.pr + 1) * per)
        self.model = RefGPT(self.cfg, self.tp_group, self.dtype, self.dev, (lo, hi), first=self.pr == 0,
                            last=self.pr == pp - 1)
        self.model.ckpt = self.ckpt
        self.ddp = None
        if self.dp > 1:
            from torch.nn.parallel import DistributedDataParallel as DDP
            self.ddp = DDP(self.model, device_ids=[self.local] if self.dev.type == "cuda" else None, process_group=self.dp_group, bucket_cap_mb=100,
                           gradient_as_bucket_view=True)
        self.opt = _Masters(self.model.parameters(), 1e-4, (0.9, 0.95), 0.1)
        self.prev = None if self.pr == 0 else self.rank - self.dp * tp
        self.next = None if self.pr == pp - 1 else self.rank + self.dp * tp
        self.global_batch = self.micro * self.nmb * tp * self.dp

    def _fwd_stage(self, ids, x_in):
        m = self.model
        x = m.embed(ids) if m.first else x_in
        x = m.run_layers(x)
        return m.loss(x, ids) / self.nmb if m.last else x

    def _step_mp(self, ids_all):
        """One optimizer step.  pp == 1: microbatch loop with gradient accumulation; pp > 1: one-forward-one-backward."""
        opt = self.opt
        opt.zero_grad()
        nmb, micro = self.nmb, self.micro
        chunks = [ids_all[i * micro:(i + 1) * micro] for i in range(nmb)]
        total = torch.zeros((), device=self.dev)
        if self.pp == 1:
            for i, ids in enumerate(chunks):
                sync = contextlib.nullcontext() if (self.ddp is None or i == nmb - 1) else self.ddp.no_sync()
                with sync:
                    loss = self._fwd_stage(ids, None)
                    loss.backward()
                total += loss.detach()
            opt.step()
            return total
        H = self.cfg["hidden_size"]
        act_shape = (micro * self.tp, self.seq, H)
        warm = min(self.pp - 1 - self.pr, nmb)
        inputs, outputs, pending = [], [], []

        def fwd(i):
            x_in = None
            if self.prev is not None:
                x_in = torch.empty(act_shape, dtype=self.dtype, device=self.dev)
                dist.recv(x_in, src=self.prev, group=self.fwd_g[(self.prev, self.rank)])
                x_in.requires_grad_(True)
            out = self._fwd_stage(chunks[i], x_in)
            if self.next is not None:
                # asynchronous: in the steady state this stage sends an activation forward while the next stage sends
                # a gradient back, and two blocking sends facing each other never complete
                pending.append((dist.isend(out.detach(), dst=self.next, group=self.fwd_g[(self.rank, self.next)]), out))
            inputs.append(x_in)
            outputs.append(out)

        def bwd(last_mb):
            x_in, out = inputs.pop(0), outputs.pop(0)
            sync = contextlib.nullcontext() if (self.ddp is None or last_mb) else self.ddp.no_sync()
            with sync:
                if self.next is not None:
                    g = torch.empty(act_shape, dtype=self.dtype, device=self.dev)
                    dist.recv(g, src=self.next, group=self.bwd_g[(self.rank, self.next)])
                    out.backward(g)
                else:
                    out.backward()
                    total.add_(out.detach())
            if self.prev is not None:
                pending.append((dist.isend(x_in.grad, dst=self.prev, group=self.bwd_g[(self.prev, self.rank)]), x_in.grad))

        for i in range(warm):
            fwd(i)
        done = 0
        for i in range(warm, nmb):
            fwd(i)
            bwd(done == nmb - 1)
            done += 1
        while done < nmb:
            bwd(done == nmb - 1)
            done += 1
        for work, _keep in pending:
            work.wait()
        opt.step()
        return total

    # ---- sharded data parallelism (ZeRO-3)
    def _build_fsdp(self):
        from torch.distributed.fsdp import MixedPrecisionPolicy, fully_shard

        assert self.pp == 1 and self.tp == 1 and self.sdp == self.n
        with torch.device("meta"):
            model = RefGPT(self.cfg, None, torch.float32, None)
        mp = MixedPrecisionPolicy(param_dtype=torch.bfloat16, reduce_dtype=torch.float32)
        for layer in model.layers:
            fully_shard(layer, mp_policy=mp)
        fully_shard(model, mp_policy=mp)
        model.to_empty(device=self.dev)
        with torch.no_grad():
            for name, p in model.named_parameters():
                if name.endswith("bias") or "bias" in name.split(".")[-1]:
                    p.zero_()
                elif ".ln" in name or name.startswith("ln_f"):
                    p.fill_(1.0)
                else:
                    p.normal_(std=0.02)
        model._rot = None
        model.ckpt = self.ckpt
        self.model = model
        self.fsdp_opt = torch.optim.AdamW(model.parameters(), lr=1e-4, betas=(0.9, 0.95), weight_decay=0.1, fused=True)
        self.global_batch = self.micro * self.nmb * self.n

    def _step_fsdp(self, ids_all):
        nmb, micro = self.nmb, self.micro
        self.fsdp_opt.zero_grad(set_to_none=True)
        total = torch.zeros((), device=self.dev)
        off = torch.autograd.graph.save_on_cpu(pin_memory=True) if self.offload else contextlib.nullcontext()
        for i in range(nmb):
            ids = ids_all[i * micro:(i + 1) * micro]
            self.model.set_requires_gradient_sync(i == nmb - 1)
            with off:
                loss = self.model(ids) / nmb
            loss.backward()
            total += loss.detach()
        torch.nn.utils.clip_grad_norm_(self.model.parameters(), 1.0)      # the reference's default sdp_gradient_clipping
        self.fsdp_opt.step()
        return total

    def step(self, ids_all):
        return self._step_fsdp(ids_all) if self.sdp > 1 else self._step_mp(ids_all)

    def per_rank_batch(self):
        return self.micro * self.nmb


def run(args, mcfg, clock_sampler_cls, flops_fn, params_fn):
    n = int(os.environ.get("WORLD_SIZE", "1"))
    r = Runner(args, mcfg, n)
    dev, seq, vocab = r.dev, r.seq, mcfg["vocab_size"]
    prb = r.per_rank_batch()
    host_ids = [torch.randint(0, vocab, (prb, seq), dtype=torch.int64).pin_memory() for _ in range(4)]
    dev_ids = torch.randint(0, vocab, (prb, seq), dtype=torch.int64, device=dev)

    def sync_all():
        torch.cuda.synchronize()
        if n > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 1)):
        r.step(dev_ids)
    sync_all()
    sampler = clock_sampler_cls(r.local)
    if r.rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    e0.record()
    for _ in range(args.steps):
        r.step(dev_ids)
    e1.record()
    sync_all()
    ms_a = e0.elapsed_time(e1) / args.steps
    sync_all()
    t0 = time.perf_counter()
    e0.record()
    last = None
    for i in range(args.steps):
        ids = host_ids[i % len(host_ids)].to(dev, non_blocking=True)
        last = float(r.step(ids).item())
    e1.record()
    sync_all()
    ms_b = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1000) / args.steps
    clocks = sampler.stop() if r.rank == 0 else None
    t = torch.tensor([ms_a, ms_b], device=dev, dtype=torch.float64)
    if n > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_a, ms_b = float(t[0]), float(t[1])
    if r.rank != 0:
        sync_all()
        return None
    tokens = r.global_batch * seq
    value = tokens / (ms_a / 1000)
    line = {
        "metric": f"training tokens/sec (device-timed, max over ranks), {args.model} bf16 seq {seq}",
        "value": round(value, 1), "unit": "tokens/s", "n_gpus": n, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_a, 2), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic tokens, random-init weights",
        "impl": "baseline (stock PyTorch: F.linear / baddbmm / torch softmax, LayerNorm, GELU; NCCL collectives on the "
                "compute stream; fused AdamW on fp32 masters; DDP / FSDP / send-recv pipeline)",
        "config": {"model": args.model, "params": params_fn(mcfg), "layers": mcfg["num_layers"],
                   "global_batch": r.global_batch, "seq_len": seq, "microbatches": r.nmb, "micro_batch": r.micro,
                   "parallelism": r.label, "activation_checkpointing": r.ckpt, "offload_activations": r.offload,
                   "valid": not bool(args.layers)},
        "model_tflops_per_gpu": round(value * flops_fn(mcfg, seq) / n / 1e12, 1), "clocks": clocks,
        "e2e": {"value": round(tokens / (ms_b / 1000), 1), "unit": "tokens/s", "h2d_bytes_per_step": prb * seq * 8,
                "d2h_bytes_per_step": 4, "ms_per_step": round(ms_b, 2), "last_loss": last},
        "gpu_launches": 0, "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 1),
    }
    sync_all()
    return line
