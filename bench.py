#!/usr/bin/env python
"""Flagship training benchmark (driver contract: see README / DESIGN.md).

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Metric: training tokens/s of GPT-J 6B (bf16, seq 2048, random-init weights, synthetic tokens) through the
library's public API - smp.init / smp.DistributedModel / smp.DistributedOptimizer / @smp.step - one
optimizer step per timed step.  Rank 0 prints ONE JSON line.

  value      device-timed (CUDA events, max over ranks) with the token batch already resident on the GPU
  e2e.value  the same step fed from pinned host memory (H2D copy of the ids every step) and read back
             (D2H of the loss every step)

`--impl reference` would run the unmodified reference; it cannot be installed here (binary-only wheel
with its native libraries missing), so that arm reports `unavailable`.
`--impl baseline` runs the reference's hot path restated with stock PyTorch + NCCL (package `baseline/`, nothing
of this library on that path) for the same model / N / batch geometry; its measured tokens/s are kept in
`baseline/measured.json` and the product line reports `vs_baseline` against them.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


METRIC_NAMES = {"gpt-j-6b": "GPT-J-6B", "gpt-neox-20b": "GPT-NeoX-20B", "gpt2-xl": "GPT-2-XL-1.5B",
                "bert-large": "BERT-large", "roberta-large": "RoBERTa-large"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "baseline"])
    ap.add_argument("--model", default=os.environ.get("BENCH_MODEL", "gpt-j-6b"))
    ap.add_argument("--seq", type=int, default=0)
    ap.add_argument("--micro-batch", type=int, default=0, help="sequences per microbatch per data-parallel rank")
    ap.add_argument("--microbatches", type=int, default=0)
    ap.add_argument("--parallelism", default="", help="e.g. tp2, pp2, pp2tp4, sdp8, dp2 (default: chosen per N)")
    ap.add_argument("--checkpoint", type=int, default=-1, help="activation checkpointing (1/0, default auto)")
    ap.add_argument("--active-microbatches", type=int, default=0, help="pipeline window (default: config default)")
    ap.add_argument("--offload", type=int, default=0, help="offload checkpointed activations to pinned host memory")
    ap.add_argument("--fp16", type=int, default=0, help="fp16 with dynamic loss scaling instead of bf16")
    ap.add_argument("--shard-optimizer-state", type=int, default=0, help="ZeRO-1 style optimizer-state sharding over dp")
    ap.add_argument("--layers", type=int, default=0, help="debug only: override layer count (marks the run invalid)")
    return ap.parse_args()


def reference_arm(args):
    why = ("reference is an unpacked cp38 binary wheel with no setup.py/pyproject and all five native libraries "
           "(smplib.so, smp_torch_cuda_lib.so, amp_C.so, ...) absent; pip install of /root/reference fails: not installable")
    if int(os.environ.get("RANK", "0")) == 0:          # one line per job, also under torchrun
        print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)
    return 0


def _roofline(model_tflops):
    """Model TFLOP/s per GPU against the measured cuBLAS bf16 throughput of this pod (driver-written
    MEASURED_PEAKS.json; B200_PROFILING.md's stated numbers when the file is absent)."""
    peaks = {"bf16_tflops": 1664.2, "bf16_tflops_sustained": 1351.7, "source": "fallback (B200_PROFILING.md)"}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            m = json.load(f)
        peaks = {"bf16_tflops": float(m["bf16_tflops"]), "bf16_tflops_sustained": float(m["bf16_tflops_sustained"]),
                 "source": "MEASURED_PEAKS.json"}
    except (OSError, ValueError, KeyError):
        pass
    return {"cublas_bf16_tflops_burst": peaks["bf16_tflops"], "cublas_bf16_tflops_sustained": peaks["bf16_tflops_sustained"],
            "fraction_of_sustained": round(model_tflops / peaks["bf16_tflops_sustained"], 3),
            "fraction_of_burst": round(model_tflops / peaks["bf16_tflops"], 3), "peaks_from": peaks["source"]}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx = float(parts[2])
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def pick_parallelism(n, spec):
    """Returns (pp, tp, sdp, label)."""
    import re

    if not spec:
        spec = {1: "dp1", 2: "tp2", 4: "tp4", 8: "pp2tp4"}.get(n, f"tp{n}")
    pp = int(m.group(1)) if (m := re.search(r"pp(\d+)", spec)) else 1
    tp = int(m.group(1)) if (m := re.search(r"tp(\d+)", spec)) else 1
    sdp = int(m.group(1)) if (m := re.search(r"sdp(\d+)", spec)) else 1
    return pp, tp, sdp, spec


def baseline_table():
    """Tokens/s of the comparison arm (``--impl baseline``) measured on B200 this round, keyed
    "model/N/parallelism"; written by hand from the logs under profiles/ (see BASELINE.md section 5)."""
    path = os.path.join(ROOT, "baseline", "measured.json")
    try:
        with open(path) as f:
            return json.load(f)
    except (OSError, ValueError):
        return {}


def run_product(args, model_name, parallelism, steps, warmup, micro=0, nmb=0, checkpoint=-1, offload=0):
    """One timed run of the library through its public API; returns the JSON line (rank 0) or None."""
    import torch
    import torch.distributed as dist

    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.models.configs import MODEL_CONFIGS, param_count, train_flops_per_token
    from smdistributed_modelparallel_b200.ops import _native as N

    n = int(os.environ.get("WORLD_SIZE", "1"))
    pp, tp, sdp, label = pick_parallelism(n, parallelism)
    dp = n // (pp * tp)
    mcfg = dict(MODEL_CONFIGS[model_name])
    if args.layers:
        mcfg["num_layers"] = args.layers
    seq = args.seq or mcfg["num_positions"]
    # per-N batch geometry (weak scaling: sequences per step grow with N)
    micro = micro or args.micro_batch or 4
    nmb = nmb or args.microbatches or (2 if pp == 1 else 8)

    half = torch.float16 if args.fp16 else torch.bfloat16
    conf = {"microbatches": nmb, "bf16": not args.fp16, "fp16": bool(args.fp16),
            "shard_optimizer_state": bool(args.shard_optimizer_state), "pipeline_parallel_degree": pp, "tensor_parallel_degree": tp,
            "ddp": n > 1, "delayed_parameter_initialization": True, "pipeline": "interleaved"}
    if sdp > 1:
        conf["sharded_data_parallel_degree"] = sdp
    if pp > 1:      # pipeline window: microbatches in flight on the leader (memory vs. overlap; 4 measured best at pp2)
        conf["active_microbatches"] = args.active_microbatches or min(4, nmb)
        # balance the stages by time, not by parameter bytes: 180 GB per GPU leaves memory to spare here, and the
        # embedding table would otherwise count like several (tensor-parallel) layers
        conf["memory_weight"] = float(os.environ.get("BENCH_MEMORY_WEIGHT", "0.0"))
    if offload:
        conf["offload_activations"] = True
    if tp > 1:
        conf["prescaled_batch"] = False
    if os.environ.get("BENCH_TP_COMM") == "nccl":
        conf["fused_tp_collectives"] = False       # ablation: NCCL collectives next to the library's own GEMMs
    os.environ.setdefault("SMP_LOG_LEVEL", "warning")
    smp.init(conf)
    from smdistributed_modelparallel_b200.parallel import tp_ops
    dev = torch.device("cuda", smp.local_rank())
    torch.manual_seed(1234 + smp.rank())

    with smp.model_creation(tensor_parallelism=tp > 1, dtype=half):
        module = smp.nn.DistributedTransformerLMHead(**mcfg)
    use_ckpt = checkpoint if checkpoint >= 0 else (args.checkpoint if args.checkpoint >= 0 else (1 if model_name == "gpt-neox-20b" else 0))
    if use_ckpt:
        smp.set_activation_checkpointing(module.transformer.seq_layers, pack_args_as_tuple=True, strategy="each")
    # pipeline stages are balanced by the partitioner's time estimate (GEMM weights per tp rank; the configuration
    # measured on 8 GPUs).  BENCH_TRACE=1 balances them on measured module times instead (first-step tracing; the
    # parameters are still on the meta device then and borrow storage for the traced passes) - measured on 4 GPUs only.
    trace = pp > 1 and os.environ.get("BENCH_TRACE", "0") != "0"
    model = smp.DistributedModel(module, trace_execution_times=trace)
    optimizer = smp.DistributedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-4, betas=(0.9, 0.95),
                                                           weight_decay=0.1),
                                         dynamic_loss_scale=bool(args.fp16))

    @smp.step
    def train_step(model, ids):
        loss, _ = model((ids, None, None, None, ids))
        model.backward(loss)
        return loss

    # batch seen by one rank per step; tp ranks each bring their own sequences (non-prescaled batch)
    per_rank_batch = micro * nmb
    global_batch = per_rank_batch * (dp * tp)
    tokens_per_step = global_batch * seq
    vocab = mcfg["vocab_size"]
    host_ids = [torch.randint(0, vocab, (per_rank_batch, seq), dtype=torch.int64).pin_memory() for _ in range(4)]
    dev_ids = torch.randint(0, vocab, (per_rank_batch, seq), dtype=torch.int64, device=dev)

    def one_step(ids):
        optimizer.zero_grad()
        out = train_step(model, ids)
        optimizer.step()
        return out

    def sync_all():
        torch.cuda.synchronize()
        if n > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(warmup, 1)):
        one_step(dev_ids)
    sync_all()
    torch.cuda.reset_peak_memory_stats()

    sampler = ClockSampler(smp.local_rank())
    if smp.rank() == 0:
        sampler.start()
    # ---- phase A: device-resident batch, CUDA events
    N.reset_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    e0.record()
    for _ in range(steps):
        out = one_step(dev_ids)
    e1.record()
    sync_all()
    ms_a = e0.elapsed_time(e1) / steps
    launches = N.launches()
    # ---- phase B: end to end (pinned host ids -> device every step, loss read back every step)
    sync_all()
    t0 = time.perf_counter()
    e0.record()
    last = None
    seen = []
    for i in range(steps):
        ids = host_ids[i % len(host_ids)].to(dev, non_blocking=True)
        out = one_step(ids)
        last = float(out.reduce_mean().item())
        seen.append(round(last, 4))
    e1.record()
    sync_all()
    ms_b = e0.elapsed_time(e1) / steps
    wall_b = (time.perf_counter() - t0) * 1000 / steps
    clocks = sampler.stop() if smp.rank() == 0 else None

    if os.environ.get("BENCH_TORCH_PROFILE"):
        # diagnosis only (after the timed regions): one step under torch.profiler, summarised per rank on the box
        from torch.profiler import ProfilerActivity, profile
        sync_all()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            one_step(dev_ids)
            sync_all()
        os.makedirs(os.environ["BENCH_TORCH_PROFILE"], exist_ok=True)
        prof.export_chrome_trace(os.path.join(os.environ["BENCH_TORCH_PROFILE"], f"trace_rank{smp.rank()}.json"))

    t = torch.tensor([ms_a, max(ms_b, wall_b)], device=dev, dtype=torch.float64)
    if n > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_a, ms_b = float(t[0]), float(t[1])
    peak = torch.cuda.max_memory_allocated() / 2 ** 30
    line = None
    if smp.rank() == 0:
        value = tokens_per_step / (ms_a / 1000)
        e2e = tokens_per_step / (ms_b / 1000)
        flops = train_flops_per_token(mcfg, seq)                  # model FLOPs: recomputation is not counted
        base = baseline_table().get(f"{model_name}/{n}/{label}")
        line = {
            "metric": f"training tokens/sec (device-timed, max over ranks), {METRIC_NAMES.get(model_name, model_name)} {'fp16' if args.fp16 else 'bf16'} seq {seq}",
            "value": round(value, 1), "unit": "tokens/s", "n_gpus": n, "steps": steps, "warmup": warmup,
            "ms_per_step": round(ms_a, 2), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": round(value / base, 3) if base else None,
            "dtype": "fp16" if args.fp16 else "bf16", "data": "synthetic tokens, random-init weights",
            "impl": "b200",
            "config": {"model": model_name, "params": param_count(mcfg), "layers": mcfg["num_layers"],
                       "global_batch": global_batch, "seq_len": seq, "microbatches": nmb, "micro_batch": micro,
                       "parallelism": label, "tp_comm": tp_ops.get_comm_mode() if tp > 1 else None, "activation_checkpointing": bool(use_ckpt), "offload_activations": bool(offload),
                       "active_microbatches": conf.get("active_microbatches", nmb) if pp > 1 else None,
                       "memory_weight": conf.get("memory_weight") if pp > 1 else None,
                       "partition": ("auto, traced module times" if trace else "auto, estimated from GEMM weights") if pp > 1 else None,
                       "optimizer": "AdamW fp32 master" + (", state sharded over dp" if args.shard_optimizer_state else ""),
                       "l2": f"working set ({2 * param_count(mcfg) / 2 ** 30:.1f} GiB of 16-bit weights read per pass) >> 126 MB L2; no explicit flush needed",
                       "baseline_tokens_per_s": base,
                       "valid": not bool(args.layers)},
            "model_tflops_per_gpu": round(value * flops / n / 1e12, 1),
            "model_tflops_note": "6 x matmul params + full (non-causal) attention products per token; the kernels skip the masked half of causal attention",
            "roofline": _roofline(value * flops / n / 1e12),
            "clocks": clocks,
            "e2e": {"value": round(e2e, 1), "unit": "tokens/s", "h2d_bytes_per_step": per_rank_batch * seq * 8,
                    "d2h_bytes_per_step": 4 * nmb, "ms_per_step": round(ms_b, 2), "last_loss": last, "losses": seen[-8:]},
            "gpu_launches": launches,
            "peak_mem_gb": round(peak, 1),
        }
    sync_all()
    # leave the process ready for another configuration
    del model, optimizer, module, train_step
    smp.reset()
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    return line


def main():
    args = parse()
    if args.impl == "reference":
        return reference_arm(args)

    if args.impl == "baseline":
        from baseline.engine import run as run_baseline
        from smdistributed_modelparallel_b200.models.configs import MODEL_CONFIGS, param_count, train_flops_per_token

        mcfg = dict(MODEL_CONFIGS[args.model])
        if args.layers:
            mcfg["num_layers"] = args.layers
        line = run_baseline(args, mcfg, ClockSampler, train_flops_per_token, param_count)
        if line is not None:
            print(json.dumps(line), flush=True)
        return 0

    n = int(os.environ.get("WORLD_SIZE", "1"))
    if n != args.gpus:
        print(f"[bench] WORLD_SIZE={n} but --gpus {args.gpus}; using WORLD_SIZE", file=sys.stderr)
    line = run_product(args, args.model, args.parallelism, args.steps, args.warmup, checkpoint=args.checkpoint,
                       offload=args.offload)
    # BASELINE.json's headline metric has two halves: GPT-J 6B PPxTP (the line above) and GPT-NeoX 20B under sharded data
    # parallelism with activation checkpointing + offload on 8 GPUs; the second rides along as "secondary"
    second = None
    spec = os.environ.get("BENCH_SECONDARY", "1")        # "0": off; "model,parallelism,ckpt,offload": override (tests)
    sec = ("gpt-neox-20b", "sdp8", 1, 1) if (n == 8 and args.model == "gpt-j-6b" and not args.parallelism and spec == "1") else None
    if spec not in ("0", "1"):
        f = spec.split(",")
        sec = (f[0], f[1], int(f[2]), int(f[3]))
    if sec is not None:
        k2 = max(3, min(args.steps, int(os.environ.get("BENCH_SECONDARY_STEPS", "8"))))
        # the headline line must print whatever happens to the second configuration: every rank arms a watchdog that
        # ends the process (rank 0 printing the line first) if the secondary run has not finished in time
        limit = float(os.environ.get("BENCH_SECONDARY_TIMEOUT", "300"))

        def give_up():
            if line is not None:
                line["secondary"] = {"error": f"secondary run ({sec[0]} {sec[1]}) exceeded {limit:.0f} s and was abandoned"}
                print(json.dumps(line), flush=True)
            os._exit(0)

        if line is not None:                                     # a copy for the log, should the process be killed
            print("[bench] headline before the secondary run: " + json.dumps(line), file=sys.stderr, flush=True)
        dog = threading.Timer(limit, give_up)
        dog.daemon = True
        dog.start()
        try:
            second = run_product(args, sec[0], sec[1], k2, min(args.warmup, 3), checkpoint=sec[2], offload=sec[3])
        except Exception as e:                                   # noqa: BLE001 - the headline line must still print
            second = {"error": f"{type(e).__name__}: {str(e)[:300]}"}
        finally:
            dog.cancel()
    if line is not None:
        if second is not None:
            line["secondary"] = second
        print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
