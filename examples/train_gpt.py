"""Training script in the shape SageMaker model-parallel users write them (the reference ships none; this follows its
public API, reference smdistributed/modelparallel/torch/__init__.py:31-163): a HuggingFace causal LM, wrapped by
``smp.DistributedModel`` / ``smp.DistributedOptimizer``, stepped with ``@smp.step``, checkpointed with
``smp.save_checkpoint`` / ``smp.resume_from_checkpoint``.  It imports the library under its original name
(``smdistributed.modelparallel.torch``), which resolves to the B200-native implementation in this repository.

One process per GPU (``torchrun --nproc-per-node N examples/train_gpt.py ...``); without a GPU it runs on CPU / gloo,
which is how ``tests/test_examples_gloo.py`` exercises it.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/train_gpt.py --model gpt-j-6b \
        --tensor-parallel-degree 4 --pipeline-parallel-degree 2 --microbatches 8 --bf16 --steps 20
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/train_gpt.py --model gpt-neox-20b \
        --sharded-data-parallel-degree 8 --activation-checkpointing --offload-activations --bf16
"""

from __future__ import annotations

import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import smdistributed.modelparallel.torch as smp  # noqa: E402

SIZES = {
    # name: (family, layers, hidden, heads, positions, vocab)
    "gpt2-tiny": ("gpt2", 2, 64, 4, 64, 256),
    "gptj-tiny": ("gptj", 2, 64, 4, 64, 256),
    "gpt2-xl": ("gpt2", 48, 1600, 25, 1024, 50257),
    "gpt-j-6b": ("gptj", 28, 4096, 16, 2048, 50400),
    "gpt-neox-20b": ("gptneox", 44, 6144, 64, 2048, 50432),
}


def build_hf_model(name):
    family, layers, hidden, heads, positions, vocab = SIZES[name]
    if family == "gpt2":
        from transformers import GPT2Config, GPT2LMHeadModel
        return GPT2LMHeadModel(GPT2Config(n_layer=layers, n_embd=hidden, n_head=heads, n_positions=positions,
                                          vocab_size=vocab, resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)), vocab, positions
    if family == "gptj":
        from transformers import GPTJConfig, GPTJForCausalLM
        return GPTJForCausalLM(GPTJConfig(n_layer=layers, n_embd=hidden, n_head=heads, n_positions=positions,
                                          vocab_size=vocab, rotary_dim=min(64, hidden // heads // 2 * 2))), vocab, positions
    from transformers import GPTNeoXConfig, GPTNeoXForCausalLM
    return GPTNeoXForCausalLM(GPTNeoXConfig(num_hidden_layers=layers, hidden_size=hidden, num_attention_heads=heads,
                                            intermediate_size=4 * hidden, max_position_embeddings=positions,
                                            vocab_size=vocab)), vocab, positions


def parse(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="gpt2-tiny", choices=sorted(SIZES))
    ap.add_argument("--tensor-parallel-degree", type=int, default=1)
    ap.add_argument("--pipeline-parallel-degree", type=int, default=1)
    ap.add_argument("--sharded-data-parallel-degree", type=int, default=1)
    ap.add_argument("--microbatches", type=int, default=2)
    ap.add_argument("--batch-per-rank", type=int, default=4)
    ap.add_argument("--seq", type=int, default=0)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--bf16", action="store_true")
    ap.add_argument("--fp16", action="store_true")
    ap.add_argument("--activation-checkpointing", action="store_true")
    ap.add_argument("--offload-activations", action="store_true")
    ap.add_argument("--shard-optimizer-state", action="store_true")
    ap.add_argument("--delayed-parameter-initialization", action="store_true")
    ap.add_argument("--checkpoint-dir", default="")
    ap.add_argument("--save-every", type=int, default=0)
    ap.add_argument("--resume", action="store_true")
    ap.add_argument("--full-checkpoint", action="store_true", help="gathered (partial=False) instead of per-rank files")
    return ap.parse_args(argv)


def main(argv=None):
    args = parse(argv)
    cuda = torch.cuda.is_available()
    cfg = {
        "tensor_parallel_degree": args.tensor_parallel_degree,
        "pipeline_parallel_degree": args.pipeline_parallel_degree,
        "microbatches": args.microbatches,
        "ddp": True,
        "bf16": args.bf16,
        "fp16": args.fp16,
        "offload_activations": args.offload_activations,
        "shard_optimizer_state": args.shard_optimizer_state,
        "delayed_parameter_initialization": args.delayed_parameter_initialization,
    }
    if args.sharded_data_parallel_degree > 1:
        cfg["sharded_data_parallel_degree"] = args.sharded_data_parallel_degree
    if not cuda:
        cfg["ddp_dist_backend"] = "gloo"
    smp.init(cfg)
    torch.manual_seed(1234)                         # the same initial weights on every rank
    half = torch.bfloat16 if args.bf16 else (torch.float16 if args.fp16 else None)
    with smp.model_creation(tensor_parallelism=args.tensor_parallel_degree > 1, dtype=half):
        net, vocab, positions = build_hf_model(args.model)
    if args.activation_checkpointing:
        blocks = net.transformer.h if hasattr(net, "transformer") else net.gpt_neox.layers
        smp.set_activation_checkpointing(blocks, strategy="each")
    model = smp.DistributedModel(net)
    optimizer = smp.DistributedOptimizer(torch.optim.AdamW(model.parameters(), lr=args.lr, weight_decay=0.01),
                                         dynamic_loss_scale=args.fp16)

    @smp.step
    def train_step(model, ids):
        loss = model(input_ids=ids, labels=ids).loss
        model.backward(loss)
        return loss

    start = 0
    if args.resume and args.checkpoint_dir:
        extra = smp.resume_from_checkpoint(args.checkpoint_dir, partial=not args.full_checkpoint)
        start = int((extra or {}).get("step", 0))
    seq = args.seq or min(positions, 32 if not cuda else positions)
    dev = torch.device("cuda", smp.local_rank()) if cuda else torch.device("cpu")
    # every data-parallel rank draws its own batches; ranks of one tensor-parallel group do too (non-prescaled batch)
    data = torch.Generator().manual_seed(100 + smp.dp_rank())
    fixed = torch.randint(0, vocab, (args.batch_per_rank, seq), generator=data)
    losses = []
    for step in range(start, start + args.steps):
        ids = fixed.to(dev)                         # one repeated batch: the loss has to fall
        t0 = time.perf_counter()
        optimizer.zero_grad()
        out = train_step(model, ids)
        optimizer.step()
        loss = out.reduce_mean().item()
        losses.append(loss)
        if smp.rank() == 0:
            print(f"step {step} loss {loss:.4f} ({(time.perf_counter() - t0) * 1e3:.0f} ms)", flush=True)
        if args.checkpoint_dir and args.save_every and (step + 1) % args.save_every == 0:
            smp.save_checkpoint(args.checkpoint_dir, tag=f"step{step + 1}", partial=not args.full_checkpoint, model=model,
                                optimizer=optimizer, user_content={"step": step + 1})
    smp.barrier()
    return {"losses": losses, "start": start, "rank": smp.rank(), "pp_rank": smp.pp_rank(), "tp_rank": smp.tp_rank()}


if __name__ == "__main__":
    main()
