"""Timed comparison arm: the reference's hot path restated with stock PyTorch and NCCL only.

Nothing in this package imports the product's kernels, ops, engine or fused collectives - only
``smdistributed_modelparallel_b200.models.configs`` (architecture numbers).  What it restates, with the reference
call sites it stands for (paths under smdistributed/modelparallel/torch/):

* tensor-parallel transformer layer in the reference's batch-gather layout: all-gather of the batch over the tp group
  at the layer stack's entry, three ``F.linear`` for q / k / v, ``baddbmm`` scores, softmax, ``matmul`` with V, a
  row-parallel ``F.linear`` followed by an **all-reduce**, GELU MLP with a second all-reduce, all-reduce of the input
  gradients in backward, narrow at exit (nn/transformer.py:1436-1546,1077-1173; nn/utils.py:540-571,701-821) - every
  collective a ``torch.distributed`` call on the compute stream;
* vocab-parallel embedding / LM head with the three-all-reduce cross-entropy (nn/cross_entropy.py:28-98);
* 16-bit parameters with fp32 master weights and a fused multi-tensor AdamW (fp16/fp16.py:380-464,597-651);
* data parallelism as bucketed all-reduce overlapped with backward (``torch.nn.parallel.DistributedDataParallel``,
  ddp_model.py:278-374);
* sharded data parallelism as ZeRO-3 all-gather / reduce-scatter (``FullyShardedDataParallel``; the reference
  delegates this to DeepSpeed stage 3, optimizers/optimizer.py:471-491), activation checkpointing and pinned-memory
  offload of the checkpoints (offload.py:76-327);
* pipeline parallelism as NCCL point-to-point between stages with a one-forward-one-backward schedule
  (pipeline.py:117-145; server_comm.py:260-302).
"""
