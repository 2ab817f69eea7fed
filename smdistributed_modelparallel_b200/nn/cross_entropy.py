"""``DistributedCrossEntropy``: loss over logits whose vocabulary dimension is split across tp ranks.

Reference: smdistributed/modelparallel/torch/nn/cross_entropy.py:100-112 (module) over
``_VocabParallelCrossEntropy`` (:28-98).  Runs the fused native kernels (ops/cross_entropy.py).
"""

from __future__ import annotations

import torch.nn as nn

from ..ops.cross_entropy import vocab_parallel_cross_entropy
from .utils import tp_group, tp_size


class DistributedCrossEntropy(nn.Module):
    def __init__(self, vocab_range=None, ignore_index: int = -100):
        super().__init__()
        self.vocab_range = vocab_range
        self.ignore_index = ignore_index

    def forward(self, vocab_parallel_logits, target):
        start = self.vocab_range[0] if self.vocab_range is not None else 0
        group = tp_group() if tp_size() > 1 else None
        return vocab_parallel_cross_entropy(vocab_parallel_logits, target, start, group, self.ignore_index)
