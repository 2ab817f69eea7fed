"""Tensor-parallel embedding: vocabulary-parallel (rows split) or hidden-parallel (columns split).

Reference: smdistributed/modelparallel/torch/nn/embedding.py:24-290 (``DistributedEmbedding``; vocab
range per rank, masked lookup + all-reduce or reduce-scatter :133-171, hidden-parallel all-gather +
all-to-all :175-197).  Output layout here is token-parallel: every rank returns the embeddings of its
own tokens, so the vocab-parallel path is ids all-gather -> masked lookup -> reduce-scatter over
tokens.
"""

from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..parallel import tp_ops
from .dist_module import DistributedModule
from .utils import (allgather_for_tp, get_local_channels, get_start_pos_for_slicing, initialize_with_input_partition,
                    initialize_with_output_partition, parameter_creation_scope, tp_group, tp_rank, tp_size)


class DistributedEmbedding(DistributedModule):
    def __init__(self, num_embeddings, embedding_dim, padding_idx=None, max_norm=None, norm_type=2.0,
                 scale_grad_by_freq=False, sparse=False, _weight=None, initializer_range=0.02, vocab_parallel=True,
                 _skip_allgather=False, _output_full_batch=False, dtype=None, device=None,
                 _skip_scatter_and_merge=False):
        super().__init__()
        assert max_norm is None and not scale_grad_by_freq and not sparse, \
            "max_norm / scale_grad_by_freq / sparse are not supported by DistributedEmbedding"
        self.num_embeddings, self.embedding_dim = num_embeddings, embedding_dim
        self.padding_idx = padding_idx
        self.vocab_parallel = vocab_parallel
        self._skip_allgather = _skip_allgather      # inputs are already identical on every tp rank
        self._output_full_batch = _output_full_batch
        # hidden-parallel only: keep the looked-up columns [tokens, H/t] (reference nn/embedding.py:195)
        self._skip_scatter_and_merge = _skip_scatter_and_merge
        if vocab_parallel:
            self.local_num_embeddings = get_local_channels(num_embeddings)
            self.vocab_start_idx = get_start_pos_for_slicing(num_embeddings)
            self.vocab_end_idx = self.vocab_start_idx + self.local_num_embeddings
            shape = (self.local_num_embeddings, embedding_dim)
            scope = initialize_with_output_partition(self, axis=0)
        else:
            self.local_embedding_dim = get_local_channels(embedding_dim)
            self.vocab_start_idx, self.vocab_end_idx = 0, num_embeddings
            shape = (num_embeddings, self.local_embedding_dim)
            scope = initialize_with_input_partition(self, axis=1)
        with parameter_creation_scope(self, scaled_batch=True, dtype=dtype):
            with scope:
                self.weight = nn.Parameter(torch.empty(*shape, device=device))
        if _weight is not None:
            self.weight.data.copy_(self._slice_full(_weight))
        else:
            nn.init.normal_(self.weight, 0.0, initializer_range)
            if padding_idx is not None and vocab_parallel and self.vocab_start_idx <= padding_idx < self.vocab_end_idx:
                self.weight.data[padding_idx - self.vocab_start_idx].zero_()
        self.weight_split_shapes = None

    def _slice_full(self, full):
        if self.vocab_parallel:
            return full[self.vocab_start_idx:self.vocab_end_idx]
        s = get_start_pos_for_slicing(self.embedding_dim)
        return full[:, s:s + self.local_embedding_dim]

    def forward(self, ids):
        t = tp_size()
        g = tp_group() if t > 1 else None
        shape = ids.shape
        flat = ids.reshape(-1)
        if t == 1:
            return F.embedding(ids, self.weight, self.padding_idx)
        full_ids = flat if self._skip_allgather else tp_ops.gather_tokens(flat.contiguous(), g)
        if self.vocab_parallel:
            local = full_ids - self.vocab_start_idx
            outside = (local < 0) | (local >= self.local_num_embeddings)
            local = local.masked_fill(outside, 0)
            part = F.embedding(local, self.weight)
            part = part.masked_fill(outside[:, None], 0.0)
            if self._output_full_batch:
                from .utils import fwd_allreduce_for_tp
                out = fwd_allreduce_for_tp(part)
                return out.reshape(-1, *shape[1:], self.embedding_dim) if self._skip_allgather else \
                    out.reshape(t * shape[0], *shape[1:], self.embedding_dim)
            out = tp_ops.ScatterSumTokens.apply(part, g)            # [T/t, H]: this rank's tokens
            if self._skip_allgather:                                 # prescaled batch: tokens were shared
                return out.reshape(-1, self.embedding_dim)
            return out.reshape(*shape, self.embedding_dim)
        # hidden-parallel: every rank looks up all tokens for its columns, then tokens <-> columns all-to-all
        part = F.embedding(full_ids, self.weight)                    # [T, H/t]
        if self._skip_scatter_and_merge:
            lead = shape if self._skip_allgather else (t * shape[0], *shape[1:])
            return part.reshape(*lead, part.shape[-1])
        from .utils import scatter_and_merge_for_tp
        out = scatter_and_merge_for_tp(part, split_dim=0, merge_dim=1)   # [T/t, H]
        if self._skip_allgather:
            return out.reshape(-1, self.embedding_dim)
        return out.reshape(*shape, self.embedding_dim)
