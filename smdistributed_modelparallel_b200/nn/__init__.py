"""``smp.nn``: tensor-parallel modules (reference smdistributed/modelparallel/torch/nn/__init__.py:24-35)."""

from .cross_entropy import DistributedCrossEntropy
from .dist_module import DistributedModule
from .embedding import DistributedEmbedding
from .layer_norm import DistributedLayerNorm, FusedLayerNorm, MixedFusedLayerNorm
from .linear import DistributedLinear
from .transformer import (DistributedAttentionLayer, DistributedTransformer, DistributedTransformerLayer,
                          DistributedTransformerLMHead, DistributedTransformerOutputLayer)

__all__ = [
    "DistributedLinear", "DistributedEmbedding", "DistributedLayerNorm", "DistributedTransformerLayer",
    "DistributedAttentionLayer", "DistributedTransformerOutputLayer", "DistributedTransformer",
    "DistributedTransformerLMHead", "DistributedModule", "FusedLayerNorm", "MixedFusedLayerNorm",
    "DistributedCrossEntropy",
]
