"""``smp.optimizers`` (reference smdistributed/modelparallel/torch/optimizers/__init__.py)."""

from .fused import FusedAdagrad, FusedAdam, FusedLAMB, FusedNovoGrad, FusedSGD
from .optimizer import DistributedOptimizer

__all__ = ["FusedAdam", "FusedLAMB", "FusedNovoGrad", "FusedSGD", "FusedAdagrad", "DistributedOptimizer"]
