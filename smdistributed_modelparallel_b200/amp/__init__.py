"""``smp.amp`` (reference smdistributed/modelparallel/torch/amp/__init__.py)."""
from .scaler import GradScaler

__all__ = ["GradScaler"]
