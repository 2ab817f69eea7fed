"""Import path of the reference's fp16 package; checkpoints pickle their loss scaler under
``smdistributed.modelparallel.torch.fp16.loss_scaler`` (reference torch/optimizers/optimizer.py:185-190)."""
from . import loss_scaler  # noqa: F401
from smdistributed_modelparallel_b200.precision.bit16 import *  # noqa: F401,F403
