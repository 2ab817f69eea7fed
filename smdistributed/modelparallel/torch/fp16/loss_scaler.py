"""``fp16_states_*.pt`` files hold a pickled ``LossScaler`` / ``DynamicLossScaler`` object whose class is looked
up at this path on load (reference torch/optimizers/optimizer.py:185-190,270-271), so the classes live here
under the reference's names."""
from smdistributed_modelparallel_b200.precision.loss_scaler import DynamicLossScaler, LossScaler, LossScalerBase  # noqa: F401
