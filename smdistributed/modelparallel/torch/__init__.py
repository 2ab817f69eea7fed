"""Drop-in import path of the SageMaker library: scripts written as

    import smdistributed.modelparallel.torch as smp

run unchanged on the B200-native implementation.  Everything is re-exported from
``smdistributed_modelparallel_b200``; the sub-packages users reach through the old path (``smp.nn``,
``smp.amp``, ``smp.optimizers``, ``...torch.nn.huggingface``) are aliased in ``sys.modules``.
"""

import sys as _sys

import smdistributed_modelparallel_b200 as _impl
from smdistributed_modelparallel_b200 import *  # noqa: F401,F403
from smdistributed_modelparallel_b200 import amp, nn, optimizers, state  # noqa: F401

__all__ = list(getattr(_impl, "__all__", []))

for _name in ("nn", "amp", "optimizers", "checkpoint", "comm", "model", "step", "state_mod", "offload", "parameter",
              "collectives"):
    _mod = getattr(_impl, _name, None)
    if _mod is None:
        try:
            _mod = __import__(f"smdistributed_modelparallel_b200.{_name}", fromlist=["_"])
        except ImportError:                                    # pragma: no cover
            continue
    if hasattr(_mod, "__path__") or hasattr(_mod, "__file__"):
        _sys.modules.setdefault(f"{__name__}.{_name}", _mod)


def __getattr__(name):
    return getattr(_impl, name)
