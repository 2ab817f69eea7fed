"""Loader of the native tensor-communication runtime (``lib/_smpb200_torch.so``, built from ``csrc/torch_ext`` by
``build_native``): uneven all-to-all / all-gather, point-to-point transfers with handles, the bucket reducer - the
tensor half of the reference's ``smplib.so``, here a PyTorch C++ extension over c10d so that
one implementation serves NCCL and gloo.

``load()`` returns the module or raises with the build command; ``SMP_NATIVE_COMM=0`` makes the callers use their
``torch.distributed`` formulation instead (kept for bisecting, not a silent fallback: a missing library is an error).
"""

from __future__ import annotations

import importlib.util
import os
from pathlib import Path

_PATH = Path(__file__).resolve().parent.parent / "lib" / "_smpb200_torch.so"
_mod = None


def enabled(part: str | None = None) -> bool:
    """``SMP_NATIVE_COMM=0`` switches the whole runtime off; ``SMP_NATIVE_COMM_OFF=a2a,reducer,p2p`` single parts."""
    if os.environ.get("SMP_NATIVE_COMM", "1") == "0":
        return False
    return part is None or part not in os.environ.get("SMP_NATIVE_COMM_OFF", "").split(",")


def load():
    global _mod
    if _mod is not None:
        return _mod
    path = Path(os.environ.get("SMPB200_TORCH_EXT", str(_PATH)))
    if not path.exists():
        raise RuntimeError(f"{path} is missing; build it with `python -m smdistributed_modelparallel_b200.build_native`")
    import torch  # noqa: F401  (libtorch must be loaded before the extension)

    spec = importlib.util.spec_from_file_location("_smpb200_torch", str(path))
    mod = importlib.util.module_from_spec(spec)
    try:
        spec.loader.exec_module(mod)
    except ImportError as e:
        raise RuntimeError(f"{path} does not load against torch {torch.__version__} ({e}); rebuild it with "
                           "`python -m smdistributed_modelparallel_b200.build_native`") from e
    _mod = mod
    return mod
