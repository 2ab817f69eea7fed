"""ctypes binding of ``libsmpb200_kernels.so`` (C ABI in ``include/smpb200_kernels.h``).

The library is loaded lazily on first use.  On a machine with a CUDA device a missing or unloadable
library is a hard error (``NativeLibraryError``) - the ops never fall back silently to eager PyTorch
for CUDA tensors.  CPU tensors (unit tests, gloo runs) take the reference implementations in the
individual op modules.

The reference loads its native half the same way (``ctypes.CDLL``; reference
smdistributed/modelparallel/backend/core.py:236).
"""

from __future__ import annotations

import ctypes
import os
import threading
from ctypes import POINTER, c_char_p, c_float, c_int, c_longlong, c_void_p
from pathlib import Path

import torch

_LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libsmpb200_kernels.so"
_lock = threading.Lock()
_lib = None

F32, F16, BF16 = 0, 1, 2
_DTYPE_CODE = {torch.float32: F32, torch.float16: F16, torch.bfloat16: BF16}


class NativeLibraryError(RuntimeError):
    pass


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DTYPE_CODE[dt]
    except KeyError:
        raise TypeError(f"unsupported dtype for native kernels: {dt}") from None


P = c_void_p
I = c_int
L = c_longlong
F = c_float
_MT = [P, I, P, P, P, I]  # stream, n, void*** lists, long long* sizes, int* dtypes, chunk

_SIGNATURES = {
    "smpk_version": ([], I),
    "smpk_device_cc": ([], I),
    "smpk_gemm": ([P, I, I, I, I, I, I, P, I, L, I, P, I, L, I, P, I, L, P, I, P, I, L, I, P, F, F, I, I], I),
    "smpk_layernorm_fwd": ([P, P, P, P, I, P, P, I, P, I, P, P, I, I, F], I),
    "smpk_ln_bwd_grid": ([I, I], I),
    "smpk_layernorm_bwd": ([P, P, I, P, I, P, P, P, I, P, I, P, P, P, I, P, I, I], I),
    "smpk_ln_partial_stats": ([P, P, I, P, I, I], I),
    "smpk_ln_apply_stats": ([P, P, I, P, P, P, P, I, P, I, I, I], I),
    "smpk_ln_bwd_local_sums": ([P, P, I, P, I, P, P, P, I, P, I, I], I),
    "smpk_ln_bwd_finish": ([P, P, I, P, I, P, P, P, I, P, P, I, I, I, I], I),
    "smpk_softmax_fwd": ([P, P, P, I, P, I, F, I, I, I, I], I),
    "smpk_softmax_bwd": ([P, P, P, P, I, I, F, I, I, I], I),
    "smpk_get_batch_per_block": ([I, I, I, I], I),
    "smpk_ce_stats": ([P, P, I, P, L, I, I, L, P, P, P], I),
    "smpk_ce_finish": ([P, P, I, P, L, I, I, L, P, P, P, F, L, P, I], I),
    "smpk_bias_gelu": ([P, P, P, I, P, P, I, L, I, I, I], I),
    "smpk_colsum_scratch_floats": ([I], I),
    "smpk_colsum": ([P, P, I, I, I, L, P, I, I, P], I),
    "smpk_qkv_rotary": ([P, P, P, P, P, I, P, P, I, I, I, I, I, I, I, I, L], I),
    "smpk_head_transpose": ([P, P, P, I, I, I, I, I, I], I),
    "smpk_mt_scale": (_MT + [F, P, P, P], I),
    "smpk_mt_axpby": (_MT + [F, F, I, P, P], I),
    "smpk_mt_adam": (_MT + [I, F, F, F, F, I, I, I, F, F, P, P], I),
    "smpk_mt_sgd": (_MT + [I, F, F, F, F, I, I, I, F, P], I),
    "smpk_mt_adagrad": (_MT + [I, F, F, F, I, F, P], I),
    "smpk_mt_lamb_stage1": (_MT + [F, F, F, I, I, F, I, I, P, F, F, P], I),
    "smpk_mt_lamb_stage2": (_MT + [I, F, F, I, P, P, P], I),
    "smpk_mt_novograd": (_MT + [I, F, F, F, F, I, I, F, I, I, P, F, P], I),
    "smpk_mt_l2norm_scratch_floats": ([], I),
    "smpk_mt_l2norm": (_MT + [P, P, P, P, P], I),
    "smpk_gemm_comm": ([P, I, I, I, I, I, P, I, I, P, I, I, P, I, P, I, P, I, I, F, P, F, I, I, P], I),
    "smpk_flags_signal": ([P, P, I, I, ctypes.c_uint], I),
    "smpk_flags_wait": ([P, P, I, I, ctypes.c_uint], I),
    "smpk_p2p_allgather": ([P, P, P, L, P, P, P, I, I, ctypes.c_uint, I], I),
    "smpk_rs_reduce": ([P, P, P, P, P, P, I, P, I, P, I, L, I, I, I, I, ctypes.c_uint], I),
    "smpk_gemm_set_pair_mode": ([I], I),
    "smpk_attn_fwd": ([P, I, I, I, I, I, I, P, L, L, L, P, L, L, L, P, L, L, L, P, L, L, L, P, F, I, I, P, L, L, L, F,
                       ctypes.c_ulonglong, ctypes.c_ulonglong], I),
    "smpk_attn_bwd": ([P, I, I, I, I, I, I, P, L, L, L, P, L, L, L, P, L, L, L, P, L, L, L, P, L, L, L, P, P,
                       P, L, L, L, P, L, L, L, P, L, L, L, F, I, I, P, L, L, L, F, ctypes.c_ulonglong,
                       ctypes.c_ulonglong], I),
    "smpk_sdp_gather": ([P, P, P, L, L, P, ctypes.c_uint, I, I, I], I),
    "smpk_sdp_reduce_adam": ([P, P, P, P, P, P, P, I, L, L, P, P, P, ctypes.c_uint, I, I, F, F, I, F, F, F, F, F, I, I,
                             I, F, I], I),
}

# entry points added after the first release are optional at load time
_OPTIONAL: dict = {}


def register_optional(name: str, argtypes, restype=I) -> None:
    _OPTIONAL[name] = (argtypes, restype)
    if _lib is not None and hasattr(_lib, name):
        fn = getattr(_lib, name)
        fn.argtypes, fn.restype = argtypes, restype


def lib_path() -> Path:
    return Path(os.environ.get("SMPB200_KERNEL_LIB", str(_LIB_PATH)))


def available() -> bool:
    """True when the shared object exists (it may still need a GPU to run anything)."""
    return lib_path().exists()


def load():
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = lib_path()
        if not path.exists():
            raise NativeLibraryError(
                f"{path} is missing - build it with `python -m smdistributed_modelparallel_b200.build_native` "
                "(or __graft_entry__.build())"
            )
        try:
            lib = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
        except OSError as e:  # pragma: no cover - environment specific
            raise NativeLibraryError(f"cannot load {path}: {e}") from e
        lib.smpk_last_error.argtypes = []
        lib.smpk_last_error.restype = c_char_p
        for name, (argtypes, restype) in list(_SIGNATURES.items()) + list(_OPTIONAL.items()):
            if not hasattr(lib, name):
                if name in _OPTIONAL:
                    continue
                raise NativeLibraryError(f"{path} lacks symbol {name}; rebuild the native library")
            fn = getattr(lib, name)
            fn.argtypes, fn.restype = argtypes, restype
        _lib = lib
    return _lib


def call(name: str, *args) -> None:
    """Invoke an int-returning entry point and raise on a non-zero status."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        msg = lib.smpk_last_error()
        raise RuntimeError(f"{name} failed: {msg.decode() if msg else 'unknown error'}")


def query(name: str, *args) -> int:
    return getattr(load(), name)(*args)


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t) -> int | None:
    return None if t is None else t.data_ptr()


def use_native(*tensors) -> bool:
    """Native kernels run iff the (first non-None) tensor lives on a CUDA device."""
    for t in tensors:
        if t is not None:
            return t.is_cuda
    return False


# launch accounting for bench.py ("gpu_launches"): counts native entry-point invocations
_launch_count = 0


def count_launch(n: int = 1) -> None:
    global _launch_count
    _launch_count += n


def launches() -> int:
    return _launch_count


def reset_launches() -> None:
    global _launch_count
    _launch_count = 0
