"""ctypes binding of ``libsmpb200_core.so`` (message bus, grad counter, bucket planner, timeline, metrics).

C ABI declared in ``include/smpb200_core.h``; sources in ``csrc/core``.  Mirrors how the reference binds
its ``smplib.so`` (reference smdistributed/modelparallel/backend/core.py:236, ``ctypes.CDLL``).
"""

from __future__ import annotations

import ctypes
import os
import threading
from ctypes import POINTER, c_char_p, c_double, c_int, c_longlong, c_void_p
from pathlib import Path

_LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libsmpb200_core.so"
_lib = None
_lock = threading.Lock()

P, I, L = c_void_p, c_int, c_longlong

_SIG = {
    "smpc_bus_listen": ([], I),
    "smpc_bus_listen_on": ([c_char_p], I),
    "smpc_bus_set_token": ([ctypes.c_ulonglong], I),
    "smpc_bus_connect": ([I, I, c_char_p, I], I),
    "smpc_bus_initialized": ([], I),
    "smpc_bus_rank": ([], I),
    "smpc_bus_size": ([], I),
    "smpc_bus_send": ([c_char_p, L, I, L, I], I),
    "smpc_bus_bcast": ([c_char_p, L, I, POINTER(I), L, I], I),
    "smpc_bus_wait_sends": ([L], I),
    "smpc_bus_poll": ([I, L], L),
    "smpc_bus_wait": ([I, L, I], L),
    "smpc_bus_retrieve": ([I, L, P, L], L),
    "smpc_bus_server_peek": ([POINTER(I), POINTER(L), I], L),
    "smpc_bus_server_pop": ([P, L], L),
    "smpc_bus_peer_failed": ([], I),
    "smpc_bus_shutdown": ([I], I),
    "smpc_gc_create": ([I, I], P),
    "smpc_gc_destroy": ([P], None),
    "smpc_gc_set_microbatch": ([P, I], None),
    "smpc_gc_get_microbatch": ([P], I),
    "smpc_gc_increment_expected": ([P, I, POINTER(I), I], I),
    "smpc_gc_mark_fwd_done": ([P, I], I),
    "smpc_gc_mark_grad_computed": ([P, I], I),
    "smpc_gc_is_ready": ([P, I], I),
    "smpc_gc_expected": ([P, I, I], I),
    "smpc_gc_seen": ([P, I, I], I),
    "smpc_gc_expected_totals": ([P, POINTER(I)], None),
    "smpc_gc_clear": ([P], None),
    "smpc_plan_buckets": ([I, POINTER(L), POINTER(I), L, L, POINTER(I)], I),
    "smpc_timeline_create": ([c_char_p, I], I),
    "smpc_timeline_enabled": ([], I),
    "smpc_timeline_start_step": ([], None),
    "smpc_timeline_end_step": ([], None),
    "smpc_timeline_record_pipeline_event": ([I, c_char_p], None),
    "smpc_timeline_num_events": ([], I),
    "smpc_timeline_flush": ([], I),
    "smpc_timeline_destroy": ([], None),
    "smpc_pool_record": ([L, L, I], None),
    "smpc_get_and_reset_memory_metrics": ([POINTER(c_double)], I),
    "smpc_get_and_reset_alloc_metrics": ([POINTER(L)], I),
    "smpc_device_count": ([], I),
    "smpc_symm_last_error": ([], c_char_p),
    "smpc_symm_alloc": ([L, POINTER(P), ctypes.c_char_p], I),
    "smpc_symm_open": ([ctypes.c_char_p, POINTER(P)], I),
    "smpc_symm_close": ([P], I),
    "smpc_symm_free": ([P], I),
    "smpc_symm_can_access_peer": ([I], I),
    "smpc_symm_selftest": ([L], I),
    "smpc_selftest_runtime": ([ctypes.c_char_p, I], I),
}


def selftest(gpu: bool = False) -> str:
    """Run the library's own self-tests (reference test/backend/d2d_test.py:11, from_blob_tensor_test.py:11 call
    the same kind of in-library checks).  Returns the summary line; raises on a failed check."""
    lib = load()
    buf = ctypes.create_string_buffer(256)
    rc = lib.smpc_selftest_runtime(buf, len(buf))
    if rc != 0:
        raise RuntimeError(f"native runtime self-test: {buf.value.decode()}")
    msg = buf.value.decode()
    if gpu:
        rc = lib.smpc_symm_selftest(1 << 20)
        if rc != 0:
            raise RuntimeError(f"symmetric-memory self-test failed ({rc}): {lib.smpc_symm_last_error().decode()}")
        msg += "; symmetric memory ok"
    return msg

_OPTIONAL: dict = {}


def register_optional(name, argtypes, restype=I):
    _OPTIONAL[name] = (argtypes, restype)
    if _lib is not None and hasattr(_lib, name):
        fn = getattr(_lib, name)
        fn.argtypes, fn.restype = argtypes, restype


def lib_path() -> Path:
    return Path(os.environ.get("SMPB200_CORE_LIB", str(_LIB_PATH)))


def load():
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            path = lib_path()
            if not path.exists():
                raise RuntimeError(f"{path} is missing - run `python -m smdistributed_modelparallel_b200.build_native`")
            lib = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
            for name, (args, res) in list(_SIG.items()) + list(_OPTIONAL.items()):
                if not hasattr(lib, name):
                    if name in _OPTIONAL:
                        continue
                    raise RuntimeError(f"{path} lacks symbol {name}; rebuild the native library")
                fn = getattr(lib, name)
                fn.argtypes, fn.restype = args, res
            _lib = lib
    return _lib
