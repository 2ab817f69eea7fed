"""The allocation-free column-sum entry point, driven through the C ABI exactly as a host program
without torch's allocator would drive it (raw pointers, raw stream handle).

`smpk_colsum_static` keeps its partial-sum scratch in a static device array inside the library, for
callers that have no way to hand the library scratch memory: plain C / ctypes hosts with no workspace
cache, plugins whose surrounding process already owns all device memory, code that must not pay the
device-wide synchronisation of cudaMalloc / cudaFree.  The test checks it against `smpk_colsum` (caller
scratch) on two back-to-back calls with different shapes - the second call reuses the array the first
one dirtied - and that an over-size request is refused with a readable message.

Named test_zz_* so pytest collects it last: with -x every other test has already run if this fails.
"""

import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

P, I, L = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong


def _lib():
    from smdistributed_modelparallel_b200.ops import _native

    lib = _native.load()
    lib.smpk_colsum_static.argtypes = [P, P, I, I, I, L, P, I, I]
    lib.smpk_colsum_static.restype = I
    lib.smpk_colsum_static_max_n.argtypes = []
    lib.smpk_colsum_static_max_n.restype = I
    lib.smpk_colsum.argtypes = [P, P, I, I, I, L, P, I, I, P]
    lib.smpk_colsum.restype = I
    lib.smpk_colsum_scratch_floats.argtypes = [I]
    lib.smpk_colsum_scratch_floats.restype = I
    lib.smpk_last_error.restype = ctypes.c_char_p
    return lib


def _code(dt):
    return {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}[dt]


def _both(lib, x, out_dtype):
    rows, n = x.shape
    stream = torch.cuda.current_stream().cuda_stream
    ref = torch.empty(n, dtype=out_dtype, device=x.device)
    scratch = torch.empty(lib.smpk_colsum_scratch_floats(n), dtype=torch.float32, device=x.device)
    rc = lib.smpk_colsum(stream, x.data_ptr(), _code(x.dtype), rows, n, x.stride(0), ref.data_ptr(), _code(out_dtype), 0,
                         scratch.data_ptr())
    assert rc == 0, lib.smpk_last_error()
    got = torch.full((n,), float("nan"), dtype=out_dtype, device=x.device)
    rc = lib.smpk_colsum_static(stream, x.data_ptr(), _code(x.dtype), rows, n, x.stride(0), got.data_ptr(),
                                _code(out_dtype), 0)
    assert rc == 0, lib.smpk_last_error()
    torch.cuda.synchronize()
    return got, ref


def test_static_scratch_colsum_matches_on_two_consecutive_calls():
    lib = _lib()
    torch.manual_seed(11)
    # call 1: wide, vector path, fills most of the static array
    x1 = torch.randn(3000, 4096, device="cuda", dtype=torch.bfloat16)
    got1, ref1 = _both(lib, x1, torch.float32)
    # call 2 straight after: narrower, odd width (scalar path), fewer rows than partial-sum slots
    x2 = torch.randn(37, 1001, device="cuda", dtype=torch.float32) * 3
    got2, ref2 = _both(lib, x2, torch.float32)
    for got, ref, x in ((got1, ref1, x1), (got2, ref2, x2)):
        assert torch.isfinite(got).all()
        rel = ((got - ref).norm() / (ref.norm() + 1e-12)).item()
        assert rel < 1e-5, f"static-scratch colsum differs from caller-scratch colsum: rel {rel:.3e}"
        # independent oracle: the same sum in float64 on the 16-/32-bit inputs as stored (both kernels accumulate in
        # fp32, so only summation order separates them from it)
        oracle = x.double().sum(0)
        rel = ((got.double() - oracle).norm() / (oracle.norm() + 1e-12)).item()
        assert rel < 1e-5, f"colsum differs from the float64 oracle: rel {rel:.3e}"


def test_static_scratch_colsum_rejects_oversize():
    lib = _lib()
    mx = lib.smpk_colsum_static_max_n()
    assert mx == 8192
    x = torch.zeros(4, mx + 8, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(mx + 8, device="cuda", dtype=torch.float32)
    rc = lib.smpk_colsum_static(torch.cuda.current_stream().cuda_stream, x.data_ptr(), 2, 4, mx + 8, x.stride(0),
                                out.data_ptr(), 0, 0)
    assert rc != 0
    msg = lib.smpk_last_error().decode()
    assert "exceeds" in msg and str(mx) in msg


def test_native_library_selftests_on_gpu():
    """The core library's own checks, including the symmetric-memory round trip on the device (the counterpart of
    the reference's in-library test_d2d(), reference test/backend/d2d_test.py:11)."""
    from smdistributed_modelparallel_b200.backend import native_core

    torch.cuda.init()
    msg = native_core.selftest(gpu=True)
    assert msg.startswith("ok:") and msg.endswith("symmetric memory ok"), msg
