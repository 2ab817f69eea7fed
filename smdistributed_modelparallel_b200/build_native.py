"""In-tree native build for the B200 libraries.

Three shared objects are produced next to this package (``smdistributed_modelparallel_b200/lib``):

* ``libsmpb200_kernels.so`` - every sm_100a CUDA kernel (``csrc/kernels/sm100/*.cu``), plain C ABI.
* ``libsmpb200_core.so``    - the host runtime (``csrc/core/*.cpp``, ``csrc/symm/*.cu``): message bus,
  timeline, grad counter, bucket planner, peer-memory handle exchange.
* ``_smpb200_torch.so``     - the tensor-communication runtime (``csrc/torch_ext/*.cpp``): a PyTorch C++ extension
  over the c10d process-group API (uneven all-to-all / all-gather, point-to-point handles, the bucket reducer),
  compiled with the host compiler against this interpreter's torch headers and imported as a Python module.

The first two are compiled with ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`` (cross-compiles
without a GPU) and loaded with ``ctypes`` - the same arrangement as the reference, which loads its
``smplib.so`` through ``ctypes.CDLL`` (reference smdistributed/modelparallel/backend/core.py:236).
The build is incremental: an object is rebuilt only when its source or a header is newer.
"""

from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
PKG = Path(__file__).resolve().parent
LIB_DIR = PKG / "lib"
BUILD_DIR = ROOT / "build" / "native"

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
          "--expt-relaxed-constexpr", "-I", str(ROOT / "include")]

KERNEL_LIB = LIB_DIR / "libsmpb200_kernels.so"
CORE_LIB = LIB_DIR / "libsmpb200_core.so"


def _newer(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(d).stat().st_mtime > t for d in deps)


def _run(cmd, verbose):
    if verbose:
        print(" ".join(str(c) for c in cmd), flush=True)
    res = subprocess.run([str(c) for c in cmd], capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError(f"native build failed: {' '.join(str(c) for c in cmd)}")
    return res.stdout + res.stderr


def _build_lib(src_dirs, exts, out: Path, extra, verbose, ptxas_v=False):
    srcs = []
    headers = list((ROOT / "include").glob("*.h"))
    for d in src_dirs:
        for ext in exts:
            srcs += sorted(Path(d).glob(f"*{ext}"))
        headers += list(Path(d).glob("*.cuh")) + list(Path(d).glob("*.hpp")) + list(Path(d).glob("*.h"))
    if not srcs:
        return None
    BUILD_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    jobs = []
    objs = []
    for s in srcs:
        o = BUILD_DIR / (out.stem + "__" + s.stem + ".o")
        objs.append(o)
        if _newer(o, [s] + headers):
            cmd = [NVCC, *ARCH_FLAGS, *COMMON, *extra, "-c", s, "-o", o]
            if s.suffix != ".cu":
                cmd.insert(1, "-x")
                cmd.insert(2, "cu")
            if ptxas_v:
                cmd += ["-Xptxas", "-v"]
            jobs.append(cmd)
    logs = []
    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            logs = list(ex.map(lambda c: _run(c, verbose), jobs))
    if jobs or _newer(out, objs):
        _run([NVCC, *ARCH_FLAGS, "-shared", "-o", out, *objs, "-lpthread", "-ldl"], verbose)
    return "\n".join(logs)


TORCH_EXT = LIB_DIR / "_smpb200_torch.so"
CXX = os.environ.get("CXX", "g++")


def _build_torch_ext(verbose):
    """csrc/torch_ext/*.cpp -> lib/_smpb200_torch.so (one translation unit per source, host compiler)."""
    import sysconfig

    import torch
    from torch.utils import cpp_extension as ce

    srcs = sorted((ROOT / "csrc" / "torch_ext").glob("*.cpp"))
    if not srcs:
        return ""
    BUILD_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    stamp = BUILD_DIR / "_smpb200_torch.torch_version"
    built_for = stamp.read_text() if stamp.exists() else ""
    if not _newer(TORCH_EXT, srcs) and built_for == torch.__version__:
        return ""
    libdir = ce.library_paths()[0]
    cmd = [CXX, "-O2", "-std=c++17", "-shared", "-fPIC", "-fvisibility=hidden", "-DTORCH_EXTENSION_NAME=_smpb200_torch",
           "-DTORCH_API_INCLUDE_EXTENSION_H", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"]
    cmd += [f"-I{i}" for i in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}"]
    cmd += [*srcs, "-o", TORCH_EXT, f"-L{libdir}", "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_python",
            f"-Wl,-rpath,{libdir}"]
    log = _run(cmd, verbose)
    stamp.write_text(torch.__version__)
    return log


def build(verbose: bool = False, ptxas_v: bool = False) -> dict:
    """Compile the libraries (incrementally). Returns {"kernels": path, "core": path, "torch_ext": path, "log": str}."""
    log = []
    log.append(_build_torch_ext(verbose))
    log.append(_build_lib([ROOT / "csrc" / "kernels" / "sm100"], [".cu"], KERNEL_LIB, [], verbose, ptxas_v) or "")
    log.append(_build_lib([ROOT / "csrc" / "core", ROOT / "csrc" / "symm"], [".cpp", ".cu"], CORE_LIB, [], verbose,
                          ptxas_v) or "")
    return {"kernels": KERNEL_LIB, "core": CORE_LIB, "torch_ext": TORCH_EXT, "log": "\n".join(log)}


if __name__ == "__main__":
    out = build(verbose="-v" in sys.argv, ptxas_v="--ptxas" in sys.argv)
    if out["log"].strip():
        print(out["log"])
    print("built:", out["kernels"], out["core"], out["torch_ext"])
