"""Clock record for the micro-benchmark tools: ``with clocks("label"):`` samples nvidia-smi (SM clock, throttle reasons)
every 100 ms while the block runs and prints one ``clocks[label]: {...}`` line, so every timing in a tool's log can be
read against the clock it was taken at (a roofline fraction without that record says nothing on a power-capped part).
Nothing is changed on the GPU: clocks are read, never set."""

from __future__ import annotations

import contextlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


@contextlib.contextmanager
def clocks(label, gpu_index=None, out=None):
    from bench import ClockSampler

    import torch

    idx = torch.cuda.current_device() if gpu_index is None and torch.cuda.is_available() else (gpu_index or 0)
    visible = os.environ.get("CUDA_VISIBLE_DEVICES")
    if visible:                                   # nvidia-smi counts physical devices
        try:
            idx = int(visible.split(",")[idx])
        except (ValueError, IndexError):
            pass
    s = ClockSampler(idx)
    s.start()
    try:
        yield
    finally:
        rec = s.stop()
        print(f"clocks[{label}]: {json.dumps(rec)}", file=out or sys.stdout, flush=True)


def calibrated_iters(fn, floor=10, target_s=0.25, cap=5000):
    """How many back-to-back calls of ``fn`` fill ``target_s`` seconds (so that a timed loop is long enough for the
    clock sampler to see the load): times three calls after the caller's warm-up."""
    import torch

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(3):
        fn()
    e1.record()
    torch.cuda.synchronize()
    per = max(e0.elapsed_time(e1) / 3e3, 1e-6)
    return int(min(cap, max(floor, target_s / per)))
