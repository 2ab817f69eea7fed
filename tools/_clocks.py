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
