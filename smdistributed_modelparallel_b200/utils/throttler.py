"""Bound on the number of collectives in flight (reference smdistributed/modelparallel/torch/
throttler.py:16-63: a ring of CUDA events, at most ``SMP_NCCL_THROTTLE_LIMIT`` = 8 outstanding).

On NVSwitch the limit is about workspace, not links: each in-flight collective pins its input and
output buffers, so the cap bounds transient memory.  CPU/gloo runs are synchronous and skip the ring.
"""

from __future__ import annotations

import os
from collections import deque

import torch


class NCCLThrottler:
    def __init__(self):
        self.set_throttle_limit()
        self._events: deque = deque()

    def set_throttle_limit(self, limit=None):
        self.limit = int(os.environ.get("SMP_NCCL_THROTTLE_LIMIT", 8)) if limit is None else int(limit)

    def run(self, fn, *args, **kwargs):
        if not torch.cuda.is_available() or self.limit <= 0:
            return fn(*args, **kwargs)
        while len(self._events) >= self.limit:
            self._events.popleft().synchronize()
        out = fn(*args, **kwargs)
        ev = torch.cuda.Event()
        ev.record()
        self._events.append(ev)
        return out

    def reset(self):
        self._events.clear()


throttler = NCCLThrottler()
