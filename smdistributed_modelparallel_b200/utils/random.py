"""Tensor-parallel-consistent random number streams.

Replicated tensors inside a tensor-parallel group (dropout on all-reduced activations, parameters that
every tp rank initialises identically) must see the same random numbers on every tp rank, while sharded
tensors must not.  As in the reference (reference smdistributed/modelparallel/torch/random.py:8-34), a
second generator state seeded from ``tensor_parallel_seed`` is kept next to the default one and swapped
in for the duration of a ``with`` block.
"""

from __future__ import annotations

from contextlib import contextmanager

import torch


class RngManager:
    def __init__(self):
        self._cpu_state = None
        self._cuda_state = None
        self._seed = 0

    def set_state(self, seed: int):
        self._seed = int(seed)
        cur_cpu = torch.get_rng_state()
        torch.manual_seed(self._seed)
        self._cpu_state = torch.get_rng_state()
        torch.set_rng_state(cur_cpu)
        self._cuda_state = None  # created lazily on first CUDA use

    def _ensure_cuda(self):
        if self._cuda_state is None and torch.cuda.is_available():
            cur = torch.cuda.get_rng_state()
            torch.cuda.manual_seed(self._seed)
            self._cuda_state = torch.cuda.get_rng_state()
            torch.cuda.set_rng_state(cur)

    @contextmanager
    def consistent_rng_state(self, seed=None):
        if self._cpu_state is None:
            self.set_state(self._seed)
        if seed is not None:
            self.set_state(seed)
        use_cuda = torch.cuda.is_available()
        if use_cuda:
            self._ensure_cuda()
        saved_cpu = torch.get_rng_state()
        saved_cuda = torch.cuda.get_rng_state() if use_cuda else None
        torch.set_rng_state(self._cpu_state)
        if use_cuda:
            torch.cuda.set_rng_state(self._cuda_state)
        try:
            yield
        finally:
            self._cpu_state = torch.get_rng_state()
            torch.set_rng_state(saved_cpu)
            if use_cuda:
                self._cuda_state = torch.cuda.get_rng_state()
                torch.cuda.set_rng_state(saved_cuda)

    def get_states(self):
        return {"cpu": self._cpu_state, "cuda": self._cuda_state, "seed": self._seed}

    def set_states(self, st):
        self._cpu_state, self._cuda_state, self._seed = st["cpu"], st["cuda"], st["seed"]
