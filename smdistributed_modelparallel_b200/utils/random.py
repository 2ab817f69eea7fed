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
        """Derive the side stream from ``seed`` without touching the process-wide generators: ``torch.manual_seed``
        would also reseed every CUDA generator, i.e. wipe a per-rank seed the user set before ``smp.init`` and make
        all ranks draw the same dropout masks."""
        self._seed = int(seed)
        self._cpu_state = torch.Generator().manual_seed(self._seed).get_state()
        self._cuda_state = None  # created lazily on first CUDA use

    def _ensure_cuda(self):
        if self._cuda_state is None and torch.cuda.is_available():
            dev = torch.device("cuda", torch.cuda.current_device())
            self._cuda_state = torch.Generator(device=dev).manual_seed(self._seed).get_state()

    @contextmanager
    def consistent_rng_state(self, seed=None):
        if self._cpu_state is None or seed is not None:
            self.set_state(self._seed if seed is None else seed)
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
