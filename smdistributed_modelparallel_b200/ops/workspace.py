"""Grow-only per-device scratch buffers for the native kernels (the kernels themselves never allocate)."""

from __future__ import annotations

import torch

_buffers: dict = {}


def get_workspace(device, tag: str, nbytes: int) -> torch.Tensor:
    """Return a uint8 CUDA buffer of at least ``nbytes`` keyed by (device, stream, tag).

    Keyed by the current stream as well so that kernels running concurrently on side streams never
    share scratch.
    """
    device = torch.device(device)
    key = (device.index if device.index is not None else torch.cuda.current_device(),
           torch.cuda.current_stream(device).cuda_stream, tag)
    buf = _buffers.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        _buffers[key] = buf
    return buf


def release_all() -> None:
    _buffers.clear()
