"""Microbatch splitting of ``@smp.step`` arguments and the ``StepOutput`` container.

Behaviour follows the reference (reference smdistributed/modelparallel/backend/split.py:13-228 and
torch/step.py:39-66): every tensor found anywhere inside args/kwargs is narrowed along its split axis
(default 0) unless its argument name is listed in ``non_split_inputs``; ``input_split_axes`` overrides
the axis per argument; objects exposing ``smp_slice(num_mb, mb, axis)`` slice themselves; a
``StepOutput`` passed back into a step function is unpacked per microbatch.
"""

from __future__ import annotations

import inspect
from collections.abc import Mapping

import torch

from .exceptions import InvalidStepOutputError, TensorSplitError
from .logger import get_logger

_QUIET_TYPES = (int, float, bool, str, bytes, type(None), complex)


def map_structure(fn, obj):
    """Apply ``fn`` to every leaf of nested lists / tuples / dicts (namedtuples preserved)."""
    if isinstance(obj, (list, tuple)) and not hasattr(obj, "smp_slice"):
        mapped = [map_structure(fn, o) for o in obj]
        if isinstance(obj, tuple) and hasattr(obj, "_fields"):
            return type(obj)(*mapped)
        return type(obj)(mapped)
    if isinstance(obj, Mapping):
        return type(obj)((k, map_structure(fn, v)) for k, v in obj.items())
    return fn(obj)


class StepOutput:
    """Per-microbatch outputs of one returned value of a step function."""

    def __init__(self, outputs):
        if not isinstance(outputs, (list, tuple)):
            raise InvalidStepOutputError(f"StepOutput only accepts list or tuple, but got {type(outputs)}")
        self._outputs = list(outputs)

    def __getitem__(self, index):
        return self._outputs[index]

    def __len__(self):
        return len(self._outputs)

    @property
    def outputs(self):
        return self._outputs

    def clear(self):
        self._outputs = None

    def add_output(self, out):
        self._outputs.append(out)

    def process_outputs(self, func):
        return func(self._outputs)

    def map(self, func):
        self._outputs = [func(o) for o in self._outputs]
        return self

    def _tensors(self, what):
        if not self._outputs or not all(isinstance(o, torch.Tensor) for o in self._outputs):
            raise InvalidStepOutputError(f"StepOutput.{what} needs every microbatch output to be a tensor")
        return self._outputs

    def reduce_mean(self):
        ts = self._tensors("reduce_mean")
        return torch.stack([t.float() if not t.is_floating_point() else t for t in ts]).mean(dim=0)

    def reduce_sum(self):
        return torch.stack(self._tensors("reduce_sum")).sum(dim=0)

    def concat(self, dim: int = 0):
        ts = self._tensors("concat")
        return torch.cat([t.unsqueeze(0) if t.dim() == 0 else t for t in ts], dim=dim)

    def stack(self, dim: int = 0):
        return torch.stack(self._tensors("stack"), dim=dim)


class TensorSplitter:
    def __init__(self, func, non_split_inputs=None, input_split_axes=None):
        if non_split_inputs is not None and not isinstance(non_split_inputs, list):
            raise TensorSplitError("Non-split inputs must be a list.")
        if input_split_axes is not None and not isinstance(input_split_axes, dict):
            raise TensorSplitError("input_split_axes must be a dict.")
        self.func = func
        self.non_split_inputs = list(non_split_inputs or [])
        self.input_split_axes = dict(input_split_axes or {})
        self._arg_names = inspect.getfullargspec(func).args
        extra = set(self.non_split_inputs) - set(self._arg_names)
        if extra:
            raise TensorSplitError(f"Non-split inputs list {self.non_split_inputs} contains a non-argument of "
                                   f"{getattr(func, '__name__', func)}.")
        extra = set(self.input_split_axes) - set(self._arg_names)
        if extra:
            raise TensorSplitError(f"Input split axes dict {self.input_split_axes} contains a non-argument key of "
                                   f"{getattr(func, '__name__', func)}.")
        self._warned = set()

    # tensor slicing: a view, never a copy
    @staticmethod
    def slice(tensor: torch.Tensor, num_mb: int, mb: int, axis: int = 0) -> torch.Tensor:
        n = tensor.size(axis)
        if n % num_mb != 0:
            raise TensorSplitError(
                f"Dimension {axis} of size {n} of a step-function input is not divisible by the number of "
                f"microbatches ({num_mb}).")
        per = n // num_mb
        return tensor.narrow(axis, mb * per, per)

    def _leaf(self, value, num_mb, mb, splittable, axis):
        if hasattr(value, "smp_slice"):
            return value.smp_slice(num_mb, mb, axis)
        if isinstance(value, StepOutput):
            return value.outputs[mb]
        if not splittable:
            return value
        if isinstance(value, torch.Tensor):
            return self.slice(value, num_mb, mb, axis) if value.dim() > 0 else value
        tname = type(value).__name__
        if tname not in self._warned and not callable(value) and not isinstance(value, _QUIET_TYPES) \
                and not isinstance(value, (torch.nn.Module, torch.optim.Optimizer)):
            get_logger().warning(
                f"Non-splittable object of type {type(value)} passed to smp.step. If it contains tensors that need "
                "to be split across microbatches, implement a 'smp_slice' method for this class.")
            self._warned.add(tname)
        return value

    def preprocess_args(self, args, kwargs, num_mb, mb):
        mb_args = []
        for i, a in enumerate(args):
            name = self._arg_names[i] if i < len(self._arg_names) else None
            splittable = name not in self.non_split_inputs
            axis = self.input_split_axes.get(name, 0)
            mb_args.append(map_structure(lambda x: self._leaf(x, num_mb, mb, splittable, axis), a))
        mb_kwargs = {}
        for k, v in kwargs.items():
            splittable = k not in self.non_split_inputs
            axis = self.input_split_axes.get(k, 0)
            mb_kwargs[k] = map_structure(lambda x: self._leaf(x, num_mb, mb, splittable, axis), v)
        return tuple(mb_args), mb_kwargs

    def preprocess_args_all_mbs(self, args, kwargs, num_mb):
        pairs = [self.preprocess_args(args, kwargs, num_mb, mb) for mb in range(num_mb)]
        return [p[0] for p in pairs], [p[1] for p in pairs]
