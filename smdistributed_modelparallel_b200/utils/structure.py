"""Nested-structure helpers used around the step function and the module server (reference
smdistributed/modelparallel/torch/utils.py:24-166, 213-300).

``flatten_structure`` turns any nesting of list / tuple / namedtuple / set / dict (incl. ``defaultdict`` and
``OrderedDict``) into a flat list of leaves plus a *spec*: the same nesting with every leaf replaced by its
position in the flat list.  The spec holds no tensors, so it can be kept (or pickled to another rank) while the
leaves travel separately; ``unflatten_structure`` rebuilds the original types from any list of the same length.
The reference registers its "phantom structures" in the global state and hands back an id; here the spec is a
plain value owned by the caller.
"""

from __future__ import annotations

import collections
from contextlib import contextmanager

import torch


def is_namedtuple(obj) -> bool:
    return isinstance(obj, tuple) and hasattr(obj, "_fields") and hasattr(obj, "_asdict")


class _Leaf(int):
    """Position of a leaf in the flat list (a distinct type, so integer leaves of the structure stay leaves)."""
    __slots__ = ()

    def __repr__(self):
        return f"#{int(self)}"


def flatten_structure(structure):
    """-> (flat list of leaves, spec)."""
    flat = []

    def walk(x):
        if isinstance(x, dict):
            if isinstance(x, collections.defaultdict):
                out = collections.defaultdict(x.default_factory)
            else:
                out = type(x)()
            for k, v in x.items():
                out[k] = walk(v)
            return out
        if is_namedtuple(x):
            return type(x)(*[walk(v) for v in x])
        if isinstance(x, (list, tuple)):
            return type(x)(walk(v) for v in x)
        if isinstance(x, (set, frozenset)):
            # a set of specs: leaves are distinct positions, so nothing collapses; order is fixed here
            return ("__set__", type(x), tuple(walk(v) for v in x))
        flat.append(x)
        return _Leaf(len(flat) - 1)

    return flat, walk(structure)


def unflatten_structure(flat, spec):
    def build(s):
        if isinstance(s, _Leaf):
            return flat[int(s)]
        if isinstance(s, dict):
            if isinstance(s, collections.defaultdict):
                out = collections.defaultdict(s.default_factory)
            else:
                out = type(s)()
            for k, v in s.items():
                out[k] = build(v)
            return out
        if isinstance(s, tuple) and len(s) == 3 and s[0] == "__set__":
            return s[1](build(v) for v in s[2])
        if is_namedtuple(s):
            return type(s)(*[build(v) for v in s])
        if isinstance(s, (list, tuple)):
            return type(s)(build(v) for v in s)
        raise TypeError(f"not a structure spec: {type(s).__name__}")

    return build(spec)


@contextmanager
def flattened(structure):
    """``with flattened(x) as flat: ...`` - edit ``flat`` in place; ``flat.rebuild()`` gives the structure back."""

    class _Flat(list):
        def rebuild(self_inner):
            return unflatten_structure(self_inner, spec)

    leaves, spec = flatten_structure(structure)
    yield _Flat(leaves)


def flatten_into_args(args, kwargs):
    """(args, kwargs) -> (all positional, keyword order), the form autograd functions need."""
    keys = list(kwargs.keys())
    return tuple(args) + tuple(kwargs[k] for k in keys), keys


def unflatten_from_args(all_args, keyword_order):
    n = len(all_args) - len(keyword_order)
    return tuple(all_args[:n]), {k: v for k, v in zip(keyword_order, all_args[n:])}


def dtype_size(dtype) -> int:
    return torch.empty((), dtype=dtype).element_size()


def get_tensor_size(t) -> int:
    return t.numel() * t.element_size()


def get_size_tensors_in_obj(obj) -> int:
    """Bytes of every tensor reachable in ``obj``."""
    return sum(get_tensor_size(x) for x in flatten_structure(obj)[0] if torch.is_tensor(x))


def get_devices_tensors_in_obj(obj, device_set=None) -> set:
    device_set = set() if device_set is None else device_set
    for x in flatten_structure(obj)[0]:
        if torch.is_tensor(x):
            device_set.add(x.device)
    return device_set


def check_requires_grad(obj) -> bool:
    return any(torch.is_tensor(x) and x.requires_grad for x in flatten_structure(obj)[0])


def convert_args_to_device(args, kwargs, device, only_convert_gpu_tensors=False):
    def move(x):
        if torch.is_tensor(x) and (not only_convert_gpu_tensors or x.is_cuda):
            return x.to(device)
        return x

    flat, spec = flatten_structure((tuple(args), dict(kwargs)))
    return unflatten_structure([move(x) for x in flat], spec)
