"""Walks over the autograd graph (reference smdistributed/modelparallel/torch/graph_utils.py:14-113).

The pipeline engine cuts the backward pass into local segments at *boundary leaves*: tensors received from
another stage that were made leaves with ``requires_grad`` (runtime/server.py, ``Boundary`` / ``ExecRecord``).
``get_ancestors`` answers, for a set of outputs, which parameters and which such leaves a backward pass from
them will reach - what the reference uses to know how many gradient contributions to expect per parameter.
``unused_parameters`` builds on it (parameters a step cannot reach, whose gradients a reducer must not wait for).
"""

from __future__ import annotations

import torch

from ..backend.logger import get_logger


def _roots(outputs):
    if torch.is_tensor(outputs):
        outputs = [outputs]
    if not isinstance(outputs, (list, tuple)):
        raise TypeError("outputs must be a tensor or a list / tuple of tensors")
    return [t.grad_fn for t in outputs if torch.is_tensor(t) and t.requires_grad and t.grad_fn is not None]


def get_ancestors(outputs, upto_inputs=None):
    """-> (parameters, leaf tensors that are not parameters) reachable from ``outputs``; the walk does not go
    past the tensors in ``upto_inputs``."""
    stop = {t.grad_fn for t in (upto_inputs or ()) if torch.is_tensor(t) and t.grad_fn is not None}
    stack = [n for n in _roots(outputs) if n not in stop]
    seen, params, leaves = set(), set(), set()
    while stack:
        node = stack.pop()
        if node in seen:
            continue
        seen.add(node)
        var = getattr(node, "variable", None)               # AccumulateGrad nodes carry their leaf
        if var is not None:
            (params if isinstance(var, torch.nn.Parameter) else leaves).add(var)
            continue
        for nxt, _ in node.next_functions:
            if nxt is not None and nxt not in stop and nxt not in seen:
                stack.append(nxt)
    return params, leaves


def unused_parameters(module, outputs):
    """Names of the parameters of ``module`` that a backward pass from ``outputs`` does not reach."""
    used, _ = get_ancestors(outputs)
    ids = {id(p) for p in used}
    return [n for n, p in module.named_parameters() if p.requires_grad and id(p) not in ids]


def count_nodes(outputs) -> dict:
    """Histogram of autograd node types reachable from ``outputs``."""
    stack, seen, hist = list(_roots(outputs)), set(), {}
    while stack:
        node = stack.pop()
        if node in seen:
            continue
        seen.add(node)
        hist[type(node).__name__] = hist.get(type(node).__name__, 0) + 1
        stack.extend(n for n, _ in node.next_functions if n is not None)
    return hist


def log_backward_graph(var_grad_fn, level=0, max_depth=50, _out=None):
    """Indented dump of the graph above ``var_grad_fn`` at debug level; returns the lines."""
    top = _out is None
    lines = [] if top else _out
    if var_grad_fn is not None and level <= max_depth:
        var = getattr(var_grad_fn, "variable", None)
        extra = f" -> leaf {tuple(var.shape)}{' (parameter)' if isinstance(var, torch.nn.Parameter) else ''}" \
            if var is not None else ""
        lines.append("  " * level + type(var_grad_fn).__name__ + extra)
        for nxt, _ in var_grad_fn.next_functions:
            log_backward_graph(nxt, level + 1, max_depth, lines)
    if top:
        log = get_logger()
        for ln in lines:
            log.debug(ln)
    return lines
