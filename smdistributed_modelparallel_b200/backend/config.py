"""Schema-driven configuration (``smp.init(config)``).

Behavioural parity with the reference's ``ModelParallelConfig`` (reference
smdistributed/modelparallel/backend/config.py:42-305): the ``SM_HP_MP_PARAMETERS`` environment JSON
overrides the dict, aliases are folded, unknown keys are rejected, values are type / choice / bound
checked, cross-constraints apply only to non-default values, defaults may be arithmetic expressions of
other keys (clamped into their bounds), and a non-interleaved pipeline is switched to interleaved when
``active_microbatches != microbatches``.

The implementation is independent: the schema (``config_schema.yaml``) is a list of entries resolved
in topological order, and expressions are evaluated on a restricted Python AST.
"""

from __future__ import annotations

import ast
import json
import operator
import os
from graphlib import CycleError, TopologicalSorter
from pathlib import Path

import yaml

from .exceptions import SMPConfigError, SMPInvalidArgumentError
from .logger import get_logger

_SCHEMA_PATH = Path(__file__).with_name("config_schema.yaml")
_KINDS = {"int": int, "float": float, "bool": bool, "str": str, "none": type(None)}
_BINOPS = {ast.Add: operator.add, ast.Sub: operator.sub, ast.Mult: operator.mul, ast.Div: operator.truediv,
           ast.FloorDiv: operator.floordiv}
_schema_cache = None


def load_schema() -> list[dict]:
    global _schema_cache
    if _schema_cache is None:
        with open(_SCHEMA_PATH) as f:
            doc = yaml.safe_load(f)
        entries = doc["parameters"]
        seen = set()
        for e in entries:
            if e["key"] in seen:
                raise SMPConfigError(f"duplicate schema key {e['key']}")
            seen.add(e["key"])
        _schema_cache = entries
    return _schema_cache


def _evaluate(expr, resolved: dict):
    """Literal, or arithmetic over already-resolved keys ("pipeline_parallel_degree + 2")."""
    if not isinstance(expr, str):
        return expr
    if expr in resolved:
        return resolved[expr]
    try:
        tree = ast.parse(expr, mode="eval").body
    except SyntaxError:
        return expr  # a plain string value

    def walk(node):
        if isinstance(node, ast.Constant) and isinstance(node.value, (int, float)):
            return node.value
        if isinstance(node, ast.Name):
            if node.id not in resolved:
                raise KeyError(node.id)
            return resolved[node.id]
        if isinstance(node, ast.BinOp) and type(node.op) in _BINOPS:
            return _BINOPS[type(node.op)](walk(node.left), walk(node.right))
        if isinstance(node, ast.UnaryOp) and isinstance(node.op, ast.USub):
            return -walk(node.operand)
        raise SMPConfigError(f"unsupported expression in config schema: {expr!r}")

    try:
        return walk(tree)
    except KeyError:
        return expr  # refers to nothing we know: treat as a literal string (e.g. "interleaved")


class ConfigEntry:
    """One schema entry bound to the values resolved so far."""

    def __init__(self, spec: dict, resolved: dict):
        self.spec = spec
        self.key = spec["key"]
        self._resolved = resolved

    def _bound(self, which):
        return _evaluate(self.spec[which], self._resolved) if which in self.spec else None

    def default(self):
        if "default" not in self.spec:
            raise SMPInvalidArgumentError(f"Config parameter {self.key} is required.")
        d = _evaluate(self.spec["default"], self._resolved)
        if isinstance(d, (int, float)) and not isinstance(d, bool):
            hi, lo = self._bound("max"), self._bound("min")
            if hi is not None:
                d = min(d, hi)
            if lo is not None:
                d = max(d, lo)
        return d

    def _accepted_types(self):
        kinds = self.spec.get("kind")
        if kinds is None:
            return None
        kinds = kinds if isinstance(kinds, list) else [kinds]
        return [_KINDS[k] for k in kinds]

    def validate(self, value):
        types = self._accepted_types()
        if types is not None and type(value) not in types:
            # a float that is integral is accepted where an int is expected ("5e8" style values)
            if int in types and isinstance(value, float) and value.is_integer():
                value = int(value)
            elif float in types and type(value) is int:
                value = float(value)
            else:
                raise SMPInvalidArgumentError(
                    f"Config parameter {self.key} type needs to be one of {[t.__name__ for t in types]}. "
                    f"Found: {type(value).__name__}.")
        if "choices" in self.spec and value not in self.spec["choices"]:
            raise SMPInvalidArgumentError(
                f"Config parameter {self.key} must be one of {self.spec['choices']}. Found: {value}.")
        lo, hi = self._bound("min"), self._bound("max")
        if lo is not None and value is not None and value < lo:
            raise SMPInvalidArgumentError(f"Config parameter {self.key} ({value}) cannot be less than {lo}.")
        if hi is not None and value is not None and value > hi:
            raise SMPInvalidArgumentError(f"Config parameter {self.key} ({value}) cannot be larger than {hi}.")
        if value != self.default():
            for other, want in self.spec.get("needs", {}).items():
                if self._resolved[other] != want:
                    raise SMPInvalidArgumentError(
                        f"Setting config parameter {self.key} to non-default value {value} requires {other} to be "
                        f"set to {want}. Found: {self._resolved[other]}")
            for other, bad in self.spec.get("forbids", {}).items():
                if self._resolved[other] == bad:
                    raise SMPInvalidArgumentError(
                        f"Setting config parameter {self.key} to non-default value {value} requires {other} to not "
                        f"be {bad}.")
            any_of = self.spec.get("any_of")
            if any_of and not any(self._resolved[o] == w for o, w in any_of.items()):
                raise SMPInvalidArgumentError(
                    f"Setting config parameter {self.key} to non-default value {value} requires one of {any_of}.")
        return value


def _resolution_order(schema):
    ts = TopologicalSorter()
    index = {e["key"]: i for i, e in enumerate(schema)}
    for e in schema:
        ts.add(e["key"], *e.get("after", []))
    try:
        order = list(ts.static_order())
    except CycleError as ex:  # pragma: no cover - schema bug
        raise SMPConfigError(f"cyclic dependencies in config schema: {ex}") from ex
    # stable: among ready keys keep schema order for reproducible dumps
    return sorted(order, key=lambda k: (_depth(k, schema, index), index[k]))


def _depth(key, schema, index, _memo={}):
    if key in _memo:
        return _memo[key]
    deps = schema[index[key]].get("after", [])
    d = 0 if not deps else 1 + max(_depth(x, schema, index) for x in deps)
    _memo[key] = d
    return d


class ModelParallelConfig:
    """Validated configuration; every schema key becomes an attribute."""

    def __init__(self, config: dict | None = None):
        config = dict(config or {})
        env = os.environ.get("SM_HP_MP_PARAMETERS")
        if env:
            config.update(json.loads(env))
        schema = load_schema()
        by_key = {e["key"]: e for e in schema}
        for e in schema:
            alias = e.get("alias")
            if alias and alias in config:
                if e["key"] in config and config[e["key"]] != config[alias]:
                    raise SMPInvalidArgumentError(
                        f"Conflicting values {config[e['key']]} and {config[alias]} are provided for config parameter "
                        f"{e['key']} and its alias {alias}.")
                config[e["key"]] = config.pop(alias)
        for k in config:
            if k not in by_key:
                raise SMPInvalidArgumentError(f"Unrecognized config parameter {k}.")

        resolved: dict = {}
        self._provided = set(config)
        for key in _resolution_order(schema):
            entry = ConfigEntry(by_key[key], resolved)
            resolved[key] = entry.validate(config[key]) if key in config else entry.default()
        self._values = resolved
        for k, v in resolved.items():
            setattr(self, k, v)
        self._input_config = config
        self._fp16_param_init = self.fp16 or self.fp16_params

        log = get_logger()
        if self.active_microbatches != self.microbatches and self.pipeline != "interleaved":
            self.pipeline = self._values["pipeline"] = "interleaved"
            log.info("Simple pipeline is only supported when 'active_microbatches' is equal to 'microbatches'. "
                     "Using interleaved pipeline instead.")
        if self.pipeline_parallel_degree > 1 and self.checkpoint_attentions:
            log.warning("Cannot checkpoint attentions when pipeline-parallel degree is more than 1, disabling it.")
            self.checkpoint_attentions = self._values["checkpoint_attentions"] = False
        self._zero2d_config_dict: dict = {}

    # ---- sharded data parallelism
    def zero2d_enabled(self) -> bool:
        return self.sharded_data_parallel_degree > 1 or bool(self._sharded_data_parallelism_config)

    def zero2d_config_dict(self) -> dict:
        return self._zero2d_config_dict

    def construct_zero2d_config_dict(self, core) -> None:
        from .zero_config import construct_zero2d_config_dict

        self._zero2d_config_dict = construct_zero2d_config_dict(self, core)

    # ---- introspection
    def get_config_dict(self) -> dict:
        return dict(self._values)

    def display_config(self) -> None:
        log = get_logger()
        log.info("Configuration parameters:")
        for e in load_schema():
            if not e.get("internal"):
                log.info(f"  {e['key']}: {self._values[e['key']]}")
        for e in load_schema():
            if e.get("deprecated") and e["key"] in self._provided:
                repl = e.get("replaced_by")
                log.warning(f'WARNING: "{e["key"]}" is a deprecated config key' +
                            (f', please use "{repl}" instead' if repl else ""))
