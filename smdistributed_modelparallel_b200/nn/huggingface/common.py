"""Shared machinery for the HuggingFace integrations: a small rule engine that renames / reshapes state-dict
entries in both directions, and helpers for the forward / return hooks.

The reference implements each family by hand (reference smdistributed/modelparallel/torch/nn/huggingface/
{gpt2,gptj,gptneo,gptneox,bert,roberta}.py, ~2.3 kLoC of key-by-key code).  Here a family is a list of
``Rule`` objects (pattern with ``{i}`` layer index -> pattern, optional tensor transform pair), which keeps
the two directions consistent by construction.
"""

from __future__ import annotations

import re
from collections import OrderedDict
from dataclasses import dataclass
from typing import Callable

import torch

from ...backend.exceptions import SMPValidationError


def _t(x):
    return x.t().contiguous()


def _same(x):
    return x


@dataclass
class Rule:
    hf: str                         # e.g. "transformer.h.{i}.ln_1.weight"
    smp: str                        # e.g. "transformer.seq_layers.{i}.attention.pre_layernorm.weight"
    to_smp: Callable = _same
    to_hf: Callable = _same

    def _regex(self, pattern):
        # an arbitrary dotted prefix is carried over, so a replaced submodule (BertModel.encoder, a GPT-2
        # wrapped in a user module) translates the same way as a top-level one
        return re.compile(r"^((?:.*\.)?)" + re.escape(pattern).replace(r"\{i\}", r"(\d+)") + "$")

    def match(self, key, direction):
        m = self._regex(self.hf if direction == "to_smp" else self.smp).match(key)
        if not m:
            return None
        target = self.smp if direction == "to_smp" else self.hf
        if len(m.groups()) > 1:
            target = target.replace("{i}", m.group(2))
        return m.group(1) + target


@dataclass
class FusedQKVRule:
    """One HF tensor holding q, k, v <-> three smp tensors.  ``split`` / ``merge`` get (tensor, config)."""
    hf: str
    smp_prefix: str                 # "...attention" -> .query/.key/.value + suffix
    suffix: str                     # "weight" or "bias"
    split: Callable = None
    merge: Callable = None


class Translator:
    def __init__(self, rules, qkv_rules=(), config=None, drop_hf=(), drop_smp=()):
        self.rules, self.qkv_rules, self.config = list(rules), list(qkv_rules), config
        self.drop_hf = [re.compile(p) for p in drop_hf]
        self.drop_smp = [re.compile(p) for p in drop_smp]

    def hf_to_smp(self, sd):
        out = OrderedDict()
        for k, v in sd.items():
            if any(p.search(k) for p in self.drop_hf):
                continue
            done = False
            for r in self.qkv_rules:
                m = re.match(r"^((?:.*\.)?)" + re.escape(r.hf).replace(r"\{i\}", r"(\d+)") + "$", k)
                if m:
                    pre = m.group(1) + r.smp_prefix.replace("{i}", m.group(2))
                    q, kk, vv = r.split(v, self.config)
                    out[f"{pre}.query.{r.suffix}"], out[f"{pre}.key.{r.suffix}"], out[f"{pre}.value.{r.suffix}"] = q, kk, vv
                    done = True
                    break
            if done:
                continue
            for r in self.rules:
                t = r.match(k, "to_smp")
                if t is not None:
                    out[t] = r.to_smp(v)
                    done = True
                    break
            if not done:
                out[k] = v
        return out

    def smp_to_hf(self, sd):
        out = OrderedDict()
        sd = OrderedDict((k, v) for k, v in sd.items() if not k.startswith("_smp_"))
        used = set()
        for r in self.qkv_rules:
            pat = re.compile(r"^((?:.*\.)?)" + re.escape(r.smp_prefix).replace(r"\{i\}", r"(\d+)")
                             + rf"\.query\.{r.suffix}$")
            for k in list(sd):
                m = pat.match(k)
                if not m:
                    continue
                pre = m.group(1) + r.smp_prefix.replace("{i}", m.group(2))
                keys = [f"{pre}.{n}.{r.suffix}" for n in ("query", "key", "value")]
                out[m.group(1) + r.hf.replace("{i}", m.group(2))] = r.merge(sd[keys[0]], sd[keys[1]], sd[keys[2]], self.config)
                used.update(keys)
        for k, v in sd.items():
            if k in used or any(p.search(k) for p in self.drop_smp):
                continue
            for r in self.rules:
                t = r.match(k, "to_hf")
                if t is not None:
                    out[t] = r.to_hf(v)
                    break
            else:
                out[k] = v
        return out


def strip_prefix(sd, prefix):
    return OrderedDict((k[len(prefix):] if k.startswith(prefix) else k, v) for k, v in sd.items())


def check_unsupported(name, exc, **kwargs):
    """Raise ``exc`` when a forward argument the distributed replacement cannot honour is actually used
    (None, False and empty sequences count as unused)."""
    def used(v):
        if v is None or v is False:
            return False
        if isinstance(v, (list, tuple)) and len(v) == 0:
            return False
        return True

    bad = [k for k, v in kwargs.items() if used(v)]
    if bad:
        raise exc(f"arguments {bad} of HuggingFace {name}.forward are not supported by the tensor-parallel replacement")


def gelu_name(act: str) -> str:
    """HF activation names -> names understood by DistributedTransformerOutputLayer."""
    table = {"gelu_new": "gelu_new", "gelu_fast": "gelu_fast", "gelu_pytorch_tanh": "gelu_pytorch_tanh",
             "gelu": "gelu", "relu": "relu"}
    if act not in table:
        raise SMPValidationError(f"activation {act!r} is not supported by the tensor-parallel transformer")
    return table[act]
