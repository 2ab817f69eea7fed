"""Checkpoint API: ``smp.save`` / ``smp.load`` / ``smp.save_checkpoint`` / ``smp.resume_from_checkpoint``.

On-disk layout and key names are the reference's (reference smdistributed/modelparallel/torch/checkpoint.py:
42-121 file-name schemes, :124-165 save/load, :180-297 save_checkpoint, :300-319 sharded-DP files, :381-484
resume, :487-536 config verification).  Model, optimizer and fp16-state files of the pipeline / tensor / data
parallel modes use the reference's keys and pickle the loss scaler under the reference's class path, so they can
be exchanged with it (``tests/test_checkpoint_api.py::test_reference_shaped_fp16_states``); the sharded-data-
parallel optimizer files keep the reference's names but hold this engine's own state (the reference's are
DeepSpeed dictionaries) and are not interchangeable:

    {path}/{tag}_partial/model_{pp}_{tp}.pt                    local state dict of (pp, tp), rdp_rank 0
    {path}/{tag}_partial/optimizer_states_{pp}_{tp}[_{rdp}].pt  (+ _{rdp} with shard_optimizer_state)
    {path}/{tag}_partial/fp16_states_{pp}_{tp}[_{rdp}].pt
    {path}/{tag}_partial/user_content.pt, smp_config.pt
    {path}/newest                                              one "{tag}_partial" line per checkpoint
    {path}/{tag}                                               full (gathered, translated) model
    {path}/{tag}_partial/model_{shard}.pt, optimizer_{shard}.pt  sharded data parallelism
"""

from __future__ import annotations

import glob
import os
import re
import shutil

import torch

from .backend.collectives import CommGroup
from .backend.exceptions import (IncompatibleCheckpointFoundError, IncompatibleCheckpointRankFoundError,
                                 MissingCheckpointFilesError, SMPCheckpointError, SMPValidationError)
from .backend.logger import get_logger, rmsg
from .state_mod import state

SMP_VERSION = "1.12.1"


def _c():
    return state.core


def _torch_load(f, **kwargs):
    kwargs.setdefault("weights_only", False)
    return torch.load(f, **kwargs)


# ------------------------------------------------------------------------------- file-name schemes
def _matching(prefix):
    return [x for x in glob.glob(f"{glob.escape(prefix)}_*") if ".sagemaker-" not in x]


def _scheme(prefix, back_compat=False):
    """3 = {name}_{pp}_{tp}_{rdp}, 2 = {name}_{pp}_{tp}, 1 = {name}_{pp}; validates the file set."""
    files = _matching(prefix)
    if not files:
        raise MissingCheckpointFilesError(prefix)
    end = "$" if back_compat else r"\.pt$"
    pats = {3: re.compile(re.escape(prefix) + r"_(\d+)_(\d+)_(\d+)" + end),
            2: re.compile(re.escape(prefix) + r"_(\d+)_(\d+)" + end),
            1: re.compile(re.escape(prefix) + r"_(\d+)" + end)}
    c = _c()
    n = len(files)
    scheme = None
    for f in files:
        for k in (3, 2, 1):
            m = pats[k].search(f)
            if m:
                scheme = k
                ids = [int(g) for g in m.groups()]
                limits = {1: [("rank", n)],
                          2: [("pp_rank", n // c.tp_size()), ("tp_rank", n // c.pp_size())],
                          3: [("pp_rank", n // (c.tp_size() * c.rdp_size())), ("tp_rank", n // (c.pp_size() * c.rdp_size())),
                              ("rdp_rank", n // (c.tp_size() * c.pp_size()))]}[k]
                for v, (what, lim) in zip(ids, limits):
                    if v >= max(lim, 1):
                        raise IncompatibleCheckpointRankFoundError(f"{what} {v} >= {lim} in {prefix}")
                break
    expected = {1: c.pp_size(), 2: c.mp_size(), 3: c.size()}.get(scheme)
    if scheme is None or n != expected:
        raise IncompatibleCheckpointFoundError(f"found {n} files for {prefix}, expected {expected}: {files}")
    return scheme


def save(obj, f, partial=True, v3=False, **kwargs):
    """``torch.save`` with rank-suffixed file names for partial checkpoints."""
    c = _c()
    if partial:
        f = f"{f}_{c.pp_rank()}_{c.tp_rank()}_{c.rdp_rank()}.pt" if v3 else f"{f}_{c.pp_rank()}_{c.tp_rank()}.pt"
    if c.rank() == 0 or partial:
        torch.save(obj, f, **kwargs)


def load(f, partial=True, back_compat=False, **kwargs):
    c = _c()
    if partial:
        k = _scheme(f, back_compat)
        f = {3: f"{f}_{c.pp_rank()}_{c.tp_rank()}_{c.rdp_rank()}", 2: f"{f}_{c.pp_rank()}_{c.tp_rank()}",
             1: f"{f}_{c.pp_rank()}"}[k]
        if not back_compat:
            f += ".pt"
    return _torch_load(f, **kwargs)


# ------------------------------------------------------------------------------- save_checkpoint
def _check_valid_tag(tag):
    if not isinstance(tag, str):
        raise SMPCheckpointError(f"checkpoint tag can only be str, getting {type(tag)}")
    tags = state.comm.gather(tag, CommGroup.WORLD, 0)
    if _c().rank() == 0 and any(t != tags[0] for t in tags):
        raise SMPCheckpointError(f"checkpoint tag should be same across ranks, getting {tags}")


def save_checkpoint(path, tag, partial=True, model=None, optimizer=None, user_content=None, translate_if_full=True,
                    num_kept_partial_checkpoints=None):
    log = get_logger()
    c, cfg = _c(), state.cfg
    if model is None and optimizer is None:
        log.warning("Both model and optimizer are None, skipping saving checkpoint...")
        return
    if num_kept_partial_checkpoints is not None and num_kept_partial_checkpoints <= 0:
        log.warning(f"Invalid num_kept_partial_checkpoints number {num_kept_partial_checkpoints}, ignoring...")
        num_kept_partial_checkpoints = None
    sdp = cfg.zero2d_enabled()
    if sdp:
        if not partial:
            log.warning("Sharded data parallelism does not support saving a full checkpoint, saving partial instead...")
            partial = True
        if state.optimizer is None:
            raise SMPValidationError("With sharded data parallelism smp.save_checkpoint may be called only after "
                                     "smp.DistributedOptimizer is created")
    kind = "partial" if partial else "full"
    log.info(rmsg(f"Saving {kind} checkpoint with tag {tag} to {path}"))
    _check_valid_tag(tag)
    if partial:
        tag = f"{tag}_{kind}"
        os.makedirs(os.path.join(path, tag), exist_ok=True)
    else:
        os.makedirs(path, exist_ok=True)
    state.comm.barrier()

    if model is not None:
        if sdp:
            model.sdp_engine.save_model_checkpoint(path, tag)
        else:
            _save_model(model, path, tag, partial, translate_if_full)
    if optimizer is not None:
        if sdp:
            model_ = model or state.model
            model_.sdp_engine.save_optimizer_checkpoint(path, tag)
        elif partial:
            _save_optimizer(optimizer, path, tag)
        else:
            log.warning("Saving optimizer for full checkpoint is not supported, skipping...")
    if user_content is not None:
        f = os.path.join(path, tag, "user_content.pt") if partial else os.path.join(path, f"user_content_{tag}")
        save(user_content, f, partial=False)
    if partial:
        conf = cfg.get_config_dict()
        conf["smp_version"] = SMP_VERSION
        save(conf, os.path.join(path, tag, "smp_config.pt"), partial=False)
    if c.rank() == 0 and partial:
        newest = os.path.join(path, "newest")
        lines = open(newest).readlines() if os.path.isfile(newest) else []
        if num_kept_partial_checkpoints is not None:
            while len(lines) >= num_kept_partial_checkpoints:
                old = os.path.join(path, lines[0].rstrip("\n"))
                if os.path.exists(old):
                    shutil.rmtree(old)
                lines = lines[1:]
        lines.append(f"{tag}\n")
        with open(newest, "w") as fd:
            fd.writelines(lines)
    state.comm.barrier()


def _save_model(model, path, tag, partial, translate_if_full, translate_function=None):
    from .model import DistributedModel

    if not isinstance(model, DistributedModel):
        raise SMPValidationError("smp.save_checkpoint can only take smp.DistributedModel as input model")
    c = _c()
    if partial:
        if c.rdp_rank() == 0:
            save(model.local_state_dict(), os.path.join(path, tag, "model"), partial=True)
    else:
        sd = model.state_dict(gather_to_rank0=True)          # collective over the model-parallel group
        if c.rank() == 0:
            if translate_if_full:
                if translate_function is not None:
                    sd = translate_function(sd)
                else:
                    for smp_to_hf, _ in state.tp_registry.active_translate_functions:
                        sd = smp_to_hf(sd)
            else:
                sd["_smp_is_partial"] = False
            save(sd, os.path.join(path, tag), partial=False)
    state.comm.barrier()


def _save_optimizer(optimizer, path, tag):
    if not hasattr(optimizer, "local_optimizer_state_dict"):
        raise SMPValidationError("smp.save_checkpoint can only take smp.DistributedOptimizer as input optimizer")
    c, cfg = _c(), state.cfg
    if cfg.shard_optimizer_state or c.rdp_rank() == 0:
        save(optimizer.local_optimizer_state_dict(), os.path.join(path, tag, "optimizer_states"),
             v3=cfg.shard_optimizer_state)
        if cfg.fp16 or cfg.bf16:
            fp16_states = optimizer.local_fp16_state_dict()
            if fp16_states is not None:
                save(fp16_states, os.path.join(path, tag, "fp16_states"), v3=cfg.shard_optimizer_state)
    state.comm.barrier()


# ------------------------------------------------------------------------------- resume
def resume_from_checkpoint(path, tag=None, partial=True, strict=True, load_optimizer=True,
                           load_sharded_optimizer_state=True, translate_function=None):
    log = get_logger()
    cfg = state.cfg
    sdp = cfg.zero2d_enabled()
    if sdp:
        if state.optimizer is None:
            raise SMPValidationError("With sharded data parallelism smp.resume_from_checkpoint must be called after "
                                     "the optimizer is wrapped with smp.DistributedOptimizer")
        if not partial:
            raise SMPValidationError("With sharded data parallelism only partial checkpoints can be loaded")
    kind = "partial" if partial else "full"
    if tag is None:
        if not partial:
            raise SMPValidationError("tag is required to load a full checkpoint!")
        newest = os.path.join(path, "newest")
        if not os.path.isfile(newest):
            raise SMPValidationError(f"Unable to find newest file at {newest}; pass an explicit tag")
        tag = open(newest).readlines()[-1].rstrip("\n")
        user_tag = tag[: -(len(kind) + 1)]
    else:
        user_tag = tag
        if partial:
            tag = f"{user_tag}_{kind}"
    if partial and kind not in tag:
        raise SMPValidationError(f"The newest checkpoint is not saved with the `{kind}` postfix")
    log.info(f"Resuming from {kind} checkpoint with tag {user_tag} from {path}")
    ckpt = os.path.join(path, tag)
    if not os.path.exists(ckpt):
        raise SMPValidationError(f"Checkpoint {ckpt} does not exist!")
    if partial and _c().rank() == 0 and os.environ.get("SMP_VERIFY_CHECKPOINT_CONFIG", "1") not in ("0", "false"):
        conf_file = os.path.join(ckpt, "smp_config.pt")
        if os.path.isfile(conf_file):
            verify_smp_config(_torch_load(conf_file), partial=partial, load_optimizer=load_optimizer)

    if state.model is None:
        state.loaded_model_state = (_LazyModelCheckpoint(path, tag, partial, translate_function), {"strict": strict})
    elif sdp:
        state.model.sdp_engine.load_model_checkpoint(path, tag, strict)
    else:
        load_model_from_checkpoint(state.model, path, tag, partial, strict, translate_function)
    if load_optimizer:
        if not partial:
            log.warning("Skipping loading optimizer for full checkpoint loading...")
        elif state.optimizer is None:
            state.loaded_optimizer_state = _LazyOptimizerCheckpoint(path, tag)
        elif sdp:
            state.model.sdp_engine.load_optimizer_checkpoint(path, tag, load_sharded_optimizer_state)
        else:
            load_optimizer_from_checkpoint(state.optimizer, path, tag, True, load_sharded_optimizer_state,
                                           tag_is_final=True)
    uc = os.path.join(ckpt, "user_content.pt") if partial else os.path.join(path, f"user_content_{tag}")
    return _torch_load(uc) if os.path.isfile(uc) else None


class _LazyModelCheckpoint(dict):
    """Stand-in state dict resolved when the DistributedModel is finally created."""

    def __init__(self, path, tag, partial, translate_function):
        super().__init__()
        self.args = (path, tag, partial, translate_function)


class _LazyOptimizerCheckpoint(dict):
    def __init__(self, path, tag):
        super().__init__()
        self.args = (path, tag)


def load_model_from_checkpoint(model, path, tag, partial, strict=True, translate_function=None):
    if partial:
        sd = load(os.path.join(path, tag, "model"), partial=True, map_location="cpu")
    else:
        sd = _torch_load(os.path.join(path, tag), map_location="cpu")
    return model.load_state_dict(sd, strict=strict, translate_function=translate_function)


def load_optimizer_from_checkpoint(optimizer, path, tag=None, partial=True, load_sharded_optimizer_state=True,
                                   tag_is_final=False):
    if tag is not None and not tag_is_final and partial and not tag.endswith("_partial"):
        tag = f"{tag}_partial"
    base = os.path.join(path, tag)
    sd = {"optimizer_state_dict": load(os.path.join(base, "optimizer_states"), partial=True, map_location="cpu")}
    try:
        sd["fp16"] = load(os.path.join(base, "fp16_states"), partial=True, map_location="cpu")
    except MissingCheckpointFilesError:
        pass
    optimizer.load_optimizer_checkpoint_state(sd)


def verify_smp_config(saved, partial=True, load_optimizer=True):
    log = get_logger()
    saved = dict(saved)
    ver = saved.pop("smp_version", SMP_VERSION)
    if tuple(int(x) for x in re.findall(r"\d+", ver)[:2]) < (1, 10):
        raise SMPValidationError(f"Checkpoint was saved with smp version {ver} < 1.10.0, which is not supported")
    if ver != SMP_VERSION:
        log.warning(f"Checkpoint was saved with a different smp version ({ver} vs {SMP_VERSION}).")
    cur = state.cfg.get_config_dict()
    mismatch = {k: (v, cur[k]) for k, v in saved.items() if k in cur and cur[k] != v}
    if (load_optimizer and partial) or state.cfg.zero2d_enabled():
        fixed = {"pipeline_parallel_degree", "tensor_parallel_degree", "shard_optimizer_state",
                 "sharded_data_parallel_degree"} & set(mismatch)
        if fixed:
            lines = "".join(f"feature {f}, loaded {mismatch[f][0]}, current {mismatch[f][1]}\n" for f in sorted(fixed))
            raise SMPValidationError(f"The following features changes are not allowed for loading partial: \n{lines}")
    if mismatch:
        lines = "".join(f"feature {f}, loaded {a}, current {b}\n" for f, (a, b) in mismatch.items())
        log.warning(f"Warning: Found mismatched features between save and load: \n{lines}")
