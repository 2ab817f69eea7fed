"""Sharded-data-parallel ("ZeRO-2D") engine configuration.

The reference maps its ``sdp_*`` knobs onto a DeepSpeed stage-3 config dict (reference
smdistributed/modelparallel/backend/zero_config.py:13-131 and ds_config_defaults.json:6-27).  This
library ships its own ZeRO-3 engine (``parallel/sdp.py``); the dict built here keeps the same key
names so user-supplied JSON files written for the reference still parse, and feeds that engine.
"""

from __future__ import annotations

import json

from .exceptions import SMPValidationError, UnsupportedShardedConfigError

DEFAULTS = {
    "zero_optimization": {
        "stage": 3,
        "overlap_comm": True,
        "contiguous_gradients": True,
        "reduce_scatter": True,
        "reduce_bucket_size": int(5e8),
        "stage3_prefetch_bucket_size": int(5e8),
        "stage3_param_persistence_threshold": int(1e6),
        "stage3_max_live_parameters": int(1e9),
        "stage3_max_reuse_distance": int(1e9),
        "zero2d_shard_size": 1,
        "zero2d_hierarchical_allgather": True,
    },
    "gradient_clipping": 1.0,
    "zero_allow_untested_optimizer": True,
    "train_micro_batch_size_per_gpu": 1,
    "steps_per_print": 1 << 62,
}


def construct_zero2d_config_dict(cfg, core) -> dict:
    if not cfg.zero2d_enabled():
        return {}
    out = json.loads(json.dumps(DEFAULTS))
    path = cfg._sharded_data_parallelism_config
    if path:
        with open(path) as f:
            user = json.load(f)
        for k, v in user.items():
            if isinstance(v, dict) and isinstance(out.get(k), dict):
                out[k].update(v)
            else:
                out[k] = v
    z = out["zero_optimization"]
    if not path or "zero2d_shard_size" not in (user.get("zero_optimization", {}) if path else {}):
        z["zero2d_shard_size"] = cfg.sharded_data_parallel_degree
    if "sdp_reduce_bucket_size" in cfg._provided or not path:
        z["reduce_bucket_size"] = int(cfg.sdp_reduce_bucket_size)
    if "sdp_param_persistence_threshold" in cfg._provided or not path:
        z["stage3_param_persistence_threshold"] = int(cfg.sdp_param_persistence_threshold)
    if "sdp_max_live_parameters" in cfg._provided or not path:
        z["stage3_max_live_parameters"] = int(cfg.sdp_max_live_parameters)
    if "sdp_hierarchical_allgather" in cfg._provided or not path:
        z["zero2d_hierarchical_allgather"] = bool(cfg.sdp_hierarchical_allgather)
    if "sdp_gradient_clipping" in cfg._provided or not path:
        out["gradient_clipping"] = float(cfg.sdp_gradient_clipping)
    out["bf16"] = {"enabled": bool(cfg.bf16)}
    out["fp16"] = {"enabled": bool(cfg.fp16)}
    validate_zero2d_config(out, core)
    return out


def validate_zero2d_config(d: dict, core) -> None:
    z = d.get("zero_optimization", {})
    if z.get("stage") != 3:
        raise UnsupportedShardedConfigError("sharded data parallelism only supports zero_optimization.stage == 3")
    shard = int(z.get("zero2d_shard_size", 1))
    if shard < 1:
        raise SMPValidationError("zero2d_shard_size must be >= 1")
    world = getattr(core, "_size", None) if core is not None else None
    if world is not None and world % shard != 0:
        raise SMPValidationError(
            f"sharded_data_parallel_degree ({shard}) must divide the number of processes ({world})")
    if z.get("offload_optimizer") or z.get("offload_param"):
        raise UnsupportedShardedConfigError("CPU/NVMe offload of sharded state is not supported")
    if d.get("gradient_clipping", 0.0) < 0:
        raise SMPValidationError("gradient_clipping must be non-negative")
