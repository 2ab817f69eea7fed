"""Config schema: defaults, type / range / dependency errors, aliases and derived values
(reference test/backend/test_config.py:13-154 tier)."""

import pytest

from smdistributed_modelparallel_b200.backend.config import ModelParallelConfig
from smdistributed_modelparallel_b200.backend.exceptions import SMPInvalidArgumentError


def test_defaults():
    c = ModelParallelConfig({})
    assert c.pipeline_parallel_degree == 1 and c.tensor_parallel_degree == 1
    assert c.microbatches == 1 and c.pipeline == "interleaved"
    assert c.placement_strategy == "cluster" and c.optimize == "speed"
    assert c.memory_weight == pytest.approx(0.8)
    assert c.sdp_gradient_clipping == pytest.approx(1.0)
    assert c.sdp_reduce_bucket_size == int(5e8)
    assert c.active_microbatches == c.microbatches
    assert not c.zero2d_enabled()


def test_partitions_alias():
    c = ModelParallelConfig({"partitions": 2, "microbatches": 4})
    assert c.pipeline_parallel_degree == 2
    assert c.active_microbatches == 4


def test_unknown_key_rejected():
    with pytest.raises(SMPInvalidArgumentError):
        ModelParallelConfig({"no_such_option": 1})


@pytest.mark.parametrize("cfg", [
    {"microbatches": 0},
    {"microbatches": "4"},
    {"pipeline": "gpipe"},
    {"memory_weight": 1.5},
    {"placement_strategy": "XYZ"},
    {"optimize": "fast"},
    {"pipeline_parallel_degree": -1},
])
def test_invalid_values(cfg):
    with pytest.raises(SMPInvalidArgumentError):
        ModelParallelConfig(cfg)


def test_dependencies():
    # tensor parallelism needs ddp
    with pytest.raises(SMPInvalidArgumentError):
        ModelParallelConfig({"tensor_parallel_degree": 2})
    # sharded data parallelism excludes optimizer-state sharding
    with pytest.raises(SMPInvalidArgumentError):
        ModelParallelConfig({"sharded_data_parallel_degree": 2, "ddp": True, "shard_optimizer_state": True})
    c = ModelParallelConfig({"tensor_parallel_degree": 2, "ddp": True, "shard_optimizer_state": True})
    assert c.tensor_parallel_degree == 2 and c.shard_optimizer_state
    # fp16 and bf16 exclude each other
    with pytest.raises(SMPInvalidArgumentError):
        ModelParallelConfig({"fp16": True, "bf16": True})


def test_sharded_data_parallel_config():
    c = ModelParallelConfig({"sharded_data_parallel_degree": 4, "ddp": True, "bf16": True})
    assert c.zero2d_enabled()
    with pytest.raises(SMPInvalidArgumentError):
        ModelParallelConfig({"sharded_data_parallel_degree": 4, "ddp": True, "tensor_parallel_degree": 2})
    with pytest.raises(SMPInvalidArgumentError):
        ModelParallelConfig({"sharded_data_parallel_degree": 4, "ddp": True, "pipeline_parallel_degree": 2})


def test_integral_float_accepted_for_int():
    c = ModelParallelConfig({"sharded_data_parallel_degree": 2, "ddp": True, "sdp_reduce_bucket_size": 2e8})
    assert c.sdp_reduce_bucket_size == 200000000 and isinstance(c.sdp_reduce_bucket_size, int)


def test_get_config_dict_round_trip():
    c = ModelParallelConfig({"microbatches": 8, "pipeline_parallel_degree": 2})
    d = c.get_config_dict()
    assert d["microbatches"] == 8 and d["pipeline_parallel_degree"] == 2
    c2 = ModelParallelConfig({k: v for k, v in d.items() if k in ("microbatches", "pipeline_parallel_degree")})
    assert c2.get_config_dict()["microbatches"] == 8


def test_sagemaker_env_overrides_the_dict(monkeypatch):
    monkeypatch.setenv("SM_HP_MP_PARAMETERS", '{"microbatches": 4, "pipeline": "simple"}')
    c = ModelParallelConfig({"microbatches": 2, "pipeline_parallel_degree": 2})
    # active_microbatches defaults to pipeline_parallel_degree + 2 = 4 = microbatches, so "simple" is kept
    assert c.microbatches == 4 and c.pipeline == "simple" and c.pipeline_parallel_degree == 2
    assert c.active_microbatches == 4


def test_alias_conflict_and_formula_defaults():
    with pytest.raises(SMPInvalidArgumentError):
        ModelParallelConfig({"partitions": 2, "pipeline_parallel_degree": 4})
    # active_microbatches defaults to microbatches; a different value forces the interleaved schedule
    c = ModelParallelConfig({"pipeline_parallel_degree": 2, "microbatches": 8, "active_microbatches": 3, "pipeline": "simple"})
    assert c.active_microbatches == 3 and c.pipeline == "interleaved"


@pytest.mark.parametrize("key", ["horovod", "herring"])
def test_foreign_data_parallel_backends_are_refused(key):
    """Horovod / Herring are not carried over; accepting the flag would train without gradient averaging."""
    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.backend.exceptions import SMPUnsupportedError

    with pytest.raises(SMPUnsupportedError, match="ddp"):
        smp.init({key: True})


def test_fused_tp_collectives_key_selects_the_comm_mode(monkeypatch):
    """`fused_tp_collectives` (default true) is what picks the peer-memory kernels over NCCL around the
    tensor-parallel GEMMs; without a CUDA device the mode is 'nccl' whatever the key says."""
    import torch

    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.parallel import tp_ops

    monkeypatch.delenv("SMP_TP_COMM", raising=False)
    on, off = ModelParallelConfig({}), ModelParallelConfig({"fused_tp_collectives": False})
    assert tp_ops.select_comm_mode(on, True) == "fused" and tp_ops.select_comm_mode(off, True) == "nccl"
    assert tp_ops.select_comm_mode(on, False) == "nccl"
    monkeypatch.setenv("SMP_TP_COMM", "nccl")
    assert tp_ops.select_comm_mode(on, True) == "nccl"
    monkeypatch.delenv("SMP_TP_COMM")
    # smp.init installs the selected mode
    tp_ops.set_comm_mode("fused")
    smp.init({"microbatches": 1, "fused_tp_collectives": False})
    assert tp_ops.get_comm_mode() == "nccl"
    smp.init({"microbatches": 1})
    assert tp_ops.get_comm_mode() == ("fused" if torch.cuda.is_available() else "nccl")


def test_sdp_adam_in_backward_is_opt_in():
    """Applying the sharded Adam update inside backward changes when parameters move, so it is off unless asked for."""
    assert ModelParallelConfig({}).sdp_adam_in_backward is False
    assert ModelParallelConfig({"sdp_adam_in_backward": True}).sdp_adam_in_backward is True
