"""Exception taxonomy.

The public names match the reference so user code that catches them keeps working (reference
smdistributed/modelparallel/backend/exceptions.py and torch/exceptions.py, 76 classes).  The classes
are generated from one table: (name, base, default message template).
"""

from __future__ import annotations


class SMPValidationError(ValueError):
    """The user supplied something invalid (configuration, arguments, call order)."""


class SMPConfigError(ValueError):
    """Internal inconsistency in the configuration schema."""


class SMPUnsupportedError(RuntimeError):
    """A valid request that this library does not support."""


class SMPRuntimeError(RuntimeError):
    """Something failed while executing."""


class UnsupportedShardedConfigError(RuntimeError):
    pass


class ScaledBatchBufNotInDistModuleError(Exception):
    pass


_V, _U, _R = SMPValidationError, SMPUnsupportedError, SMPRuntimeError

_TABLE = [
    # ---- validation
    ("CommGroupConfigError", _V, "invalid communication group configuration"),
    ("WorkerSizeError", _V, "world size is not compatible with the requested parallelism degrees"),
    ("LoggingConfigError", _V, "invalid logging configuration"),
    ("InvalidEnvironmentError", _V, "the launch environment is missing rank information"),
    ("NotInitializedError", _V, "smp.init() must be called first"),
    ("InvalidTransactionIDError", _V, "invalid transaction id"),
    ("InvalidLinkIDError", _V, "invalid link id"),
    ("InvalidStepOutputError", _V, "invalid StepOutput operation"),
    ("SMPInvalidArgumentError", _V, "invalid argument"),
    ("DDPNotEnabledError", _V, "this feature needs ddp=True in the smp config"),
    ("InvalidCommGroupError", _V, "unknown communication group"),
    ("DDPConfigError", _V, "invalid DDP configuration"),
    ("HorovodConfigError", _V, "invalid Horovod configuration"),
    ("CheckpointingConfigError", _V, "invalid activation checkpointing configuration"),
    ("DistEmbeddingConfigError", _V, "invalid DistributedEmbedding configuration"),
    ("DistTransformerConfigError", _V, "invalid DistributedTransformer configuration"),
    ("HFBertConfigError", _V, "unsupported HuggingFace BERT configuration"),
    ("HFGPT2ConfigError", _V, "unsupported HuggingFace GPT-2 configuration"),
    ("HFGPTJConfigError", _V, "unsupported HuggingFace GPT-J configuration"),
    ("HFGPTNeoConfigError", _V, "unsupported HuggingFace GPT-Neo configuration"),
    ("HFGPTNeoxConfigError", _V, "unsupported HuggingFace GPT-NeoX configuration"),
    ("HFRobertaConfigError", _V, "unsupported HuggingFace RoBERTa configuration"),
    ("HFT5ConfigError", _V, "unsupported HuggingFace T5 configuration"),
    ("HFViTConfigError", _V, "unsupported HuggingFace ViT configuration"),
    ("ModelTooSmallError", _V, "the model has fewer partitionable modules than pipeline stages"),
    ("InvalidSeqLenPrescaledBatchError", _V, "sequence length must be divisible by the tensor-parallel degree"),
    ("MemoryWeightError", _V, "memory_weight must lie in [0, 1]"),
    ("CostListError", _V, "invalid cost list for partitioning"),
    ("MultipleDistributedModelError", _V, "only one DistributedModel may exist per process"),
    ("BwdExecutionNotInStepFnError", _V, "model.backward must be called inside an @smp.step function"),
    ("FwdExecutionNotInStepFnError", _V, "a DistributedModel must be called inside an @smp.step function"),
    ("ModelNotPartitionedError", _V, "the model has not been partitioned yet"),
    ("HiddenDimError", _V, "hidden dimension mismatch"),
    ("SplitShapeLenError", _V, "split shapes must match the tensor-parallel degree"),
    ("SplitShapeMismatchError", _V, "split shapes do not add up to the full dimension"),
    ("ShiftValueError", _V, "invalid shift value"),
    ("PaddingSizeError", _V, "invalid padding size"),
    ("DistributedModelNotWrappedError", _V, "wrap the model with smp.DistributedModel first"),
    ("DistributedModelWrappedError", _V, "this must happen before wrapping with smp.DistributedModel"),
    ("StepFunctionCalledError", _V, "this must happen before the first @smp.step call"),
    ("InvalidPartitionIDError", _V, "partition id out of range"),
    ("HFNotAvailableError", _V, "HuggingFace transformers is not importable"),
    ("SMPCheckpointError", _V, "checkpoint error"),
    # ---- unsupported
    ("FusedLAMBError", _U, "unsupported FusedLAMB configuration"),
    ("CheckpointingError", _U, "unsupported activation checkpointing use"),
    ("DelayedParamDeviceError", _U, "delayed parameter initialization supports CPU construction only"),
    ("UnsupportedTorchVersionError", _U, "unsupported PyTorch version"),
    ("UnsupportedReducerTypeError", _U, "unsupported reducer type"),
    ("UnsupportedCommunicationVolumeUnitError", _U, "unsupported communication volume unit"),
    ("UnsupportedTPModuleError", _U, "no tensor-parallel counterpart registered for this module"),
    ("TPModuleRegisterError", _U, "invalid tensor-parallel registration"),
    ("MultipleDtypeOptShardingError", _U, "optimizer state sharding needs a single parameter dtype"),
    ("SequentialBackwardBrokenError", _U, "backward through this nn.Sequential split is not supported"),
    ("NotSupportedByFastModeError", _U, "not supported in fast mode"),
    ("RecursionDepthExceededDuringSerializationError", _U, "object nesting too deep to serialise"),
    ("UnsupportedMessageError", _U, "unsupported message type"),
    # ---- runtime
    ("TensorSplitError", _R, "cannot split tensor into microbatches"),
    ("InitializationError", _R, "initialisation failed"),
    ("SMPSegFault", _R, "native library fault"),
    ("InvalidHandleError", _R, "invalid communication handle"),
    ("MissingOutputForModuleError", _R, "no recorded output for module"),
    ("MissingParentModuleError", _R, "module has no recorded parent"),
    ("MissingModuleError", _R, "module is not registered"),
    ("ParentNodeExistingError", _R, "partition-tree node already has a parent"),
    ("ChildNodeExistingError", _R, "partition-tree node already has this child"),
    ("UnrecognizedHFKeyError", _R, "unrecognised HuggingFace state-dict key"),
    ("CustomSoftmaxKernelDtypeError", _R, "the fused softmax kernel needs fp16 or bf16 input"),
    ("MissingGradientError", _R, "expected gradient is missing"),
    ("GradRequireGradError", _R, "gradient tensors must not require grad"),
    ("SMPAMPError", _R, "mixed-precision error"),
    ("PipelineParallelBWDError", _R, "pipeline backward failed"),
    ("InvalidRequestError", _R, "invalid execution request"),
    ("InvalidExecutorError", _R, "request routed to the wrong executor"),
    ("InvalidBwdCountError", _R, "inconsistent backward count"),
    ("NonDummyTensorError", _R, "placeholder tensor was consumed by computation"),
    ("InvalidParentModuleError", _R, "invalid parent module"),
    ("MissingPathFromComputationToModuleOutputError", _R,
     "a module output does not depend on the module's computation (unsupported dataflow)"),
    ("MissingPathFromModuleInputToModuleOutputError", _R,
     "a module input that requires grad does not reach any module output (unsupported dataflow)"),
    ("NumParametersNotMatchError", _R, "parameter count mismatch"),
]


def _make(name, base, default):
    def __init__(self, *args):
        base.__init__(self, *(args if args else (default,)))

    return type(name, (base,), {"__init__": __init__, "__doc__": default, "default_message": default})


for _n, _b, _d in _TABLE:
    globals()[_n] = _make(_n, _b, _d)

_CKPT = globals()["SMPCheckpointError"]
for _n, _d in [
    ("MissingCheckpointFilesError", "checkpoint files are missing"),
    ("IncompatibleCheckpointRankFoundError", "checkpoint was written by an incompatible rank layout"),
    ("IncompatibleCheckpointFoundError", "checkpoint is incompatible with the current configuration"),
    ("MissingKeysInCheckpointError", "checkpoint lacks required keys"),
    ("RemoteBufferShouldNotExistError", "unexpected remote buffer in checkpoint"),
    ("RemoteBufferShouldExistError", "remote buffer missing from checkpoint"),
    ("InvalidReturnTypeFromCheckpointedModuleError", "checkpointed modules must return tensors or tuples of tensors"),
]:
    globals()[_n] = _make(_n, _CKPT, _d)

__all__ = [n for n in list(globals()) if n.endswith("Error") or n == "SMPSegFault"]
