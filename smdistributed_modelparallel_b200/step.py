"""``@smp.step``: split a minibatch into microbatches, run them through the engine, return StepOutputs.

Reference: smdistributed/modelparallel/torch/step.py:39-357 (StepFunction.__call__ :252-289; StepOutput
:39-50; per-step memory-metrics file :69-115; one-time metrics upload :295-312).  With pp == 1 the
microbatches run back to back on this rank; with pp > 1 the leader / follower loops of the execution
server (``runtime/server.py``) interleave them according to the configured schedule.
"""

from __future__ import annotations

import functools
import os
import time

import torch

from .backend.exceptions import NotInitializedError, StepFunctionCalledError
from .backend.split import StepOutput, TensorSplitter, map_structure
from .backend.utils import upload_metrics_to_studio
from .state_mod import state


class PTTensorSplitter(TensorSplitter):
    pass


class StepMemoryMetricsCollector:
    """Appends per-step memory numbers to ``{fn}_memory_metrics_dp{}_pp{}_tp{}_{ts}`` when
    ``SMP_WRITE_STEP_MEMORY_METRICS=1`` (format of reference step.py:69-115)."""

    def __init__(self, fn_name):
        self.enabled = os.environ.get("SMP_WRITE_STEP_MEMORY_METRICS", "0") == "1"
        self.fn_name = fn_name
        self.path = None
        self.step = 0

    def collect(self):
        if not self.enabled:
            return
        c = state.core
        if self.path is None:
            self.path = f"{self.fn_name}_memory_metrics_dp{c.dp_rank()}_pp{c.pp_rank()}_tp{c.tp_rank()}_{int(time.time())}"
            with open(self.path, "w") as f:
                f.write("step,torch_peak_allocated_mb,torch_peak_reserved_mb,gpu_free_mb,gpu_total_mb,"
                        "d2d_peak_allocated_mb,d2d_peak_reserved_mb\n")
        native = c.get_and_reset_memory_stats()
        if torch.cuda.is_available():
            alloc = torch.cuda.max_memory_allocated() / 2 ** 20
            resv = torch.cuda.max_memory_reserved() / 2 ** 20
            torch.cuda.reset_peak_memory_stats()
        else:
            alloc = resv = 0.0
        with open(self.path, "a") as f:
            f.write(f"{self.step},{alloc:.2f},{resv:.2f},{native['gpu_free_mb']:.2f},{native['gpu_total_mb']:.2f},"
                    f"{native['d2d_peak_allocated_mb']:.2f},{native['d2d_peak_reserved_mb']:.2f}\n")
        self.step += 1


def _detach(obj):
    return map_structure(lambda x: x.detach() if torch.is_tensor(x) else x, obj)


class StepFunction:
    def __init__(self, func, non_split_inputs=None, input_split_axes=None, detach_outputs=True):
        functools.update_wrapper(self, func)
        self.func = func
        self.detach_outputs = detach_outputs
        self.splitter = PTTensorSplitter(func, non_split_inputs, input_split_axes)
        self.id = id(self)
        self.memory = StepMemoryMetricsCollector(getattr(func, "__name__", "step"))

    def __call__(self, *args, **kwargs):
        if not state.initialized:
            raise NotInitializedError()
        model = state.model
        core = state.core
        num_mb = state.num_microbatches()
        state.current_step_func_id = self.id
        core.timeline_start_step()
        ctx = model._step() if model is not None else _null()
        with ctx:
            state.in_step_func = True
            try:
                if core.pp_size() > 1 and model is not None:
                    outputs = state.exec_server.run_step(self, args, kwargs, num_mb)
                else:
                    outputs = []
                    for mb in range(num_mb):
                        state.microbatch = mb
                        core.timeline_record_pipeline_event(mb, "fwd_bwd_start")
                        mb_args, mb_kwargs = self.splitter.preprocess_args(args, kwargs, num_mb, mb)
                        out = self.func(*mb_args, **mb_kwargs)
                        outputs.append(_detach(out) if self.detach_outputs else out)
                        core.timeline_record_pipeline_event(mb, "fwd_bwd_end")
            finally:
                state.in_step_func = False
                state.clear_minibatch_state()
        core.timeline_end_step()
        self.memory.collect()
        if model is not None and not state.metrics_uploaded and model.partitioned:
            state.metrics_uploaded = True
            upload_metrics_to_studio(state.module_manager.get_metrics(core.pp_rank()))
        return self._package(outputs)

    @staticmethod
    def _package(outputs):
        """Transpose [mb][returned value...] into one StepOutput per returned value."""
        if not outputs:
            return None
        first = outputs[0]
        if first is None:
            return None
        if isinstance(first, tuple):
            return tuple(StepOutput([o[i] for o in outputs]) for i in range(len(first)))
        if isinstance(first, list):
            return [StepOutput([o[i] for o in outputs]) for i in range(len(first))]
        return StepOutput(outputs)


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def step(*args, non_split_inputs=None, input_split_axes=None, detach_outputs=True):
    """Decorator: ``@smp.step`` or ``@smp.step(non_split_inputs=[...], input_split_axes={...})``."""
    if len(args) == 1 and callable(args[0]) and non_split_inputs is None and input_split_axes is None:
        return StepFunction(args[0], detach_outputs=detach_outputs)

    def deco(fn):
        return StepFunction(fn, non_split_inputs, input_split_axes, detach_outputs)

    return deco
