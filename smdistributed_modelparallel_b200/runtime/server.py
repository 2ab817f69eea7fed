"""Pipeline-parallel execution engine.

Semantics follow the reference's module server (reference smdistributed/modelparallel/torch/server.py:
145-478, worker.py:144-520, patches/execution.py:165-440, ops.py:226-388, sequential.py:94-281): any
``nn.Module`` of the tree may live on any pipeline rank; the step function runs on pp_rank 0 of each
pipeline group (the leader); when execution reaches a module owned by another rank, a request carries
the arguments there, the owner runs the module (recursively dispatching *its* remote children) and
returns the outputs; backward crosses the same boundaries in reverse.  Runs of consecutive
``nn.Sequential`` children on one rank travel as a single request.

What is different (by design):

* **No worker threads.**  A task that must wait (for a forward result, a backward result) pumps the
  message bus re-entrantly: incoming requests are served on the same Python stack, and the leader may
  start the next microbatch while it waits - that is what fills the pipeline.  One thread, no
  condition-variable hand-offs, deterministic order.
* **Segment-wise backward.**  Outputs of a remote call enter the caller's autograd graph as leaves.
  ``model.backward`` runs autograd on the local segment; gradients that reach a boundary leaf are
  shipped to the owner, which back-propagates its own segment and returns the gradients of the call's
  inputs; the caller then continues from those inputs.  Nothing ever blocks inside the autograd
  engine, and gradient contributions from several paths simply accumulate (linearity), which is what
  the reference's GradCounter / dummy-backward bookkeeping (server.py:162-252) works to guarantee.
* ``model.backward`` inside the step function *registers* the backward; the engine schedules it by the
  pipeline policy (``simple``: after all forwards; ``interleaved``: as soon as possible, before new
  forwards) within the ``active_microbatches`` budget.
* Activations and gradients cross ranks with NCCL (gloo on CPU) point-to-point on one dedicated 2-rank
  communicator per direction; the control messages ride the TCP bus of libsmpb200_core.so.
"""

from __future__ import annotations

import contextlib
import os
import threading
import time
import traceback
from dataclasses import dataclass, field

import torch
import torch.distributed as dist
import torch.nn as nn

from ..backend.exceptions import PipelineParallelBWDError, SMPRuntimeError
from ..backend.logger import get_logger
from ..backend.split import map_structure
from .messages import (BackwardRequest, BackwardResult, ForwardRequest, ForwardResult, StepDone, destubify,
                       stubify, stubs_of)
from .pipeline import MbStatus


@dataclass
class ExecRecord:
    """Executor side: what a served forward call left behind for its backward."""
    mb: int
    input_leaves: list            # leaves created from received tensors (requires_grad per stub)
    outputs: list                 # output tensors attached to the local graph
    requester: int
    prev_hop: int | None = None   # fast mode: the chain stage that fed this call (its gradients go there)


@dataclass
class Boundary:
    """Requester side: one remote call made while running microbatch ``mb``."""
    mb: int
    call_id: tuple
    owner: int                    # global rank that executed the call
    out_leaves: list              # leaves standing for the remote outputs in the local graph
    in_tensors: list              # the argument tensors (attached to the local graph) in stub order
    in_requires: list


_HOST_TX_BASE = 1 << 40       # transaction ids of tensors that travel over the host bus (even: not a user-API id)


class _HostRecv:
    """A tensor reception over the host message bus, shaped like a ``dist.Work``."""

    def __init__(self, comm, src, tx, buf):
        self.comm, self.src, self.tx, self.buf, self.done = comm, src, tx, buf, False

    def is_completed(self):
        return self.done or self.comm.poll_recv_from_global(self.src, self.tx)

    def wait(self):
        if self.done:
            return
        raw = self.comm.wait_recv_from_global(self.src, self.tx, depickle=False)
        if self.buf.numel():
            flat = torch.frombuffer(bytearray(raw), dtype=torch.uint8)
            self.buf.copy_(flat.view(self.buf.dtype).reshape(self.buf.shape))
        self.done = True


class _HostSent:
    def __init__(self, comm, tx):
        self.comm, self.tx = comm, tx

    def is_completed(self):
        return True                     # the bus copied the bytes when the send was queued

    def wait(self):
        self.comm.wait_transaction(self.tx)


class _NativeTransfer:
    """A send / receive issued through the native runtime's ``P2PEngine`` (csrc/torch_ext/comm_runtime.cpp: the
    handle API of the reference's tensor transport - send / recv return a handle, poll / wait / clear act on it,
    reference torch/ops.py:34-117), presented with the two methods the handle manager uses."""

    def __init__(self, engine, handle):
        self.engine, self.handle, self.done = engine, handle, False

    def _finish(self):
        self.engine.clear(self.handle)          # releases the engine's reference to the tensor
        self.done = True

    def is_completed(self):
        if not self.done and self.engine.poll(self.handle):
            self._finish()
        return self.done

    def wait(self):
        if not self.done:
            self.engine.wait(self.handle)
            self._finish()


class PipelineP2P:
    """Tensor transport between pipeline ranks: one 2-rank group per ordered (src, dst) pair.

    With ``SMP_DISABLE_D2D=1`` (the reference's switch for its GPU-to-GPU transport, backend/core.py:239-242) or
    for a pair without a process group, tensors fall back to the host message bus: staged through host memory
    and sent as raw bytes under per-pair transaction ids (the role of the reference's MPI tensor fallback,
    test/backend/test_d2d_send_recv.py:63-86)."""

    def __init__(self, state):
        from .links import HandleManager, LinkManager

        self.state = state
        self.me = state.core.rank()
        self.links = LinkManager(state.core.pp_rank())     # one link per transfer slot, recycled every step
        self.handles = HandleManager()                     # keeps in-flight sends (and their tensors) alive
        self.host_only = os.environ.get("SMP_DISABLE_D2D", "0") not in ("0", "", "false")
        self._host_tx = {}                                 # (src, dst) -> tensors moved over the bus so far
        self.host_stats = {"sent": 0, "received": 0, "bytes": 0}
        # static mode (reference backend/config.yaml:191-197, torch/server_queue.py:419-490): the first step of
        # a step function records, per source rank, the sequence of tensor receptions; later steps post those
        # receives ahead of the control message that announces them, so a sender's transfer never waits for this
        # rank's Python to get round to the message
        self._plans = {}            # key -> {src: [(metas key, device)]}
        self._open = False
        self._static = False
        self._recording = False
        self._key = None
        self._seen, self._cursor, self._posted = {}, {}, {}
        self.lookahead = max(1, int(os.environ.get("SMP_STATIC_LOOKAHEAD", "2")))
        self.stats = {"recorded_steps": 0, "replayed_steps": 0, "preposted": 0}
        # A receive is ordered after the compute already enqueued on the current stream (ProcessGroupNCCL semantics),
        # i.e. it starts when its consumer is about to run.  Posting it from a dedicated stream instead lets it start
        # at once - measured on 2 x B200 (GPT-J 6B pp2, profiles/pp2_p2p_variants.log) that is much slower (33.3k vs
        # 44.3k tokens/s): an NCCL receive kernel that waits for its sender spins on the GPU for the whole wait and
        # holds up work of other streams queued behind it on the same hardware connection.  So the side stream is
        # an opt-in (SMP_P2P_SIDE_STREAM=1) for platforms where early posting pays.
        self._pool_step = [0, 0]            # bytes (allocated, reserved) received into device buffers this step
        # transfers go through the native runtime's handle engine on gloo; on NCCL that is an opt-in
        # (SMP_NATIVE_P2P=1) until it has been timed against the torch.distributed calls on GPUs
        from ..backend import native_torch

        mode = os.environ.get("SMP_NATIVE_P2P", "auto")
        on_gloo = dist.is_initialized() and dist.get_backend() == "gloo"
        self._engine = native_torch.load().P2PEngine() if (
            native_torch.enabled("p2p") and (mode == "1" or (mode == "auto" and on_gloo))) else None
        self._recv_stream = None
        dev = getattr(state, "device", None)
        if dev is not None and dev.type == "cuda" and os.environ.get("SMP_P2P_SIDE_STREAM", "0") == "1":
            self._recv_stream = torch.cuda.Stream(device=dev)

    def _wire(self, t):
        t = t.detach()
        if t.dtype == torch.bool:
            t = t.to(torch.uint8)
        return t if t.is_contiguous() else t.contiguous()

    def _next_host_tx(self, src, dst):
        n = self._host_tx.get((src, dst), 0)
        self._host_tx[(src, dst)] = n + 1
        # unique per ordered pair and per tensor; both ends count the same sequence
        return _HOST_TX_BASE + 2 * ((n << 16) | (src & 0xFF) << 8 | (dst & 0xFF))

    def _via_host(self, a, b):
        return self.host_only or (a, b) not in self.state.pp_p2p_groups

    def send(self, tensors, dst):
        host = self._via_host(self.me, dst)
        group = None if host else self.state.pp_p2p_groups[(self.me, dst)]
        nbytes = 0
        for t in tensors:
            w = self._wire(t)
            n = w.numel() * w.element_size()
            if host:
                tx = self._next_host_tx(self.me, dst)
                raw = w.reshape(-1).view(torch.uint8).cpu().numpy().tobytes() if n else b""
                self.state.comm.async_send_global(raw, dst, tx, do_pickle=False)
                work, keep = _HostSent(self.state.comm, tx), None
                self.host_stats["sent"] += 1
                self.host_stats["bytes"] += n
            elif self._engine is not None:
                work = _NativeTransfer(self._engine, self._engine.send(group, w, dist.get_group_rank(group, dst), 0))
                keep = None                                  # the engine holds the tensor until the handle is cleared
            else:
                work, keep = dist.isend(w, dst=dst, group=group), w
            self.handles.register_handle(work, self.handles.SEND, keep, self.links.get_link(n))
            nbytes += n
        self.state.module_manager.record_hop(nbytes)
        self.handles.reap(self.handles.SEND)

    def recv(self, metas, src, device, defer=False):
        """metas: [(shape, dtype)] -> list of tensors.  The receives are posted now (posting order must follow
        message order); the current stream waits for them now, or - with ``defer`` - when the caller later runs
        the returned ``finish()`` (results stashed for a task that resumes later must not stall whatever the
        compute stream is doing meanwhile)."""
        metas = [(tuple(shape), dtype) for shape, dtype in metas]
        if self._static and self._open and not self._recording:
            out, works = self._take_preposted(tuple(metas), src)
        else:
            if self._static and self._open:
                self._seen.setdefault(src, []).append((tuple(metas), device))
            out, works = self._post(metas, src, device)

        def finish():
            for w in works:
                w.wait()                    # NCCL: the current stream waits for the transfer; gloo / host: blocks
            if self._recv_stream is not None:
                cur = torch.cuda.current_stream(device)
                for b in out:
                    b.record_stream(cur)    # allocated on the receive stream, consumed on this one
            return [b.to(torch.bool) if dt == torch.bool else b for b, (_, dt) in zip(out, metas)]

        return finish if defer else finish()

    def _post(self, metas, src, device):
        host = self._via_host(src, self.me)
        group = None if host else self.state.pp_p2p_groups[(src, self.me)]
        out, works = [], []
        side = self._recv_stream if (not host and self._recv_stream is not None) else None
        ctx = torch.cuda.stream(side) if side is not None else contextlib.nullcontext()
        with ctx:
            for shape, dtype in metas:
                wire_dtype = torch.uint8 if dtype == torch.bool else dtype
                buf = self._pool_alloc(shape, wire_dtype, device)
                if host:
                    works.append(_HostRecv(self.state.comm, src, self._next_host_tx(src, self.me), buf))
                    self.host_stats["received"] += 1
                elif self._engine is not None:
                    works.append(_NativeTransfer(self._engine,
                                                 self._engine.recv(group, buf, dist.get_group_rank(group, src), 0)))
                else:
                    works.append(dist.irecv(buf, src=src, group=group))
                out.append(buf)
        return out, works

    def _pool_alloc(self, shape, dtype, device):
        """A receive buffer, counted in the native core's pool metrics (``smp.get_and_reset_memory_stats`` /
        ``get_and_reset_alloc_stats``; the reference counts its receiver-side D2D pool, backend/core.py:296-312).
        Allocated = tensor bytes, reserved = what the caching allocator sets aside for them (512 B granules, 2 MiB
        blocks above 1 MiB).  The buffers of a step are counted as released when the next step begins: received
        activations live until their microbatch's backward, so the peak is that of one whole step."""
        n = 1
        for d in shape:
            n *= int(d)
        n *= torch.empty((), dtype=dtype).element_size()
        gran = (2 << 20) if n > (1 << 20) else 512
        reserved = -(-n // gran) * gran
        try:
            buf = torch.empty(shape, dtype=dtype, device=device)
        except RuntimeError:            # out of memory: count the failure, let the caller see the error
            self.state.core.pool_record(0, reserved, ok=False)
            self.state.core.pool_record(0, -reserved, ok=True)
            raise
        self.state.core.pool_record(n, reserved, ok=True)
        self._pool_step[0] += n
        self._pool_step[1] += reserved
        return buf

    # ------------------------------------------------------------------ static mode
    def begin_step(self, key, schedule_settled=True):
        cfg = self.state.cfg
        if self._pool_step[0] or self._pool_step[1]:
            self.state.core.pool_record(-self._pool_step[0], -self._pool_step[1], ok=True)
            self._pool_step = [0, 0]
        # posting ahead needs the order of receptions from each source to repeat exactly.  It does when stages
        # only answer the leader; with fast-mode chains over three or more stages a stage is fed by two sources and
        # the order of its answers depends on timing, so the dynamic protocol stays in charge there.
        # ... and it does only once the leader's schedule is final: while warm-up depths are still being tried
        # (SMP_SCHEDULE_RECORD_STEPS) every step orders its receptions differently, so nothing is recorded or
        # replayed until the choice is made (the reference, too, registers the receptions of the step it picked,
        # torch/server_queue.py:596-606)
        self._static = bool(getattr(cfg, "static_mode", False)) and schedule_settled and not (
            getattr(cfg, "fast_mode", False) and self.state.core.pp_size() > 2)
        if not self._static:
            return
        if self._open:                      # the previous step never reached drain(): what was learnt is suspect
            self._plans.clear()
        self._open, self._key = True, key
        plan = self._plans.get(key)
        self._recording = plan is None
        self._seen, self._cursor, self._posted = {}, {}, {}
        if plan is not None:
            for src in plan:
                self._prepost(src)

    def _prepost(self, src):
        plan = self._plans[self._key][src]
        posted = self._posted.setdefault(src, [])
        nxt = self._cursor.get(src, 0) + len(posted)
        while len(posted) < self.lookahead and nxt < len(plan):
            metas, device = plan[nxt]
            posted.append(self._post(list(metas), src, device))
            self.stats["preposted"] += 1
            nxt += 1

    def _take_preposted(self, metas, src):
        plan = self._plans[self._key].get(src, [])
        i = self._cursor.get(src, 0)
        if i >= len(plan) or plan[i][0] != metas:
            raise SMPRuntimeError(
                f"static_mode: reception {i} from rank {src} differs from the recorded step "
                f"({metas if i >= len(plan) else (plan[i][0], metas)}); static mode needs identical steps")
        out, works = self._posted[src].pop(0)
        self._cursor[src] = i + 1
        self._prepost(src)
        return out, works

    def drain(self):
        """End of step: every send has left; the links are free for the next step."""
        for h in self.handles.handles():
            self.handles.wait(h)
            self.handles.clear_handle(h)
        self.links.reset()
        if self._static and self._open:
            self._open = False
            if self._recording:
                self._plans[self._key] = self._seen
                self.stats["recorded_steps"] += 1
            else:
                for src, plan in self._plans[self._key].items():
                    if self._cursor.get(src, 0) != len(plan):
                        self._plans.clear()
                        raise SMPRuntimeError(
                            f"static_mode: {self._cursor.get(src, 0)} receptions from rank {src} this step, "
                            f"{len(plan)} in the recorded one; static mode needs identical steps")
                self.stats["replayed_steps"] += 1


def _autocast_state(device_type):
    for dt in dict.fromkeys((device_type, "cuda", "cpu")):
        try:
            if torch.is_autocast_enabled(dt):
                return (dt, torch.get_autocast_dtype(dt))
        except RuntimeError:
            continue
    return None


def _autocast_ctx(state):
    import contextlib

    if not state:
        return contextlib.nullcontext()
    return torch.autocast(state[0], dtype=state[1])


_task_local = threading.local()


class _Task:
    """A microbatch running in its own thread.  Exactly one thread holds the baton at any time: the scheduler
    (main thread) or one task, so the engine stays logically single-threaded; threads only give every
    microbatch its own suspendable Python stack."""

    def __init__(self, sched, mb, phase):
        self.sched, self.mb, self.phase = sched, mb, phase
        self.waiting_key = None
        self.finished = False
        self.error = None
        self.resume_evt = threading.Event()
        self.thread = threading.Thread(target=self._main, name=f"smp-mb{mb}", daemon=True)

    def _main(self):
        srv = self.sched.server
        self.resume_evt.wait()
        self.resume_evt.clear()
        try:
            if srv.state.device.type == "cuda":
                torch.cuda.set_device(srv.state.device)          # device and stream selection are per thread
            _task_local.task = self
            srv._run_microbatch(self.mb, self.phase)
        except BaseException as e:                               # noqa: BLE001 - re-raised on the main thread
            self.error = e
        finally:
            _task_local.task = None
            self.finished = True
            self.sched.baton.set()

    def suspend(self, key):
        self.waiting_key = key
        self.sched.baton.set()                                   # hand the baton back ...
        self.resume_evt.wait()                                   # ... until the scheduler picks this task again
        self.resume_evt.clear()
        self.waiting_key = None


class _Scheduler:
    """Deterministic microbatch scheduling on the pipeline leader.

    Rule: start ``warm`` microbatches, then alternate - resume the task that has been suspended longest
    (blocking until the result it waits for has arrived), start the next microbatch - never exceeding ``window``
    microbatches in flight.  Nothing depends on message timing, so every tensor-parallel rank of a stage issues the same
    sequence of kernels and collectives (the reference gets this from its deterministic server queue, reference
    test/torch/mpi_4ps/test_deterministic.py:17-122), and the host runs ahead of the GPUs: the order in which
    work is *enqueued* is a 1F1B-style interleave, and the streams + NCCL p2p dependencies pipeline it."""

    def __init__(self, server, next_job, window, warm=None):
        # ``next_job()`` -> (microbatch, phase) to start now, or None when nothing can start yet (the policy object
        # state.pipeline decides: runtime/pipeline.py, reference torch/pipeline.py:9-145).
        self.server, self.next_job, self.window = server, next_job, window
        # ``warm``: microbatches started back to back before the first resume; after that a new microbatch is
        # started only right after a resume (1F1B-style alternation).  Filling the whole window first would
        # park the first microbatch's loss / backward request behind every later forward on the stream and
        # leave the later stages idle (simulated for pp2, 8 microbatches: 84 % -> 93 % of the no-bubble time).
        self.warm = window if warm is None else max(1, min(window, warm))
        self.baton = threading.Event()

    def _switch_to(self, task):
        self.baton.clear()
        task.resume_evt.set()
        self.baton.wait()
        if task.error is not None:
            raise task.error

    def run(self):
        srv = self.server
        waiting = []                      # suspended tasks, oldest first
        resumed_since_start = 0
        while True:
            job = None
            if len(waiting) < self.window and (len(waiting) < self.warm or resumed_since_start >= 1 or not waiting):
                job = self.next_job()
            if job is not None:
                task = _Task(self, *job)
                resumed_since_start = 0
                task.thread.start()
            elif waiting:
                task = waiting.pop(0)
                resumed_since_start += 1
                while task.waiting_key not in srv.stash:
                    srv._pump_one(50)
            else:
                break
            self._switch_to(task)
            if not task.finished:
                waiting.append(task)
            else:
                task.thread.join()


class _ScheduleTuner:
    """Record-and-pick-fastest for the leader's schedule (the reference's deterministic queue records the task order of
    its first ``record_step`` = 5 steps, each timed with a device synchronisation on rank 0, then replays the order of
    the fastest one for the rest of the job: torch/server_queue.py:226-299,575-626).  The order here is a function of
    two numbers - the window (``active_microbatches``, fixed by the user: it bounds memory) and ``warm``, how many
    microbatches are started before the first suspended one is resumed - so what is recorded is the step time under
    each candidate ``warm``: step 0 (partitioning, allocator warm-up) is not timed, steps 1..record_steps cycle
    through the candidates, every leader tries the same candidate at the same step, and rank 0's fastest is handed to
    every leader.  Off unless ``SMP_SCHEDULE_RECORD_STEPS`` > 0: the timed steps synchronise the device twice."""

    def __init__(self, candidates, record_steps):
        self.candidates = list(candidates)
        self.record_steps = int(record_steps)
        self.times = {c: [] for c in self.candidates}
        self.step = 0
        self.choice = None

    @property
    def recording(self):
        return self.choice is None and 1 <= self.step <= self.record_steps

    def candidate(self):
        if self.choice is not None:
            return self.choice
        if self.step == 0:
            return self.candidates[0]
        return self.candidates[(self.step - 1) % len(self.candidates)]

    def record(self, seconds):
        if self.recording:
            self.times[self.candidate()].append(float(seconds))
        self.step += 1

    @property
    def must_decide(self):
        return self.choice is None and self.step > self.record_steps

    def best(self):
        timed = {c: min(v) for c, v in self.times.items() if v}
        return min(timed, key=timed.get) if timed else self.candidates[0]


class ExecutionServer:
    def __init__(self, model):
        from ..state_mod import state

        self.state = state
        self.model = model
        self.core = state.core
        self.comm = state.comm
        self.me = self.core.rank()
        self.pp_rank = self.core.pp_rank()
        self.is_leader = self.pp_rank == 0
        self.p2p = PipelineP2P(state)
        self.partitioned = False
        self.records: dict = {}            # call_id -> ExecRecord (executor side)
        self.boundaries: dict = {}         # mb -> [Boundary] (requester side)
        self.stash: dict = {}              # ("F"|"B", call_id) -> (msg, tensors)
        self.deferred_bwd: dict = {}       # mb -> (tensors, grads)
        self._counter = 0
        self._step = 0
        self._step_ctx = None
        self._done_msg = None
        self._orig_forward = {}
        self._early = []                   # control messages that arrived while waiting for the partition
        self._tuners = {}                  # (step function id, grad enabled) -> _ScheduleTuner
        self._steps_of = {}                # (step function id, grad enabled) -> steps run so far (every rank)
        self.record_steps = int(os.environ.get("SMP_SCHEDULE_RECORD_STEPS", "0"))

    # ================================================================== partitioning (first step)
    def _trace(self, step_fn, args, kwargs, num_mb):
        """First-step tracing on the leaders (reference torch/patches/tracing.py:41-86, worker.py:248-278): run the
        step function on the first microbatch without gradients, wherever the unpartitioned model lives, recording
        per module the execution order, device time, allocator growth and bytes of tensors produced.  One warm-up
        pass and ``SMP_TRACE_PASSES - 1`` (default 2) timed passes, averaged (the reference runs five).  The
        partitioner weighs ``memory_weight * memory + (1 - memory_weight) * time`` with these.

        Parameters that are not materialised yet (delayed initialisation) borrow storage for the duration of the
        trace: the values do not matter for timing, the storage is returned before partitioning.  Skipped when
        tracing is off, nothing was requested, or the borrowed storage would exceed ``SMP_TRACE_MAX_BYTES``
        (default 60 GB; the reference skips tracing above 120 GB, worker.py:45,257-263)."""
        import time as _time

        st = self.state
        model = self.model
        want_t, want_m = model.trace_execution_times, model.trace_memory_usage
        if st.cfg.skip_tracing or not (want_t or want_m) or not self.is_leader:
            return None, None
        root = model.module
        dev = st.device
        borrowed = []
        meta_bytes = sum(p.numel() * p.element_size() for p in root.parameters() if p.is_meta)
        if meta_bytes:
            # trace_device="cpu" (reference model.py:111-124): the traced passes run on the host - for models whose
            # unpartitioned parameters do not fit one GPU; the budget is then the host's (SMP_TRACE_MAX_BYTES_CPU)
            on_cpu = getattr(model, "trace_device", "gpu") == "cpu"
            if on_cpu:
                dev = torch.device("cpu")
            limit = float(os.environ.get("SMP_TRACE_MAX_BYTES_CPU", 400e9)) if on_cpu else \
                float(os.environ.get("SMP_TRACE_MAX_BYTES", 60e9))
            if meta_bytes > limit:          # a size rule, so that every leader takes the same decision
                get_logger().info(f"tracing skipped: {meta_bytes / 1e9:.1f} GB of parameters are not materialised "
                                  f"(limit {limit / 1e9:.0f} GB)")
                return None, None
            borrowed = self._borrow_storage(root, dev)
        else:
            dev = next((p.device for p in root.parameters()), torch.device("cpu"))
        cuda = dev.type == "cuda"
        mm = st.module_manager
        trace = mm.trace
        passes = max(2, int(os.environ.get("SMP_TRACE_PASSES", "3")))
        started = {}
        hooks = []
        state_ = {"timed": False}

        def tensor_bytes(obj):
            total = [0]
            map_structure(lambda t: total.__setitem__(0, total[0] + t.numel() * t.element_size())
                          if torch.is_tensor(t) else None, obj)
            return total[0]

        def pre(mod, a):
            name = mm.get_module_name(mod)
            if name not in trace.order:
                trace.order.append(name)
            if cuda:
                torch.cuda.synchronize(dev)
            started[mod] = (_time.perf_counter(), torch.cuda.memory_allocated(dev) if cuda else 0)
            if state_["timed"] == 1:
                trace.in_bytes[name] = trace.in_bytes.get(name, 0) + tensor_bytes(a)

        def post(mod, a, out):
            if cuda:
                torch.cuda.synchronize(dev)
            if not state_["timed"]:
                started.pop(mod, None)
                return
            name = mm.get_module_name(mod)
            t0, m0 = started.pop(mod, (_time.perf_counter(), 0))
            trace.times[name] = trace.times.get(name, 0.0) + (_time.perf_counter() - t0) / (passes - 1)
            if state_["timed"] == 1:
                nbytes = tensor_bytes(out)
                grown = max(0, (torch.cuda.memory_allocated(dev) if cuda else 0) - m0)
                trace.out_bytes[name] = trace.out_bytes.get(name, 0) + nbytes
                trace.memory[name] = trace.memory.get(name, 0) + max(nbytes, grown)

        for m in root.modules():
            hooks.append(m.register_forward_pre_hook(pre))
            hooks.append(m.register_forward_hook(post))
        a, k = step_fn.splitter.preprocess_args(args, kwargs, num_mb, 0)
        a = map_structure(lambda t: t.to(dev) if torch.is_tensor(t) else t, a)
        k = map_structure(lambda t: t.to(dev) if torch.is_tensor(t) else t, k)
        st.is_tracing_flag = True
        prev_mb = st.microbatch
        st.microbatch = 0
        failed = None
        try:
            with torch.no_grad():
                for i in range(passes):
                    state_["timed"] = i              # 0: warm-up; 1: first timed pass (also sizes); 2..: timing only
                    step_fn.func(*a, **k)
        except Exception as e:                       # noqa: BLE001 - tracing is an aid: partition untraced instead
            failed = e
        finally:
            st.is_tracing_flag = False
            st.microbatch = prev_mb
            for h in hooks:
                h.remove()
            self._return_storage(borrowed)
        if failed is not None:
            get_logger().warning(f"first-step tracing failed ({type(failed).__name__}: {failed}); the model is "
                                 "partitioned from the untraced estimates")
            trace.times.clear()
            trace.memory.clear()
            return None, None
        times = {mm.get_module(n): t for n, t in trace.times.items()} if want_t else None
        memory = {mm.get_module(n): float(b) for n, b in trace.memory.items()} if want_m else None
        return times, memory

    @staticmethod
    def _borrow_storage(root, dev):
        """Give every meta parameter / buffer real (small random) storage on ``dev``; returns what to undo."""
        undo = []
        seen = {}
        for m in root.modules():
            for name, p in list(m._parameters.items()):
                if p is None or not p.is_meta:
                    continue
                if id(p) not in seen:
                    t = torch.empty(p.shape, dtype=p.dtype, device=dev)
                    if t.is_floating_point():
                        t.normal_(0.0, 0.02) if t.dim() > 1 else t.fill_(0.5)
                    new = nn.Parameter(t, requires_grad=False)
                    for tag in ("_smp_dist_axis", "_smp_scaled_batch", "_smp_one_rank"):
                        if hasattr(p, tag):
                            setattr(new, tag, getattr(p, tag))
                    seen[id(p)] = new
                undo.append((m._parameters, name, p))
                m._parameters[name] = seen[id(p)]
            for name, b in list(m._buffers.items()):
                if b is None or not b.is_meta:
                    continue
                undo.append((m._buffers, name, b))
                m._buffers[name] = torch.zeros(b.shape, dtype=b.dtype, device=dev)
        return undo

    @staticmethod
    def _return_storage(undo):
        for table, name, original in undo:
            table[name] = original
        if undo and torch.cuda.is_available():
            torch.cuda.empty_cache()

    def _partition(self, step_fn=None, args=None, kwargs=None, num_mb=1):
        from ..pipeline.partition import partition_model

        st = self.state
        mm = st.module_manager
        times = memory = None
        saved = mm.loaded_partition()
        path = st.cfg.partition_file or "./partition.data"
        if saved is None and st.cfg.load_partition:
            # config keys load_partition / partition_file (the reference reads them in its TensorFlow partitioner,
            # tensorflow/auto.py:181-233; here they do the same job for the module tree)
            import pickle
            with open(path, "rb") as f:
                saved = {str(k): int(v) for k, v in pickle.load(f).items()}
        if saved is not None:
            assignment = mm.assignment_from_saved(saved)
        else:
            if step_fn is not None and st.cfg.auto_partition:
                times, memory = self._trace(step_fn, args, kwargs, num_mb)
            assignment = partition_model(self.model.module, mm, st.cfg, module_times=times, module_memory=memory)
        # one assignment for the whole job: measured times differ from leader to leader, and the tensor-parallel
        # and data-parallel peers of a stage must all cut the model in the same place, so rank 0's copy is
        # authoritative for everyone (the reference also partitions on rank 0 and broadcasts, torch/server.py:254-268)
        if self.core.size() > 1 and self.core.pp_size() > 1:
            if self.me == 0:
                for r in range(self.core.size()):
                    if r != self.me:
                        self.comm.server_send(("partition", assignment), r)
            else:
                while True:
                    src, tx, obj = self.comm.server_next_message(-1)
                    if isinstance(obj, tuple) and obj and obj[0] == "partition":
                        assignment = obj[1]
                        break
                    self._early.append((src, tx, obj))      # a stage that already has its copy may be ahead of us
        mm.assign_partitions(assignment)
        if st.cfg.partition_file and not st.cfg.load_partition and self.me == 0:
            import pickle
            with open(st.cfg.partition_file, "wb") as f:
                pickle.dump(dict(assignment), f)
        self.model.post_partition()
        self._install_dispatchers()
        self.partitioned = True
        if self.core.rank() == 0:
            by_rank = {}
            for n, r in assignment.items():
                by_rank.setdefault(r, 0)
                by_rank[r] += 1
            get_logger().info(f"pipeline partition: modules per rank {dict(sorted(by_rank.items()))}")

    def _owner_rank(self, module) -> int:
        return self.core.pp_rank_to_rank(self.state.module_manager.get_partition(module))

    def _install_dispatchers(self):
        mm = self.state.module_manager
        root = self.model.module
        part = mm.get_partition

        def mixed_sequential(m):
            if not isinstance(m, nn.Sequential):
                return False
            ranks = {part(c) for c in m.children()}
            return len(ranks) > 1 or (ranks and ranks != {part(m)})

        seqs = {m for m in root.modules() if mixed_sequential(m)}
        for m in seqs:
            self._orig_forward[m] = m.forward
            m.forward = self._make_sequential_forward(m)
        for parent in root.modules():
            if parent in seqs:
                continue            # children of a mixed Sequential are dispatched chunk-wise by its forward
            for child in parent.children():
                if child in seqs or child in self._orig_forward:
                    continue
                if part(child) != part(parent):
                    self._orig_forward[child] = child.forward
                    child.forward = self._make_dispatch_forward(child)

    def _make_dispatch_forward(self, module):
        orig = self._orig_forward[module]
        name = self.state.module_manager.get_module_name(module)

        def forward(*args, **kwargs):
            if self.state.module_manager.get_partition(module) == self.pp_rank:
                return orig(*args, **kwargs)
            return self._remote_forward(name, None, self._owner_rank(module), args, kwargs)

        return forward

    def _make_sequential_forward(self, seq):
        return lambda inp: self.run_sequential(seq, inp)

    def _span_ckpt(self, seq, ckpt):
        """(strategy, preserve_rng_state, pack_args_as_tuple) in force for a run of ``seq``'s layers, or None."""
        if ckpt is not None:
            return tuple(ckpt)
        cfg = self.state.module_manager.get_checkpoint_config(seq)
        if cfg is not None and cfg.enabled:
            return (cfg.strategy, cfg.preserve_rng_state, cfg.pack_args_as_tuple)
        return None

    @staticmethod
    def _run_layers(layers, x, ckpt):
        if ckpt is not None and torch.is_grad_enabled():
            from .checkpointing import checkpoint_layers
            return checkpoint_layers(layers, x, *ckpt)
        for c in layers:
            x = c(x)
        return x

    def run_sequential(self, seq, inp, ckpt=None):
        """Forward of an ``nn.Sequential`` whose children live on several pipeline ranks: consecutive children
        of one rank run as one chunk, here or as one remote call; activation checkpointing (configured for the
        module, or asked for through ``smp.checkpoint_sequential``) applies to every chunk where it runs."""
        name = self.state.module_manager.get_module_name(seq)
        part = self.state.module_manager.get_partition
        ckpt = self._span_ckpt(seq, ckpt)
        kids = list(seq.children())
        i, n = 0, len(kids)
        while i < n:
            r = part(kids[i])
            j = i
            while j < n and part(kids[j]) == r:
                j += 1
            if r == self.pp_rank:
                inp = self._run_layers(kids[i:j], inp, ckpt)
            else:
                # fast mode (reference backend/config.yaml:183-190, torch/serialization.py:365-425): when
                # the chunks that follow live on further ranks, the output of one goes straight to the
                # next instead of child -> parent -> child; only the last one answers
                chain, seen, first_end = [], {r}, j
                while self.state.cfg.fast_mode and j < n:
                    r2 = part(kids[j])
                    if r2 == self.pp_rank or r2 in seen:
                        break
                    k = j
                    while k < n and part(kids[k]) == r2:
                        k += 1
                    chain.append((self.core.pp_rank_to_rank(r2), (j, k)))
                    seen.add(r2)
                    j = k
                inp = self._remote_forward(name, (i, first_end), self.core.pp_rank_to_rank(r), (inp,), {},
                                           tuple(chain), ckpt)
            i = j
        return inp

    def _run_local(self, module_name, span, args, kwargs, ckpt=None):
        mm = self.state.module_manager
        m = mm.get_module(module_name)
        if span is not None:
            return self._run_layers(list(m.children())[span[0]:span[1]], args[0], ckpt)
        return m(*args, **kwargs)     # the dispatcher sees partition == this rank and runs the real forward

    # ================================================================== message pump
    def _new_call_id(self, mb):
        self._counter += 1
        return (self.me, mb, self._counter)

    def _pump_one(self, timeout_ms) -> bool:
        """Handle one incoming control message; False if none arrived within the timeout."""
        got = self._early.pop(0) if self._early else self.comm.server_next_message(timeout_ms)
        if got is None:
            return False
        src, _tx, msg = got
        self._handle(src, msg)
        return True

    def _handle(self, src, msg):
        dev = self.state.device
        if isinstance(msg, ForwardRequest):
            metas = [(s.shape, s.dtype) for s in stubs_of(msg.skeleton)]
            try:
                tensors = self.p2p.recv(metas, src, dev)
            except Exception:
                self.comm.server_send(ForwardResult(msg.mb, msg.call_id, None, traceback.format_exc()), msg.requester)
                raise
            self._serve_forward(src, msg, tensors)
        elif isinstance(msg, ForwardResult):
            metas = [(s.shape, s.dtype) for s in stubs_of(msg.skeleton)] if msg.error is None else []
            self.stash[("F", msg.call_id)] = (msg, self.p2p.recv(metas, src, dev, defer=True))
        elif isinstance(msg, BackwardRequest):
            grads = self.p2p.recv([(shape, dt) for _, shape, dt in msg.grad_meta], src, dev)
            self._serve_backward(src, msg, grads)
        elif isinstance(msg, BackwardResult):
            grads = self.p2p.recv([(shape, dt) for _, shape, dt in msg.grad_meta], src, dev, defer=True) \
                if msg.error is None else (lambda: [])
            self.stash[("B", msg.call_id)] = (msg, grads)
        elif isinstance(msg, StepDone):
            self._done_msg = msg
        elif isinstance(msg, tuple) and msg and msg[0] == "partition":
            pass
        elif isinstance(msg, tuple) and msg and msg[0] == "mb_done":
            # the leader finished this microbatch's backward: its graphs can go
            mb = msg[1]
            for cid in [c for c, r in self.records.items() if r.mb == mb]:
                del self.records[cid]
            self.boundaries.pop(mb, None)
        else:
            raise SMPRuntimeError(f"unexpected pipeline message {type(msg).__name__}")

    def _wait_for(self, key):
        """Block until the result ``key`` has arrived and return (message, tensors).

        Inside a leader task (see ``_Scheduler``) the task is suspended and the scheduler decides what runs
        next; elsewhere (a non-leader stage waiting for a nested call) requests are served meanwhile."""
        task = getattr(_task_local, "task", None)
        if task is not None:
            task.suspend(key)
        else:
            while key not in self.stash:
                if not self._pump_one(0):
                    self._pump_one(20)
        msg, finish = self.stash.pop(key)
        return msg, finish()

    # ================================================================== forward across ranks
    def _remote_forward(self, module_name, span, owner, args, kwargs, chain=(), ckpt=None):
        st = self.state
        mb = st.microbatch
        skel, tensors = stubify((args, kwargs))
        call_id = self._new_call_id(mb)
        grad_on = torch.is_grad_enabled()
        req = ForwardRequest(mb, call_id, module_name if span is None else (module_name, span, ckpt), skel, grad_on,
                             self.me, chain, None, _autocast_state(st.device.type))
        self.core.timeline_record_pipeline_event(mb, f"fwd_req {module_name}")
        self.comm.server_send(req, owner)
        self.p2p.send(tensors, owner)
        msg, outs = self._wait_for(("F", call_id))
        st.microbatch = mb
        if msg.error is not None:
            raise SMPRuntimeError(f"remote execution of {module_name} on rank {owner} failed:\n{msg.error}")
        leaves = []
        for stub, t in zip(stubs_of(msg.skeleton), outs):
            if grad_on and stub.requires_grad and t.is_floating_point():
                t.requires_grad_(True)
            leaves.append(t)
        if grad_on:
            # the output gradients go to whoever produced the outputs: the last stage of a chain
            last = chain[-1][0] if chain else owner
            self.boundaries.setdefault(mb, []).append(
                Boundary(mb, call_id, last, leaves, tensors, [bool(t.requires_grad) for t in tensors]))
        return destubify(msg.skeleton, leaves)

    def _serve_forward(self, src, req: ForwardRequest, tensors):
        st = self.state
        prev_mb = st.microbatch
        st.microbatch = req.mb
        name, span, ckpt = (req.module_name, None, None) if isinstance(req.module_name, str) else req.module_name
        try:
            leaves = []
            for stub, t in zip(stubs_of(req.skeleton), tensors):
                if req.grad_enabled and stub.requires_grad and t.is_floating_point():
                    t.requires_grad_(True)
                leaves.append(t)
            args, kwargs = destubify(req.skeleton, leaves)
            self.core.timeline_record_pipeline_event(req.mb, f"fwd_exec {name}")
            # the caller's autocast region extends over the modules it calls, wherever they live
            with torch.set_grad_enabled(req.grad_enabled), _autocast_ctx(req.autocast):
                out = self._run_local(name, span, args, kwargs, ckpt)
            if req.chain:
                (nxt, nxt_span), rest = req.chain[0], tuple(req.chain[1:])
                skel, outs = stubify(((out,), {}))
            else:
                skel, outs = stubify(out)
            if req.grad_enabled:
                self.records[req.call_id] = ExecRecord(req.mb, leaves, outs, src, req.prev_hop)
            if req.chain:
                self.comm.server_send(ForwardRequest(req.mb, req.call_id, (name, nxt_span, ckpt), skel, req.grad_enabled,
                                                     req.requester, rest, self.me, req.autocast), nxt)
                self.p2p.send(outs, nxt)
            else:
                self.comm.server_send(ForwardResult(req.mb, req.call_id, skel), req.requester)
                self.p2p.send(outs, req.requester)
        except Exception:
            self.comm.server_send(ForwardResult(req.mb, req.call_id, None, traceback.format_exc()), req.requester)
            raise
        finally:
            st.microbatch = prev_mb

    # ================================================================== backward across ranks
    def _backward_segment(self, mb, tensors, grads):
        """Autograd over the local segment, then ship boundary gradients and continue from the returned
        input gradients until nothing is outstanding."""
        pairs = [(t, g) for t, g in zip(tensors, grads) if t.requires_grad and g is not None]
        if pairs:
            torch.autograd.backward([t for t, _ in pairs], [g for _, g in pairs], retain_graph=True)
        outstanding = []
        for b in self.boundaries.get(mb, []):
            meta, send = [], []
            for i, leaf in enumerate(b.out_leaves):
                if leaf.requires_grad and leaf.grad is not None:
                    meta.append((i, tuple(leaf.grad.shape), leaf.grad.dtype))
                    send.append(leaf.grad)
                    leaf.grad = None
            if send:
                self.comm.server_send(BackwardRequest(mb, b.call_id, meta, self.me), b.owner)
                self.p2p.send(send, b.owner)
                outstanding.append(b)
        for b in outstanding:
            msg, in_grads = self._wait_for(("B", b.call_id))
            self.state.microbatch = mb
            if msg.error is not None:
                raise PipelineParallelBWDError(f"remote backward on rank {b.owner} failed:\n{msg.error}")
            ts = [b.in_tensors[i] for i, _, _ in msg.grad_meta]
            if ts:
                self._backward_segment(mb, ts, in_grads)

    def _serve_backward(self, src, req: BackwardRequest, grads):
        st = self.state
        prev_mb = st.microbatch
        st.microbatch = req.mb
        try:
            rec = self.records.get(req.call_id)
            if rec is None:
                raise PipelineParallelBWDError(f"no forward record for call {req.call_id}")
            outs = [rec.outputs[i] for i, _, _ in req.grad_meta]
            self.core.timeline_record_pipeline_event(req.mb, "bwd_exec")
            self._backward_segment(req.mb, outs, grads)
            meta, send = [], []
            for i, leaf in enumerate(rec.input_leaves):
                if leaf.requires_grad and leaf.grad is not None:
                    meta.append((i, tuple(leaf.grad.shape), leaf.grad.dtype))
                    send.append(leaf.grad)
                    leaf.grad = None
            if rec.prev_hop is not None:            # fast mode: the gradients walk the chain backwards
                self.comm.server_send(BackwardRequest(req.mb, req.call_id, meta, req.requester), rec.prev_hop)
                self.p2p.send(send, rec.prev_hop)
            else:
                self.comm.server_send(BackwardResult(req.mb, req.call_id, meta), rec.requester)
                self.p2p.send(send, rec.requester)
        except Exception:
            self.comm.server_send(BackwardResult(req.mb, req.call_id, [], traceback.format_exc()), req.requester)
            raise
        finally:
            st.microbatch = prev_mb

    # ================================================================== DistributedModel entry points
    def run_model_forward(self, model, args, kwargs):
        root = model.module
        mm = self.state.module_manager
        if self.state.is_tracing_flag:               # unpartitioned model, executed where it lives
            return root(*args, **kwargs)
        if mm.get_partition(root) == self.pp_rank:
            return root(*args, **kwargs)
        return self._remote_forward(mm.get_module_name(root), None, self._owner_rank(root), args, kwargs)

    def run_model_backward(self, model, tensors, grads):
        """Called from inside the step function: register; the scheduler decides when it runs."""
        if self.state.is_tracing_flag:
            return
        mb = self.state.microbatch
        self.deferred_bwd[mb] = (tuple(tensors), tuple(grads))

    # ================================================================== the step
    def _run_microbatch(self, mb, phase):
        """One microbatch on the leader: the step function (forward; ``model.backward`` only registers the
        backward) and / or the registered backward, per ``phase`` in {"both", "fwd", "bwd"}."""
        ctx = self._step_ctx
        pipe = self.state.pipeline
        st = self.state
        st.microbatch = mb
        if phase in ("both", "fwd"):
            pipe.promote_status(mb)
            self.core.timeline_record_pipeline_event(mb, "fwd_start")
            fn, a, k = ctx["func"], ctx["args"][mb], ctx["kwargs"][mb]
            out = fn(*a, **k)
            st.microbatch = mb
            ctx["outputs"][mb] = map_structure(lambda x: x.detach() if torch.is_tensor(x) else x, out) \
                if ctx["detach"] else out
            if mb in self.deferred_bwd:
                pipe.mark_ready_for_backward(mb)
            else:
                pipe.mark_done(mb)
        if phase in ("both", "bwd") and mb in self.deferred_bwd:
            pipe.promote_status(mb)
            self.core.timeline_record_pipeline_event(mb, "bwd_start")
            tensors, grads = self.deferred_bwd.pop(mb)
            self._backward_segment(mb, tensors, grads)
            self.boundaries.pop(mb, None)
            del tensors, grads
            for r in self.core.get_pp_group():
                if r != self.me:
                    self.comm.server_send(("mb_done", mb), r)
            pipe.mark_done(mb)
            self.core.timeline_record_pipeline_event(mb, "bwd_end")

    # ---- record-and-pick-fastest (leaders only)
    def _tuner(self, window):
        if self.record_steps <= 0 or self.core.pp_size() < 2 or window < 2:
            return None
        key = (self.state.current_step_func_id, torch.is_grad_enabled())
        t = self._tuners.get(key)
        if t is None:
            pp = self.core.pp_size()
            cands = [c for c in dict.fromkeys((pp, pp + 1, max(1, pp - 1), window)) if 1 <= c <= window]
            t = self._tuners[key] = _ScheduleTuner(cands, self.record_steps)
        return t

    def _device_sync(self):
        dev = getattr(self.state, "device", None)
        if dev is not None and dev.type == "cuda":
            torch.cuda.synchronize(dev)

    def _agree_on_schedule(self, tuner):
        """Rank 0's fastest candidate becomes every leader's schedule (reference :586-596: rank 0 finds the best
        step and broadcasts it)."""
        from ..backend.collectives import RankType

        leaders = self.core.get_dp_group(pp_rank=0)
        if self.me == leaders[0]:
            choice = tuner.best()
            for r in leaders[1:]:
                self.comm.send(("schedule", choice), r, RankType.WORLD_RANK)
        else:
            tag, choice = self.comm.recv_from(leaders[0], RankType.WORLD_RANK)
            if tag != "schedule":
                raise SMPRuntimeError(f"expected the schedule decision from rank {leaders[0]}, got {tag!r}")
        tuner.choice = choice
        if self.me == leaders[0]:
            shown = {c: round(min(v) * 1e3, 2) for c, v in tuner.times.items() if v}
            get_logger().info(f"pipeline schedule: warm-up of {choice} microbatches picked (step ms by candidate: {shown})")

    def run_step(self, step_fn, args, kwargs, num_mb):
        st = self.state
        if not self.partitioned:
            self._partition(step_fn, args, kwargs, num_mb)
        self._done_msg = None
        key = (st.current_step_func_id, torch.is_grad_enabled(), num_mb)
        # every rank counts the steps of a step function the way the leaders' tuner does (runs with and without
        # gradients apart): the schedule is final from step ``record_steps + 1`` on
        seen = self._steps_of.get(key[:2], 0)
        self._steps_of[key[:2]] = seen + 1
        tuning = (self.record_steps > 0 and self.core.pp_size() >= 2 and st.cfg.pipeline == "interleaved"
                  and int(st.cfg.active_microbatches) >= 2 and seen <= self.record_steps)
        self.p2p.begin_step(key, schedule_settled=not tuning)
        pipe = st.pipeline
        pipe.init_step()
        group = self.core.get_pp_group()
        if self.is_leader:
            a, k = step_fn.splitter.preprocess_args_all_mbs(args, kwargs, num_mb)
            self._step_ctx = {"func": step_fn.func, "args": a, "kwargs": k, "outputs": [None] * num_mb,
                              "detach": step_fn.detach_outputs}
            try:
                window = max(1, int(st.cfg.active_microbatches))
                interleaved = st.cfg.pipeline == "interleaved"

                def next_job():
                    # the pipeline policy names the microbatch: ``interleaved`` - one ready for backward before a
                    # new forward (a microbatch started here runs its backward in the same task, as soon as its
                    # forward has returned); ``simple`` - every forward, and backwards only once no forward is in
                    # flight; ``_only_forward`` - forwards only
                    mb = pipe.get_next_microbatch()
                    if mb is None:
                        return None
                    if pipe.get_status(mb) is MbStatus.READY_FOR_FWD:
                        return (mb, "both" if interleaved else "fwd")
                    return (mb, "bwd")

                warm = self.core.pp_size() if interleaved else None
                tuner = self._tuner(window) if interleaved else None
                if tuner is not None:
                    if tuner.must_decide:
                        self._agree_on_schedule(tuner)
                    warm = tuner.candidate()
                    timed = tuner.recording
                    if timed:
                        self._device_sync()
                        t0 = time.perf_counter()
                _Scheduler(self, next_job, window, warm=warm).run()
                if tuner is not None:
                    if timed:
                        self._device_sync()
                    tuner.record(time.perf_counter() - t0 if timed else 0.0)
                outputs = self._step_ctx["outputs"]
            except BaseException as exc:
                # the other stages are waiting for this step to end: tell them it ended badly, so that the
                # job fails everywhere instead of hanging on the ranks that saw nothing wrong
                for r in group:
                    if r != self.me:
                        self.comm.server_send(StepDone(self._step, None, f"{type(exc).__name__}: {exc}"), r)
                # sends are asynchronous: the notice has to be on the wire before the exception travels on and
                # possibly ends this process
                self.comm.wait_all_transactions()
                raise
            finally:
                self._step_ctx = None
            # the outputs go to the other stages as tensor stubs + point-to-point transfers (no device-to-host copy,
            # so the leader's host does not wait for the step's GPU work before it can start the next step)
            skel, tensors = stubify(outputs)
            for r in group:
                if r != self.me:
                    self.comm.server_send(StepDone(self._step, skel), r)
                    self.p2p.send(tensors, r)
        else:
            while self._done_msg is None:
                self._pump_one(50)
            if self._done_msg.error is not None:
                raise SMPRuntimeError(f"the step failed on pipeline rank 0: {self._done_msg.error}")
            skel = self._done_msg.outputs
            metas = [(sb.shape, sb.dtype) for sb in stubs_of(skel)]
            leader = self.core.pp_rank_to_rank(0)
            outputs = destubify(skel, self.p2p.recv(metas, leader, st.device))
        self.p2p.drain()
        self.records.clear()
        self.boundaries.clear()
        self.stash.clear()
        self.deferred_bwd.clear()
        self._step += 1
        return outputs
