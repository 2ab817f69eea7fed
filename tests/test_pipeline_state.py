"""Microbatch schedule state machines stepped by hand, tensor stub round trip and module-manager bookkeeping
(reference test/torch/test_pipeline.py:11-206, test_serialization.py, test_module_manager.py tier)."""

import pytest
import torch
import torch.nn as nn

from smdistributed_modelparallel_b200.pipeline.module_manager import ModuleManager
from smdistributed_modelparallel_b200.runtime.messages import TensorStub, destubify, stubify, stubs_of
from smdistributed_modelparallel_b200.runtime.pipeline import (InterleavedPipeline, MbStatus, OnlyForwardPipeline,
                                                               SimplePipeline)


def _drive(p, after_fwd):
    """Run the schedule to completion; every forward finishes immediately and becomes ready for backward."""
    order = []
    while p.has_more_ticks():
        mb = p.get_next_microbatch()
        assert mb is not None
        st = p.get_status(mb)
        p.promote_status(mb)
        if st is MbStatus.READY_FOR_FWD:
            order.append(("F", mb))
            after_fwd(p, mb)
        else:
            order.append(("B", mb))
            p.mark_done(mb)
    return order


def test_simple_runs_all_forwards_then_backwards():
    p = SimplePipeline(num_microbatches=4)
    order = _drive(p, lambda p, mb: p.mark_ready_for_backward(mb))
    assert order == [("F", 0), ("F", 1), ("F", 2), ("F", 3), ("B", 0), ("B", 1), ("B", 2), ("B", 3)]


def test_simple_waits_for_outstanding_forwards():
    p = SimplePipeline(num_microbatches=2)
    assert p.get_next_microbatch() == 0
    p.promote_status(0)                                  # forward of 0 in flight on a later stage
    assert p.get_next_microbatch() == 1
    p.promote_status(1)
    assert p.get_next_microbatch() is None               # nothing to do until a forward returns
    p.mark_ready_for_backward(1)
    assert p.get_next_microbatch() is None               # mb 0 still in forward: simple schedule does not start backward
    p.mark_ready_for_backward(0)
    assert p.get_next_microbatch() == 0


def test_interleaved_prefers_backward():
    p = InterleavedPipeline(num_microbatches=3)
    order = _drive(p, lambda p, mb: p.mark_ready_for_backward(mb))
    assert order == [("F", 0), ("B", 0), ("F", 1), ("B", 1), ("F", 2), ("B", 2)]
    # 1F1B steady state when results come back one tick late
    p = InterleavedPipeline(num_microbatches=4)
    assert p.get_next_microbatch() == 0
    p.promote_status(0)
    assert p.get_next_microbatch() == 1
    p.promote_status(1)
    p.mark_ready_for_backward(0)
    assert p.get_next_microbatch() == 0 and p.get_status(0) is MbStatus.READY_FOR_BWD
    p.promote_status(0)
    assert p.in_flight() == 2
    p.mark_done(0)
    assert p.get_next_microbatch() == 2


def test_only_forward():
    p = OnlyForwardPipeline(num_microbatches=3)
    seen = []
    for _ in range(3):
        mb = p.get_next_microbatch()
        seen.append(mb)
        p.promote_status(mb)
        p.mark_done(mb)
    assert seen == [0, 1, 2] and not p.has_more_ticks()


def test_get_next_mb_with_status_scans_cyclically():
    p = SimplePipeline(num_microbatches=4)
    p.mark_ready_for_backward(1)
    p.mark_ready_for_backward(3)
    assert p.get_next_mb_with_status(MbStatus.READY_FOR_BWD) == 1
    assert p.get_next_mb_with_status(MbStatus.READY_FOR_BWD, after=1) == 3
    assert p.get_next_mb_with_status(MbStatus.READY_FOR_BWD, after=3) == 1
    assert p.get_next_mb_with_status([MbStatus.DONE]) is None


def test_stub_round_trip():
    a, b = torch.randn(2, 3, requires_grad=True), torch.arange(4)
    obj = ((a, {"k": [b, 5, "s"]}), {"flag": None})
    skel, tensors = stubify(obj)
    assert len(tensors) == 2 and tensors[0] is a and tensors[1] is b
    stubs = stubs_of(skel)
    assert [s.index for s in stubs] == [0, 1]
    assert stubs[0] == TensorStub(0, (2, 3), torch.float32, True)
    assert stubs[1].dtype == torch.int64 and not stubs[1].requires_grad
    back = destubify(skel, [t.clone() for t in tensors])
    assert torch.equal(back[0][0], a) and torch.equal(back[0][1]["k"][0], b)
    assert back[0][1]["k"][1:] == [5, "s"] and back[1] == {"flag": None}


def test_module_manager_names_partitions_and_marks():
    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.a = nn.Linear(4, 4)
            self.b = nn.Sequential(nn.Linear(4, 4), nn.ReLU())

    mm = ModuleManager()

    from smdistributed_modelparallel_b200.backend.config import ModelParallelConfig

    mm.set_config(ModelParallelConfig({"pipeline_parallel_degree": 2, "microbatches": 2}))
    net = Net()
    mm.name_modules_and_create_parent_map(net)
    assert mm.get_module_name(net) == "main" or mm.get_module_name(net) == ""
    assert mm.get_module(mm.get_module_name(net.b[0])) is net.b[0]
    assert net.b in mm.get_parent_modules(net.b[0])
    name = mm.get_module_name
    mm.assign_partitions({name(net): 0, name(net.a): 0, name(net.b): 1, name(net.b[0]): 1, name(net.b[1]): 1})
    with pytest.raises(Exception):
        mm.assign_partitions({name(net): 7})
    assert mm.get_partition(net.b[0]) == 1 and mm.is_executor(net.a, 0) and not mm.is_executor(net.b, 0)
    assert mm.get_module_name(net.b) in mm.local_module_names(1)
    mm.set_activation_checkpointing(net.b, pack_args_as_tuple=True, strategy="each")
    assert mm.should_checkpoint_activations(net.b) and not mm.should_checkpoint_activations(net.a)
    mm.set_tensor_parallelism(net.a, True, foo=1)
    assert mm.is_marked_for_tp(net.a) and mm.get_tp_config(net.a) == {"foo": 1}
    with pytest.raises(Exception):
        mm.set_partition(net.a, 5)


def test_leader_scheduler_order_is_static():
    """The leader's task order depends only on (microbatches, window, warm): tasks are suspendable threads, the
    scheduler starts ``warm`` of them, then alternates resume-oldest / start-next inside the window - regardless
    of when results arrive (here they are all available at once)."""
    from smdistributed_modelparallel_b200.runtime import server as S

    class FakeServer:
        def __init__(self):
            self.stash, self.log = {}, []
            import torch
            self.state = type("St", (), {"device": torch.device("cpu")})()

        def _pump_one(self, timeout_ms):
            raise AssertionError("every awaited key is already stashed in this test")

        def _run_microbatch(self, mb, phase):
            for seg in range(3):                       # forward chunk, head + backward request, backward chunk
                self.log.append((mb, seg))
                if seg < 2:
                    key = ("K", mb, seg)
                    self.stash[key] = True
                    S._task_local.task.suspend(key)
                    self.stash.pop(key)

    def order(nmb, window, warm):
        srv = FakeServer()
        jobs = iter([(mb, "both") for mb in range(nmb)])
        S._Scheduler(srv, lambda: next(jobs, None), window, warm=warm).run()
        return srv.log

    got = order(4, 4, 2)
    assert got == [(0, 0), (1, 0), (0, 1), (2, 0), (1, 1), (3, 0), (0, 2), (2, 1), (1, 2), (3, 1), (2, 2), (3, 2)], got
    assert order(4, 4, 2) == got                                    # repeatable
    fill = order(3, 3, 3)                                           # warm == window: fill first, then oldest-first
    assert fill[:3] == [(0, 0), (1, 0), (2, 0)] and fill[3] == (0, 1)
    one = order(3, 1, 1)                                            # window 1: strictly one microbatch at a time
    assert one == [(mb, seg) for mb in range(3) for seg in range(3)]
    every = sorted(order(8, 4, 2))
    assert every == [(mb, seg) for mb in range(8) for seg in range(3)]


def test_scheduler_takes_its_microbatch_order_from_the_pipeline_policy():
    """The leader's scheduler asks the policy object (SimplePipeline / InterleavedPipeline) which microbatch to start
    (reference torch/pipeline.py:96-145 get_next_microbatch): ``simple`` starts no backward while a forward is in
    flight; ``interleaved`` runs a microbatch's backward before starting the next forward once it is ready."""
    import torch

    from smdistributed_modelparallel_b200.runtime import server as S
    from smdistributed_modelparallel_b200.runtime.pipeline import InterleavedPipeline, MbStatus, SimplePipeline

    def drive(pipe, window, phase_of):
        log = []

        class FakeServer:
            stash = {}
            state = type("St", (), {"device": torch.device("cpu")})()

            def _pump_one(self, timeout_ms):
                raise AssertionError("nothing to wait for")

            def _run_microbatch(self, mb, phase):
                log.append((mb, phase))
                pipe.promote_status(mb)
                if phase == "fwd":
                    key = ("K", mb)
                    self.stash[key] = True
                    S._task_local.task.suspend(key)        # a forward that waits for another stage
                    self.stash.pop(key)
                    log.append((mb, "fwd_end"))
                    pipe.mark_ready_for_backward(mb)
                else:
                    pipe.mark_done(mb)

        def next_job():
            mb = pipe.get_next_microbatch()
            if mb is None:
                return None
            return (mb, phase_of(pipe.get_status(mb)))

        S._Scheduler(FakeServer(), next_job, window).run()
        return log

    phase = lambda st: "fwd" if st is MbStatus.READY_FOR_FWD else "bwd"
    simple = drive(SimplePipeline(3, 3), 3, phase)
    first_bwd = next(i for i, e in enumerate(simple) if e[1] == "bwd")
    assert [e for e in simple[:first_bwd] if e[1] == "fwd_end"] == [(0, "fwd_end"), (1, "fwd_end"), (2, "fwd_end")]
    assert [e for e in simple if e[1] == "bwd"] == [(0, "bwd"), (1, "bwd"), (2, "bwd")]
    inter = drive(InterleavedPipeline(3, 3), 3, phase)
    # microbatch 0's backward starts as soon as its forward has ended, before microbatch 2's forward ends
    assert inter.index((0, "bwd")) < inter.index((2, "fwd_end"))
    assert sorted(e for e in inter if e[1] == "bwd") == [(0, "bwd"), (1, "bwd"), (2, "bwd")]


def test_tensors_inside_plain_objects_are_stubbed():
    """An instance of a user class that carries tensors crosses a stage boundary as the same class with stubs in place
    of the tensors; the caller's instance is not modified (reference test/torch/test_serialization.py:13-40)."""
    import torch

    from smdistributed_modelparallel_b200.runtime.messages import TensorStub, destubify, stubify, stubs_of

    class Holder:
        def __init__(self, x):
            self.tensor = torch.tensor([x])
            self.name = "h"
            self.inner = {"t": torch.ones(2, 3), "n": 3}

    class Plain:
        def __init__(self, x):
            self.x = x

    a = Holder(3.2)
    skel, tensors = stubify(([a, Plain(5)], {"k": a.tensor}))
    h = skel[0][0]
    assert type(h) is Holder and isinstance(h.tensor, TensorStub) and isinstance(h.inner["t"], TensorStub)
    assert h.name == "h" and h.inner["n"] == 3
    assert isinstance(a.tensor, torch.Tensor) and isinstance(a.inner["t"], torch.Tensor)     # untouched
    assert skel[0][1].x == 5 and len(tensors) == 3
    assert [s.index for s in stubs_of(skel)] == [0, 1, 2] and stubs_of(skel)[1].shape == (2, 3)
    back = destubify(skel, tensors)
    assert torch.equal(back[0][0].tensor, a.tensor) and torch.equal(back[0][0].inner["t"], a.inner["t"])
    assert back[1]["k"] is tensors[2]
    import pickle
    pickle.loads(pickle.dumps(stubify(({"k": torch.ones(1)},))[0]))                          # skeletons travel pickled


def test_stubbing_survives_reference_cycles():
    import torch

    from smdistributed_modelparallel_b200.runtime.messages import TensorStub, stubify

    class Node:
        def __init__(self):
            self.t = torch.ones(2)
            self.me = self                      # a cycle through an attribute
            self.peers = [self]

    n = Node()
    skel, tensors = stubify({"n": n, "again": n})
    assert len(tensors) == 2                    # the object is reached twice from the top, each time walked once
    assert isinstance(skel["n"].t, TensorStub) and skel["n"].me is n and skel["n"].peers[0] is n
