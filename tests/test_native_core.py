"""Native host runtime (libsmpb200_core.so) through its C ABI: gradient readiness counter, bucket planner,
timeline, and the TCP message bus between two processes (reference tier: native self-tests in
test/backend/*.py; GradCounter checks in test/torch/mpi_hybrid/smp_test_base.py)."""

import ctypes
import json
import os
import tempfile

import pytest

from dist_utils import run_distributed
from smdistributed_modelparallel_b200.backend import native_core


@pytest.fixture(scope="module")
def lib():
    return native_core.load()


def _ids(*v):
    return (ctypes.c_int * len(v))(*v), len(v)


def test_grad_counter_lifecycle(lib):
    h = lib.smpc_gc_create(3, 2)
    try:
        for mb in range(2):
            arr, n = _ids(0, 1, 1)                      # parameter 1 is used twice per microbatch
            assert lib.smpc_gc_increment_expected(h, mb, arr, n) == 0
            assert lib.smpc_gc_mark_fwd_done(h, mb) == 0
        assert lib.smpc_gc_expected(h, 0, 1) == 2 and lib.smpc_gc_expected(h, 1, 0) == 1
        tot = (ctypes.c_int * 3)()
        lib.smpc_gc_expected_totals(h, tot)
        assert list(tot) == [2, 4, 0]
        lib.smpc_gc_set_microbatch(h, 0)
        assert lib.smpc_gc_get_microbatch(h) == 0
        assert lib.smpc_gc_mark_grad_computed(h, 0) == 0
        assert lib.smpc_gc_mark_grad_computed(h, 1) == 0
        assert lib.smpc_gc_mark_grad_computed(h, 1) == 0
        assert lib.smpc_gc_mark_grad_computed(h, 1) == -2   # more contributions than announced
        assert lib.smpc_gc_is_ready(h, 0) == 0
        lib.smpc_gc_set_microbatch(h, 1)
        assert lib.smpc_gc_mark_grad_computed(h, 0) == 1     # final contribution of the step
        assert lib.smpc_gc_mark_grad_computed(h, 1) == 0
        assert lib.smpc_gc_mark_grad_computed(h, 1) == 1
        assert lib.smpc_gc_is_ready(h, 0) == 1 and lib.smpc_gc_is_ready(h, 1) == 1
        assert lib.smpc_gc_seen(h, 1, 1) == 2
        assert lib.smpc_gc_mark_grad_computed(h, 7) == -1
        lib.smpc_gc_clear(h)
        assert lib.smpc_gc_expected(h, 0, 1) == 0 and lib.smpc_gc_is_ready(h, 1) == 0
    finally:
        lib.smpc_gc_destroy(h)


def test_bucket_planner(lib):
    mb = 1 << 20
    sizes = [mb // 2, mb // 2, mb // 2, 40 * mb, 70 * mb, 10 * mb, 30 * mb]
    keys = [0] * len(sizes)
    out = (ctypes.c_int * len(sizes))()
    n = lib.smpc_plan_buckets(len(sizes), (ctypes.c_longlong * len(sizes))(*sizes), (ctypes.c_int * len(keys))(*keys),
                              1 * mb, 100 * mb, out)
    # first bucket closes at 1 MB, later ones at 100 MB
    assert list(out) == [0, 0, 1, 1, 2, 2, 3] and n == 4
    # tensors of different dtype never share a bucket
    sizes, keys = [10, 10, 10, 10], [0, 1, 0, 1]
    out = (ctypes.c_int * 4)()
    n = lib.smpc_plan_buckets(4, (ctypes.c_longlong * 4)(*sizes), (ctypes.c_int * 4)(*keys), 1000, 1000, out)
    assert n == 2 and out[0] == out[2] and out[1] == out[3] and out[0] != out[1]


def test_timeline_writes_chrome_trace(lib):
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "timeline.json")
        assert lib.smpc_timeline_create(path.encode(), 0) == 0
        assert lib.smpc_timeline_enabled() == 1
        lib.smpc_timeline_start_step()
        lib.smpc_timeline_record_pipeline_event(0, b"fwd_mb0")
        lib.smpc_timeline_record_pipeline_event(0, b"fwd_mb0")
        lib.smpc_timeline_record_pipeline_event(1, b"bwd_mb1")
        lib.smpc_timeline_end_step()
        assert lib.smpc_timeline_num_events() >= 2
        assert lib.smpc_timeline_flush() == 0
        lib.smpc_timeline_destroy()
        text = open(path).read()
        assert "fwd_mb0" in text and "bwd_mb1" in text
        data = json.loads(text)
        names = [e["name"] for e in data["traceEvents"]]
        assert names[0] == "step" and names[-1] == "step"


def test_memory_metrics(lib):
    lib.smpc_pool_record(1 << 20, 2 << 20, 1)
    lib.smpc_pool_record(1 << 10, 0, 0)
    out = (ctypes.c_longlong * 2)()
    assert lib.smpc_get_and_reset_alloc_metrics(out) == 0
    assert out[0] >= 1 and out[1] >= 1                  # successes, failures
    assert lib.smpc_get_and_reset_alloc_metrics(out) == 0
    assert out[0] == 0 and out[1] == 0


def _bus_worker(rank, world):
    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.backend.collectives import CommGroup, RankType

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 2})
    comm = smp.state.comm
    peer = 1 - rank
    # point to point, including a payload larger than one socket buffer
    big = bytes(range(256)) * 8192
    comm.send({"from": rank, "blob": big}, peer, RankType.WORLD_RANK)
    got = comm.recv_from(peer, RankType.WORLD_RANK)
    assert got["from"] == peer and got["blob"] == big
    # collectives on the control plane
    assert comm.allgather(rank * 10, CommGroup.WORLD) == [0, 10]
    if rank == 0:                                        # broadcast is sender-side; receivers call recv_from
        comm.broadcast("hello", CommGroup.WORLD)
    else:
        assert comm.recv_from(0, RankType.WORLD_RANK) == "hello"
    g = comm.gather(rank + 1, CommGroup.WORLD, 0)
    if rank == 0:
        assert g == [1, 2]
    comm.barrier()
    return True


def test_message_bus_two_processes():
    assert all(run_distributed(_bus_worker, 2))


def _failure_worker(rank, world):
    import time

    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.backend.collectives import RankType
    from smdistributed_modelparallel_b200.backend.exceptions import SMPRuntimeError

    smp.init({"pipeline_parallel_degree": 2, "microbatches": 2})
    if rank == 1:
        time.sleep(0.5)
        smp.state.core.shutdown(success=False)            # what the exit hook does when this rank dies
        return "failed-on-purpose"
    t0 = time.time()
    try:
        smp.state.comm.recv_from(1, RankType.WORLD_RANK)   # would block forever without failure propagation
    except SMPRuntimeError as e:
        return f"aborted after {time.time() - t0:.1f}s: {e}"
    return "received?"


def test_peer_failure_aborts_blocking_receive():
    """Failure propagation (reference backend/core.py:165-188,227-232; test/backend/shutdown_test.py): a rank that
    shuts down unsuccessfully makes its peers' blocking receives raise instead of hanging."""
    res = run_distributed(_failure_worker, 2, timeout=120)
    assert res[1] == "failed-on-purpose"
    assert res[0].startswith("aborted"), res[0]


def test_in_library_selftest():
    """The native library checks its own counter / planner / metrics through the C ABI."""
    from smdistributed_modelparallel_b200.backend import native_core

    msg = native_core.selftest()
    assert msg.startswith("ok:"), msg
    assert int(msg.split()[1]) >= 20


def test_link_manager_reuses_slots_by_size():
    """Link ids: smallest free link that fits, a fresh one otherwise, everything free again after reset; each
    pipeline rank draws from its own id range (reference torch/links.py:21-84)."""
    from smdistributed_modelparallel_b200.runtime.links import IDS_PER_RANK, HandleManager, LinkManager

    lm = LinkManager(pp_rank=2)
    a = lm.get_link(1000)
    b = lm.get_link(1000)
    c = lm.get_link(4000)
    assert len({a, b, c}) == 3 and all(2 * IDS_PER_RANK < x < 3 * IDS_PER_RANK for x in (a, b, c))
    assert lm.in_use() == 3
    lm.reset()
    assert lm.in_use() == 0 and lm.wasted_bytes() == 6000
    # a smaller request takes the smallest link that is large enough
    assert lm.get_link(900) in (a, b)
    assert lm.get_link(900) in (a, b)
    assert lm.get_link(900) == c               # both 1000-byte links are busy: the 4000-byte one fits
    d = lm.get_link(900)                       # nothing free: a new link of exactly that size
    assert d not in (a, b, c) and lm.link_size(d) == 900
    assert lm.stats[c] == [2, 4900]
    lm.reset()
    assert LinkManager(pp_rank=0).get_link(8) < IDS_PER_RANK

    class _Work:
        def __init__(self, done):
            self.done = done

        def is_completed(self):
            return self.done

        def wait(self):
            self.done = True

    hm = HandleManager()
    h1 = hm.register_handle(_Work(True), hm.SEND, "t1", link=a)
    h2 = hm.register_handle(_Work(False), hm.RECV, "t2")
    assert hm.is_valid_handle(h1) and hm.get_handle_direction(h2) == "recv" and hm.get_handle_output(h1) == "t1"
    assert hm.get_handle_link(h1) == a
    assert hm.reap() == 1 and not hm.is_valid_handle(h1)
    hm.wait(h2)
    assert hm.poll(h2)
    hm.clear_handle(h2)
    assert hm.handles() == []


def test_message_bus_handshake_token_source_stamp_and_lost_peer():
    """Robustness of the control bus, driven through the C ABI with a raw socket playing rank 1: a connection that
    presents the wrong job token is closed and does not take the rank's slot; frames are attributed to the socket's
    rank whatever their header claims; an oversized length is treated as a broken peer; and a connection that ends
    without a goodbye frame wakes blocked receivers with the failing rank (a killed process sends nothing)."""
    import socket
    import struct
    import threading

    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.backend import native_core

    smp.reset()
    lib = native_core.load()
    lib.smpc_bus_shutdown(1)
    magic = 0x534D5042
    try:
        port = lib.smpc_bus_listen_on(b"127.0.0.1")
        assert port > 0
        assert lib.smpc_bus_listen_on(b"127.0.0.1") == -1            # already listening
        assert lib.smpc_bus_set_token(1234) == 0
        rc = []
        t = threading.Thread(target=lambda: rc.append(lib.smpc_bus_connect(0, 2, f"127.0.0.1:{port},127.0.0.1:1".encode(), 30)))
        t.start()
        rogue = socket.create_connection(("127.0.0.1", port))
        rogue.sendall(struct.pack("<IIQ", 1, magic, 999))             # right rank, wrong token
        rogue.settimeout(10)
        assert rogue.recv(1) == b""                                   # closed by the listener
        rogue.close()
        assert t.is_alive()                                           # still waiting for the real rank 1
        peer = socket.create_connection(("127.0.0.1", port))
        peer.sendall(struct.pack("<IIQ", 1, magic, 1234))
        t.join(30)
        assert rc == [0]
        # a frame that claims to come from rank 0 is filed under the rank of its socket
        peer.sendall(struct.pack("<IIQQ", 0, 0, 77, 3) + b"abc")
        assert lib.smpc_bus_wait(1, 77, 5000) == 3
        assert lib.smpc_bus_poll(0, 77) == -1
        assert lib.smpc_bus_peer_failed() == -1
        # the peer vanishes without saying goodbye
        peer.close()
        assert lib.smpc_bus_wait(1, 78, 10000) == -2
        assert lib.smpc_bus_peer_failed() == 1
    finally:
        lib.smpc_bus_shutdown(1)
    # a clean goodbye is not a failure; an absurd frame length is
    for goodbye, expect in ((True, -1), (False, 1)):
        port = lib.smpc_bus_listen_on(b"127.0.0.1")
        lib.smpc_bus_set_token(5)
        rc = []
        t = threading.Thread(target=lambda: rc.append(lib.smpc_bus_connect(0, 2, f"127.0.0.1:{port},127.0.0.1:1".encode(), 30)))
        t.start()
        peer = socket.create_connection(("127.0.0.1", port))
        peer.sendall(struct.pack("<IIQ", 1, magic, 5))
        t.join(30)
        assert rc == [0]
        if goodbye:
            peer.sendall(struct.pack("<IIQQ", 1, 2, 2, 0))            # control channel, CTL_GOODBYE
            peer.close()
            assert lib.smpc_bus_wait(1, 1, 1500) == -1                # plain timeout
        else:
            peer.sendall(struct.pack("<IIQQ", 1, 0, 9, 1 << 60))
            assert lib.smpc_bus_wait(1, 9, 10000) == -2
        assert lib.smpc_bus_peer_failed() == expect
        peer.close()
        lib.smpc_bus_shutdown(1)
