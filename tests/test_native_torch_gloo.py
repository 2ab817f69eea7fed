"""The native tensor-communication runtime (csrc/torch_ext/comm_runtime.cpp, a PyTorch C++ extension over c10d) on four
gloo ranks: uneven split / merge all-to-all and variable-count all-gather against their definitions (the expectations
of reference test/torch/mpi_4ps/test_collectives.py:13-114), point-to-point transfers through handles (reference
torch/ops.py:34-117), and the bucket reducer - early launches, uncovered tails, per-region groups, divisor (reference
torch/ddp_model.py:336-374)."""

import torch

from dist_utils import run_distributed


def _worker(rank, world, cuda=False):
    import torch.distributed as dist

    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.backend import native_torch

    smp.init({"tensor_parallel_degree": 2, "ddp": True, **({} if cuda else {"ddp_dist_backend": "gloo"})})
    if cuda:
        torch.set_default_device(torch.device("cuda", smp.local_rank()))      # every tensor below lives on this GPU
    nt = native_torch.load()
    pg = smp.get_world_process_group()
    out = {}
    # ---- split / merge all-to-all: every rank holds [6 + rank, 7, 3]; dim 1 is split 2,2,2,1; dim 0 merged
    x = torch.arange((6 + rank) * 7 * 3, dtype=torch.float32).reshape(6 + rank, 7, 3) + 1000 * rank
    y = nt.scatter_and_merge(pg, x, 1, 0)
    out["a2a"] = y
    # explicit shapes, negative dims, merge along the last dim
    z = nt.scatter_and_merge(pg, x.transpose(0, 2).contiguous(), -2, -1, [1, 3, 2, 1], [6, 7, 8, 9])
    out["a2a_explicit"] = z
    # ---- variable-count all-gather
    counts = [3, 0, 5, 2]
    flat = torch.full((10,), -1.0)
    off = sum(counts[:rank])
    flat[off:off + counts[rank]] = float(rank + 1)
    nt.allgatherv(pg, flat, counts)
    out["agv"] = flat
    even = torch.zeros(8)
    even[2 * rank:2 * rank + 2] = rank + 1
    nt.allgatherv(pg, even, [2, 2, 2, 2])
    out["agv_even"] = even
    # ---- p2p handles: a ring, tags keep two messages apart
    eng = nt.P2PEngine()
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    a, b = torch.full((5,), float(rank)), torch.full((3, 2), float(rank) + 0.5)
    ra, rb = torch.empty(5), torch.empty(3, 2)
    if cuda and rank % 2 == 0:
        # NCCL runs unbatched transfers in issue order on one stream: a ring in which everybody receives first
        # would wait for ever, so even ranks send first
        hs1, hr1 = eng.send(pg, a, nxt, 1), eng.recv(pg, ra, prv, 1)
        hs2, hr2 = eng.send(pg, b, nxt, 2), eng.recv(pg, rb, prv, 2)
    elif cuda:
        hr1, hs1 = eng.recv(pg, ra, prv, 1), eng.send(pg, a, nxt, 1)
        hr2, hs2 = eng.recv(pg, rb, prv, 2), eng.send(pg, b, nxt, 2)
    else:
        hr1, hr2 = eng.recv(pg, ra, prv, 1), eng.recv(pg, rb, prv, 2)
        hs1, hs2 = eng.send(pg, a, nxt, 1), eng.send(pg, b, nxt, 2)
    got_b = eng.get_ready_tensor(hr2)
    got_a = eng.get_ready_tensor(hr1)
    eng.wait(hs1)
    eng.wait(hs2)
    if cuda:
        torch.cuda.synchronize()            # NCCL's wait orders streams; completion needs the kernels to have run
    assert eng.poll(hs1) and eng.poll(hr1)
    out["p2p"] = (got_a.clone(), got_b.clone(), got_a.data_ptr() == ra.data_ptr())
    st = eng.stats()
    out["p2p_stats"] = (st["sent"], st["received"], st["bytes_sent"], st["pending"])
    assert eng.reap_sends() == 0
    eng.clear(hr1)
    eng.clear(hr2)
    out["p2p_pending_after"] = eng.pending()
    try:
        eng.poll(hr1)
        out["cleared_handle_refused"] = False
    except RuntimeError:
        out["cleared_handle_refused"] = True
    # ---- bucket reducer: region 0 [0, 40) over the tp group (2 ranks), region 1 [40, 100) over the world
    buf = torch.arange(100, dtype=torch.float32) * (rank + 1)
    red = nt.BucketReducer(buf, 16)
    r0 = red.add_region(0, 40, smp.get_tp_process_group())
    r1 = red.add_region(40, 100, pg)
    r_none = red.add_region(100, 100, None)
    assert red.launch(r1, 60, 80) is True
    assert red.launch(r1, 70, 90) is False          # overlaps a slice already on its way
    assert red.launch(r0, 0, 8) is True
    assert red.launch(r_none, 100, 100) is False
    assert red.in_flight() == 2
    red.finish(0.5)
    out["reduced"] = buf.clone()
    out["red_stats"] = dict(red.stats())
    # ---- flat bucket of separate gradient tensors
    g1, g2 = torch.full((2, 3), float(rank)), torch.full((4,), 10.0 * rank)
    fb = nt.FlatBucket(pg, [g1, g2])
    fb.finish(0.25)
    out["flat"] = (g1.clone(), g2.clone(), fb.numel())
    dist.barrier()
    return out


def test_native_runtime_on_four_ranks():
    _check_runtime(run_distributed(_worker, 4, timeout=300))


def _gpu_gate():
    import os

    import pytest
    if torch.cuda.device_count() < 4:
        pytest.skip("needs 4 GPUs")
    if os.environ.get("SMP_RUN_NATIVE_GPU", "0") != "1":
        # written after the round's last multi-GPU session: not yet run over NCCL, so a -m gpu run on a larger box
        # does not meet it unannounced
        pytest.skip("set SMP_RUN_NATIVE_GPU=1 to run the native communication runtime tests over NCCL")


import pytest  # noqa: E402


@pytest.mark.gpu
def test_native_runtime_on_four_gpus():
    """The same checks with CUDA tensors over NCCL (grouped all-to-all, NCCL send / recv handles, all-reduce buckets)."""
    _gpu_gate()
    _check_runtime(run_distributed(_worker, 4, True, timeout=300, use_cuda=True))


@pytest.mark.gpu
def test_all_to_all_with_a_gradient_all_reduce_in_flight_on_gpus():
    _gpu_gate()
    assert all(run_distributed(_concurrent_worker, 4, True, timeout=300, use_cuda=True))


def _check_runtime(res):
    n = 4
    xs = [torch.arange((6 + r) * 7 * 3, dtype=torch.float32).reshape(6 + r, 7, 3) + 1000 * r for r in range(n)]
    split = [2, 2, 2, 1]
    for r, out in enumerate(res):
        lo = sum(split[:r])
        want = torch.cat([x[:, lo:lo + split[r]] for x in xs], dim=0)
        assert torch.equal(out["a2a"], want)
        xt = [x.transpose(0, 2).contiguous() for x in xs]                 # [3, 7, 6 + k]
        es = [1, 3, 2, 1]
        lo = sum(es[:r])
        want = torch.cat([t[:, lo:lo + es[r]] for t in xt], dim=2)
        assert torch.equal(out["a2a_explicit"], want)
        assert out["agv"].tolist() == [1, 1, 1, 3, 3, 3, 3, 3, 4, 4]
        assert out["agv_even"].tolist() == [1, 1, 2, 2, 3, 3, 4, 4]
        prv = (r - 1) % n
        ga, gb, same = out["p2p"]
        assert torch.equal(ga, torch.full((5,), float(prv))) and torch.equal(gb, torch.full((3, 2), prv + 0.5)) and same
        assert tuple(out["p2p_stats"]) == (2, 2, 5 * 4 + 6 * 4, 4)
        assert out["p2p_pending_after"] == 0 and out["cleared_handle_refused"]   # sends reaped, receives cleared
        base = torch.arange(100, dtype=torch.float32)
        tp_sum = (1 + 2) if r < 2 else (3 + 4)                            # tp groups are {0, 1} and {2, 3}
        want = torch.cat([base[:40] * tp_sum, base[40:] * 10]) * 0.5
        assert torch.equal(out["reduced"], want), (out["reduced"] - want).abs().max()
        assert out["red_stats"] == {"overlapped_elems": 28, "tail_elems": 32 + 40}
        g1, g2, numel = out["flat"]
        assert torch.equal(g1, torch.full((2, 3), 6 * 0.25)) and torch.equal(g2, torch.full((4,), 60 * 0.25)) and numel == 10


def _concurrent_worker(rank, world, cuda=False):
    import torch.distributed as dist

    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.backend import native_torch

    smp.init({"tensor_parallel_degree": world, "ddp": True, **({} if cuda else {"ddp_dist_backend": "gloo"})})
    if cuda:
        torch.set_default_device(torch.device("cuda", smp.local_rank()))
    nt = native_torch.load()
    pg = smp.get_tp_process_group()           # with tp = world this is the group the gradient all-reduce uses too
    x = torch.arange(4 * 8 * 3, dtype=torch.float32).reshape(4, 8, 3) + rank
    for it in range(60):
        g1, g2 = torch.full((64,), float(rank)), torch.full((3, 5), float(it))
        bucket = nt.FlatBucket(pg, [g1, g2])                      # asynchronous all-reduce in flight ...
        other = dist.all_reduce(torch.ones(16), group=pg, async_op=True)
        y = nt.scatter_and_merge(pg, x, 1, 0)                     # ... while the all-to-all runs on the same group
        z = nt.scatter_and_merge(pg, y, 0, 1, [4] * world, [2] * world)
        assert torch.equal(z, x), it
        bucket.finish(1.0)
        other.wait()
        assert g1[0].item() == sum(range(world)) and g2[0, 0].item() == it * world
    return True


def test_all_to_all_with_a_gradient_all_reduce_in_flight():
    """The uneven all-to-all must not block against an asynchronous all-reduce that was launched earlier on the same
    group (what a backward pass under ``optimize: memory`` does); gloo's own all-to-all did - one run in ten
    deadlocked - so on CPU tensors the runtime moves the pieces point to point."""
    assert all(run_distributed(_concurrent_worker, 4, timeout=180))
