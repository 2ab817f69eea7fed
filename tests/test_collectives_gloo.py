"""Uneven tensor collectives on 4 CPU/gloo ranks (reference test/torch/mpi_4ps/test_collectives.py:13-114):
all-to-all split/merge with divisible and non-divisible extents, variable-count all-gather."""

import torch

from dist_utils import run_distributed


def _worker(rank, world):
    import smdistributed_modelparallel_b200 as smp
    from smdistributed_modelparallel_b200.collectives import allgatherv_tensor, scatter_and_merge_tensor

    smp.init({"tensor_parallel_degree": world, "ddp": True, "ddp_dist_backend": "gloo"})
    out = {}
    # all-to-all: every rank holds [6 + rank, 10]; split dim 1 (10 = 3,3,2,2), merge dim 0
    x = torch.arange((6 + rank) * 10, dtype=torch.float32).reshape(6 + rank, 10) + 1000 * rank
    out["a2a"] = scatter_and_merge_tensor(x, split_axis=1, merge_axis=0, group=smp.TP_GROUP)
    # negative axes, divisible extents
    y = torch.full((2, 8, 3), float(rank))
    out["a2a_neg"] = scatter_and_merge_tensor(y, split_axis=-2, merge_axis=-1, group=smp.TP_GROUP)
    counts = [3, 1, 4, 2]
    buf = torch.zeros(sum(counts))
    off = sum(counts[:rank])
    buf[off:off + counts[rank]] = rank + 1
    out["agv"] = allgatherv_tensor(buf, counts, group=smp.TP_GROUP)
    even = torch.zeros(8)
    even[2 * rank:2 * rank + 2] = rank + 1
    out["agv_even"] = allgatherv_tensor(even, [2, 2, 2, 2], group=smp.TP_GROUP)
    return out


def test_uneven_collectives():
    res = run_distributed(_worker, 4)
    xs = [torch.arange((6 + r) * 10, dtype=torch.float32).reshape(6 + r, 10) + 1000 * r for r in range(4)]
    cols = [(0, 3), (3, 6), (6, 8), (8, 10)]
    for i, r in enumerate(res):
        want = torch.cat([x[:, cols[i][0]:cols[i][1]] for x in xs], dim=0)
        assert torch.equal(r["a2a"], want), i
        want_neg = torch.cat([torch.full((2, 2, 3), float(k)) for k in range(4)], dim=-1)
        assert torch.equal(r["a2a_neg"], want_neg)
        assert r["agv"].tolist() == [1, 1, 1, 2, 3, 3, 3, 3, 4, 4]
        assert r["agv_even"].tolist() == [1, 1, 2, 2, 3, 3, 4, 4]
