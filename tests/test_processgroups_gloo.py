"""Process groups against their rank lists on eight gloo ranks (pp2 x tp2 x rdp2): what ``smp.get_*_process_group``
gathers is what ``smp.get_*_group`` lists and what the object collective over the matching ``CommGroup`` returns; group
ranks equal ``smp.*_rank()`` (reference test/torch/mpi_hybrid/test_processgroups.py:13-60, extended to the tp / rdp / mp
groups and to both placement strategies)."""

import pytest
import torch

from dist_utils import run_distributed


def _worker(rank, world, placement):
    import torch.distributed as dist

    import smdistributed_modelparallel_b200 as smp

    smp.init({"pipeline_parallel_degree": 2, "tensor_parallel_degree": 2, "ddp": True, "microbatches": 2,
              "ddp_dist_backend": "gloo", "placement_strategy": placement})
    assert smp.rank() == dist.get_rank() == rank and smp.size() == world
    out = {"coords": (smp.pp_rank(), smp.tp_rank(), smp.rdp_rank())}
    checks = [("dp", smp.get_dp_process_group(), smp.get_dp_group(), smp.CommGroup.DP_GROUP, smp.dp_rank()),
              ("pp", smp.get_pp_process_group(), smp.get_pp_group(), smp.CommGroup.PP_GROUP, smp.pp_rank()),
              ("tp", smp.get_tp_process_group(), smp.get_tp_group(), smp.CommGroup.TP_GROUP, smp.tp_rank()),
              ("rdp", smp.get_rdp_process_group(), smp.get_rdp_group(), smp.CommGroup.RDP_GROUP, smp.rdp_rank()),
              ("mp", smp.get_mp_process_group(), smp.get_mp_group(), smp.CommGroup.MP_GROUP, smp.mp_rank()),
              ("world", smp.get_world_process_group(), list(range(world)), smp.CommGroup.WORLD, rank)]
    for name, pg, ranks, comm_group, my_group_rank in checks:
        got = [torch.zeros(1) for _ in ranks]
        dist.all_gather(got, torch.ones(1) * (rank + 1), group=pg)
        seen = [int(t.item()) - 1 for t in got]
        assert seen == list(ranks), (name, seen, ranks)
        assert smp.allgather(rank, comm_group) == list(ranks), name
        assert dist.get_rank(group=pg) == my_group_rank == list(ranks).index(rank), name
        out[name] = list(ranks)
    # the getters also answer for other coordinates
    assert smp.get_dp_group(pp_rank=smp.pp_rank()) == out["dp"]
    assert smp.get_tp_group(pp_rank=smp.pp_rank(), rdp_rank=smp.rdp_rank()) == out["tp"]
    other = smp.get_dp_group(pp_rank=1 - smp.pp_rank())
    assert len(other) == 4 and not set(other) & set(out["dp"])
    smp.barrier()
    return out


@pytest.mark.parametrize("placement", ["cluster", "spread"])
def test_groups_match_their_rank_lists(placement):
    res = run_distributed(_worker, 8, placement, timeout=420)
    for kind, size, count in (("pp", 2, 4), ("tp", 2, 4), ("rdp", 2, 4), ("dp", 4, 2), ("mp", 4, 2)):
        groups = {tuple(r[kind]) for r in res}
        assert len(groups) == count and all(len(g) == size for g in groups), (kind, groups)
        assert sorted(x for g in groups for x in g) == list(range(8))          # a partition of the world
    assert len({r["coords"] for r in res}) == 8
    if placement == "cluster":        # tensor-parallel neighbours are adjacent ranks
        assert all(abs(r["tp"][0] - r["tp"][1]) == 1 for r in res)
