"""Rank algebra vs hand-computed tables (reference test/backend/test_ranker.py:9-74: pp=10, tp=6, rdp=4
for the PDT and TPD placements)."""

import pytest

from smdistributed_modelparallel_b200.backend.topology import Ranker

PP, TP, RDP = 10, 6, 4


def test_pdt_tables():
    r = Ranker("PDT", RDP, PP, TP)          # T fastest, then D, then P
    assert r.size == 240
    rank = 100                               # 100 = p*24 + d*6 + t -> p=4, d=0, t=4
    assert (r.get_pp_rank(rank), r.get_rdp_rank(rank), r.get_tp_rank(rank)) == (4, 0, 4)
    assert r.get_tp_group(rank) == [96, 97, 98, 99, 100, 101]
    assert r.get_rdp_group(rank) == [100, 106, 112, 118]
    assert r.get_pp_group(rank) == [4 + 24 * p for p in range(10)]
    assert r.get_dp_group(rank) == list(range(96, 120))
    assert r.get_dp_rank(rank) == 4
    mp = r.get_mp_group(rank)
    assert len(mp) == PP * TP and rank in mp
    assert mp[:6] == [0, 1, 2, 3, 4, 5] and mp[6:12] == [24, 25, 26, 27, 28, 29]
    assert r.get_mp_rank(rank) == 4 * 6 + 4
    assert r.translate(4, 4, 0) == rank


def test_tpd_tables():
    r = Ranker("TPD", RDP, PP, TP)          # D fastest, then P, then T
    rank = 100                               # 100 = t*40 + p*4 + d -> t=2, p=5, d=0
    assert (r.get_tp_rank(rank), r.get_pp_rank(rank), r.get_rdp_rank(rank)) == (2, 5, 0)
    assert r.get_rdp_group(rank) == [100, 101, 102, 103]
    assert r.get_pp_group(rank) == [80 + 4 * p for p in range(10)]
    assert r.get_tp_group(rank) == [20 + 40 * t for t in range(6)]
    dp = r.get_dp_group(rank)
    assert dp == [40 * t + 20 + d for t in range(6) for d in range(4)]
    assert r.get_dp_rank(rank) == 2 * 4 + 0


@pytest.mark.parametrize("strategy", ["cluster", "spread", "PDT", "PTD", "DPT", "DTP", "TPD", "TDP"])
def test_groups_partition_the_world(strategy):
    r = Ranker(strategy, 2, 3, 2)
    for kind in ("pp", "tp", "rdp", "dp", "mp"):
        groups = r.all_groups(kind)
        flat = sorted(x for g in groups for x in g)
        assert flat == list(range(12)), (strategy, kind)
    for rank in range(12):
        assert r.translate(r.get_pp_rank(rank), r.get_tp_rank(rank), r.get_rdp_rank(rank)) == rank
        dp_rank = r.get_dp_rank(rank)
        assert r.get_rdp_rank_from_dp_rank(dp_rank) == r.get_rdp_rank(rank)
        assert r.get_tp_rank_from_dp_rank(dp_rank) == r.get_tp_rank(rank)
        mp_rank = r.get_mp_rank(rank)
        assert r.get_pp_rank_from_mp_rank(mp_rank) == r.get_pp_rank(rank)
        assert r.get_tp_rank_from_mp_rank(mp_rank) == r.get_tp_rank(rank)
        assert r.get_dp_group(rank)[dp_rank] == rank
        assert r.get_mp_group(rank)[mp_rank] == rank


def test_aliases_and_errors():
    assert Ranker("cluster", 2, 2, 2).order == "DPT"
    assert Ranker("spread", 2, 2, 2).order == "TPD"
    with pytest.raises(ValueError):
        Ranker("PPT", 2, 2, 2)
