"""Plumbing of the comparison arm (``baseline/``: the reference's hot path in stock PyTorch) on CPU / gloo: the
batch-gather tensor-parallel layers with their all-reduces, the vocabulary-parallel loss, the send / receive pipeline
with one 2-rank group per direction, DDP and the fp32-master optimizer step.  Timing happens on GPUs (bench.py --impl
baseline); here the point is that every parallel layout steps, agrees across ranks and learns."""

import types

import pytest
import torch

from dist_utils import run_distributed


def _worker(rank, world, spec):
    from baseline.engine import Runner
    from smdistributed_modelparallel_b200.models.configs import MODEL_CONFIGS

    cfg = dict(MODEL_CONFIGS["tiny"])
    args = types.SimpleNamespace(parallelism=spec, seq=16, micro_batch=2, microbatches=4, checkpoint=0, offload=0,
                                 model="tiny", layers=0)
    torch.manual_seed(0)
    r = Runner(args, cfg, world, device="cpu", dtype=torch.float32)
    r.opt.opt.param_groups[0]["lr"] = 3e-2
    g = torch.Generator().manual_seed(1 + rank)
    ids = torch.randint(0, cfg["vocab_size"], (r.per_rank_batch(), 16), generator=g)
    losses = [float(r.step(ids)) for _ in range(4)]
    return {"losses": losses, "pp_rank": r.pr, "last": r.pr == r.pp - 1}


@pytest.mark.parametrize("world,spec", [(1, "dp1"), (2, "tp2"), (2, "pp2"), (2, "dp2"), (4, "pp2tp2")])
def test_baseline_layouts_step_and_learn(world, spec):
    res = run_distributed(_worker, world, spec, timeout=300)
    last = [r for r in res if r["last"]]
    assert last
    for r in last:                         # the loss lives on the last pipeline stage
        l = r["losses"]
        assert all(x == x for x in l) and l[-1] < l[0], l
