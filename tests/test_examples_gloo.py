"""The user-facing example (examples/train_gpt.py - a HuggingFace model trained through the reference's import path
``smdistributed.modelparallel.torch``) on CPU / gloo: tensor x pipeline parallel run with a checkpoint, then a resumed
run that continues where the first stopped; a sharded-data-parallel run with activation checkpointing."""

import os
import sys
import tempfile

import pytest

transformers = pytest.importorskip("transformers")

from dist_utils import ROOT, run_distributed


def _run(rank, world, argv):
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import train_gpt

    return train_gpt.main(argv)


def test_example_tp2_pp2_checkpoint_and_resume():
    with tempfile.TemporaryDirectory() as d:
        base = ["--model", "gpt2-tiny", "--tensor-parallel-degree", "2", "--pipeline-parallel-degree", "2",
                "--microbatches", "2", "--checkpoint-dir", d]
        first = run_distributed(_run, 4, base + ["--steps", "4", "--save-every", "2"], timeout=420)
        again = run_distributed(_run, 4, base + ["--steps", "2", "--resume"], timeout=420)
        straight = run_distributed(_run, 4, ["--model", "gpt2-tiny", "--tensor-parallel-degree", "2",
                                             "--pipeline-parallel-degree", "2", "--microbatches", "2", "--steps", "6"],
                                   timeout=420)
    for r in first:
        assert r["losses"][-1] < r["losses"][0]
    for a, s in zip(again, straight):
        assert a["start"] == 4                                   # newest checkpoint: step 4
        # resumed steps 4 and 5 equal steps 4 and 5 of an uninterrupted run (weights and Adam state restored)
        assert a["losses"] == pytest.approx(s["losses"][4:6], abs=1e-4), (a["losses"], s["losses"])


def test_example_sharded_data_parallel_with_checkpointing():
    res = run_distributed(_run, 2, ["--model", "gptj-tiny", "--sharded-data-parallel-degree", "2",
                                    "--activation-checkpointing", "--steps", "4"], timeout=420)
    for r in res:
        assert r["losses"][-1] < r["losses"][0] - 0.05, r["losses"]
