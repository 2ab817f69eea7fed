"""Microbatch splitting of step arguments and StepOutput reductions (reference test/torch/test_split.py,
test_utils.py)."""

import pytest
import torch

from smdistributed_modelparallel_b200.backend.split import StepOutput, TensorSplitter, map_structure


def _fn(model, x, y, flag=True, meta=None):
    return None


def test_split_tensors_and_nested_structures():
    sp = TensorSplitter(_fn)
    x = torch.arange(8 * 3).reshape(8, 3)
    y = {"a": torch.arange(8), "b": [torch.ones(8, 2), "text"]}
    for mb in range(4):
        args, kwargs = sp.preprocess_args((object(), x, y), {"flag": False, "meta": (1, 2)}, 4, mb)
        assert torch.equal(args[1], x[2 * mb:2 * mb + 2])
        assert torch.equal(args[2]["a"], torch.arange(8)[2 * mb:2 * mb + 2])
        assert args[2]["b"][0].shape == (2, 2) and args[2]["b"][1] == "text"
        assert kwargs == {"flag": False, "meta": (1, 2)}


def test_non_split_inputs_and_axes():
    sp = TensorSplitter(_fn, non_split_inputs=["y"], input_split_axes={"x": 1})
    x = torch.arange(2 * 6).reshape(2, 6)
    y = torch.arange(5)
    args, _ = sp.preprocess_args((None, x, y), {}, 3, 1)
    assert torch.equal(args[1], x[:, 2:4])
    assert torch.equal(args[2], y)


def test_indivisible_batch_is_an_error():
    sp = TensorSplitter(_fn)
    with pytest.raises(Exception):
        sp.preprocess_args((None, torch.zeros(7, 2), torch.zeros(7)), {}, 4, 0)


def test_step_output_reductions():
    outs = StepOutput([torch.tensor([1.0, 2.0]), torch.tensor([3.0, 4.0])])
    assert len(outs) == 2 and torch.equal(outs[1], torch.tensor([3.0, 4.0]))
    assert torch.equal(outs.reduce_mean(), torch.tensor([2.0, 3.0]))
    assert torch.equal(outs.reduce_sum(), torch.tensor([4.0, 6.0]))
    assert torch.equal(outs.concat(), torch.tensor([1.0, 2.0, 3.0, 4.0]))
    assert outs.stack().shape == (2, 2)
    doubled = outs.map(lambda t: t * 2)
    assert torch.equal(doubled.outputs[0], torch.tensor([2.0, 4.0]))


def test_map_structure_preserves_containers():
    obj = {"k": [1, (2, 3)], "z": 4}
    assert map_structure(lambda v: v + 1, obj) == {"k": [2, (3, 4)], "z": 5}
