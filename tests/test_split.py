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


def test_split_then_concat_is_the_identity_for_random_nestings():
    """Property: for random nested structures of tensors (random split axes, non-split arguments, custom objects with
    ``smp_slice``), the microbatches of every tensor, concatenated along its axis, give the tensor back, and
    non-split values arrive untouched in every microbatch."""
    import random

    class Sliced:                                  # an argument that slices itself
        def __init__(self, t):
            self.t = t

        def smp_slice(self, num_mb, mb, axis):
            n = self.t.shape[axis] // num_mb
            return Sliced(self.t.narrow(axis, mb * n, n))

    def fn(model, a, b, c, keep, obj):
        return None

    for seed in range(40):
        rng = random.Random(seed)
        num_mb = rng.choice([1, 2, 3, 4])
        batch = num_mb * rng.randint(1, 3)
        axis_b = rng.choice([0, 1])

        def tensor(axis=0):
            shape = [rng.randint(1, 4), rng.randint(1, 4), rng.randint(1, 3)]
            shape[axis] = batch
            return torch.randn(*shape)

        a = {"x": tensor(), "y": [tensor(), (tensor(), "s", 3)], "z": None}
        b = tensor(axis_b)
        c = [tensor(), {"deep": {"deeper": tensor()}}]
        keep = torch.randn(5, 2)                   # declared non-split
        obj = Sliced(tensor())
        sp = TensorSplitter(fn, non_split_inputs=["keep"], input_split_axes={"b": axis_b})
        pieces = [sp.preprocess_args((None, a, b, c), {"keep": keep, "obj": obj}, num_mb, mb) for mb in range(num_mb)]

        def leaves(o, out):
            map_structure(lambda t: out.append(t) if torch.is_tensor(t) else None, o)
            return out

        for idx, (orig, ax) in enumerate(((a, 0), (b, axis_b), (c, 0))):
            want = leaves(orig, [])
            got = [leaves(p[0][idx + 1], []) for p in pieces]
            for j, w in enumerate(want):
                assert torch.equal(torch.cat([g[j] for g in got], dim=ax), w), (seed, idx, j)
        for p in pieces:
            assert p[1]["keep"] is keep
            assert p[0][1]["y"][1][1:] == ("s", 3) and p[0][1]["z"] is None
        assert torch.equal(torch.cat([p[1]["obj"].t for p in pieces], dim=0), obj.t)
        all_a, all_k = sp.preprocess_args_all_mbs((None, a, b, c), {"keep": keep, "obj": obj}, num_mb)
        for mb in range(num_mb):
            assert torch.equal(all_a[mb][2], pieces[mb][0][2]) and all_k[mb]["keep"] is keep
