"""Structure and autograd-graph helpers (reference test/torch/test_utils.py tier: flatten / unflatten of nested
containers, graph ancestor walks)."""

import collections

import torch

from smdistributed_modelparallel_b200.utils.graph import count_nodes, get_ancestors, log_backward_graph, \
    unused_parameters
from smdistributed_modelparallel_b200.utils.structure import check_requires_grad, convert_args_to_device, \
    dtype_size, flatten_into_args, flatten_structure, flattened, get_devices_tensors_in_obj, \
    get_size_tensors_in_obj, unflatten_from_args, unflatten_structure

Point = collections.namedtuple("Point", ["x", "y"])


def test_flatten_round_trip_keeps_container_types():
    dd = collections.defaultdict(list)
    dd["k"] = [1, 2]
    s = {"a": [1, (2, 3)], "p": Point(4, {"z": 5}), "o": collections.OrderedDict(b=6, a=7), "d": dd, "n": None,
         "e": [], "s": frozenset({8})}
    flat, spec = flatten_structure(s)
    assert flat == [1, 2, 3, 4, 5, 6, 7, 1, 2, None, 8]
    back = unflatten_structure(flat, spec)
    assert back == s
    assert isinstance(back["p"], Point) and isinstance(back["o"], collections.OrderedDict)
    assert isinstance(back["d"], collections.defaultdict) and back["d"].default_factory is list
    assert isinstance(back["a"][1], tuple) and isinstance(back["s"], frozenset)
    # the spec carries no leaves: any list of the same length can be poured into it
    doubled = unflatten_structure([None if x is None else 2 * x for x in flat], spec)
    assert doubled["a"] == [2, (4, 6)] and doubled["p"].y == {"z": 10}


def test_integer_leaves_are_not_confused_with_positions():
    flat, spec = flatten_structure([5, [0, 1]])
    assert flat == [5, 0, 1]
    assert unflatten_structure(["a", "b", "c"], spec) == ["a", ["b", "c"]]


def test_flattened_context_and_arg_packing():
    with flattened({"a": (1, 2), "b": 3}) as flat:
        for i in range(len(flat)):
            flat[i] += 10
        assert flat.rebuild() == {"a": (11, 12), "b": 13}
    all_args, order = flatten_into_args((1, 2), {"x": 3, "y": 4})
    assert all_args == (1, 2, 3, 4) and order == ["x", "y"]
    assert unflatten_from_args(all_args, order) == ((1, 2), {"x": 3, "y": 4})
    assert unflatten_from_args((1,), []) == ((1,), {})


def test_tensor_accounting():
    a, b = torch.zeros(4, 3), torch.zeros(5, dtype=torch.bfloat16, requires_grad=True)
    obj = {"a": a, "n": [b, "text", 3]}
    assert get_size_tensors_in_obj(obj) == 4 * 3 * 4 + 5 * 2
    assert get_devices_tensors_in_obj(obj) == {torch.device("cpu")}
    assert check_requires_grad(obj) and not check_requires_grad([a, 1])
    assert dtype_size(torch.float16) == 2 and dtype_size(torch.int64) == 8
    args, kwargs = convert_args_to_device((a,), {"k": [b]}, torch.device("cpu"))
    assert args[0] is a and kwargs["k"][0] is b


def test_graph_ancestors_and_unused_parameters():
    torch.manual_seed(0)
    net = torch.nn.ModuleDict({"used": torch.nn.Linear(4, 4), "also": torch.nn.Linear(4, 2),
                               "unused": torch.nn.Linear(4, 4)})
    x = torch.randn(3, 4, requires_grad=True)            # a boundary leaf, as a pipeline stage would create
    h = net["used"](x)
    y = net["also"](torch.relu(h))
    params, leaves = get_ancestors(y)
    assert {id(p) for p in params} == {id(p) for n in ("used", "also") for p in net[n].parameters()}
    assert [id(t) for t in leaves] == [id(x)]
    # stopping at an intermediate tensor hides everything above it
    params, leaves = get_ancestors([y], upto_inputs=[h])
    assert {id(p) for p in params} == {id(p) for p in net["also"].parameters()} and not leaves
    assert unused_parameters(net, y) == ["unused.weight", "unused.bias"]
    hist = count_nodes(y)
    assert hist["AccumulateGrad"] == 5 and hist.get("AddmmBackward0", 0) == 2
    lines = log_backward_graph(y.grad_fn)
    assert lines[0].startswith("AddmmBackward0") and any("(parameter)" in ln for ln in lines)
    assert get_ancestors(torch.zeros(2)) == (set(), set())


def test_rng_manager_leaves_the_global_generators_alone():
    """smp.init seeds the tensor-parallel side stream; the default generators keep the seed the user gave them."""
    import torch

    from smdistributed_modelparallel_b200.utils.random import RngManager

    torch.manual_seed(4321)
    want = torch.rand(3)
    torch.manual_seed(4321)
    mgr = RngManager()
    mgr.set_state(99)
    assert torch.equal(torch.rand(3), want)                     # set_state drew nothing and reseeded nothing
    torch.manual_seed(4321)
    with mgr.consistent_rng_state():
        a = torch.rand(3)
    assert torch.equal(torch.rand(3), want)                     # the block used its own stream
    assert torch.equal(a, torch.rand(3, generator=torch.Generator().manual_seed(99)))
    with mgr.consistent_rng_state():
        b = torch.rand(3)
    assert not torch.equal(a, b)                                # the side stream advances between blocks
    with mgr.consistent_rng_state(seed=99):
        assert torch.equal(torch.rand(3), a)                    # reseeding restarts it


def test_gemm_operand_repacking_keeps_the_product():
    """ops.gemm.aligned_operands: K-major copies with the reduction dimension zero-padded to a multiple of 8 give the
    same product for every storage layout, 2-D and batched (the re-packed operands are what the tcgen05 kernel gets
    when a row is not a 16-byte multiple)."""
    import torch

    from smdistributed_modelparallel_b200.ops.gemm import aligned_operands, gemm

    torch.manual_seed(0)
    for batch in ((), (3,)):
        A, B = torch.randn(*batch, 5, 13), torch.randn(*batch, 7, 13)
        want = A @ B.transpose(-1, -2)
        for a_mn in (False, True):
            for b_mn in (False, True):
                a = A.transpose(-1, -2).contiguous() if a_mn else A
                b = B.transpose(-1, -2).contiguous() if b_mn else B
                Ap, Bp = aligned_operands(a, b, a_mn, b_mn)
                assert Ap.shape == (*batch, 5, 16) and Bp.shape == (*batch, 7, 16)
                assert Ap.is_contiguous() and Bp.is_contiguous()
                assert torch.allclose(Ap @ Bp.transpose(-1, -2), want, atol=1e-6)
                # on CPU the front-end answers with the plain formulation, whatever the alignment
                assert torch.allclose(gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn), want, atol=1e-5)
