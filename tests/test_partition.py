"""Auto-partitioner pieces on mock inputs (reference test/torch/test_child_partition.py, test_partition.py):
d'Hondt seat allocation, min-max contiguous segmentation, whole-model partitioning of toy modules and
manual pins."""

import torch
import torch.nn as nn

from smdistributed_modelparallel_b200.pipeline.partition import ModulePartitioner, dhondt_allocate, get_segments


def test_dhondt_proportional():
    assert dhondt_allocate([100, 50, 25], 7) == [4, 2, 1]
    assert dhondt_allocate([1, 1, 1, 1], 4) == [1, 1, 1, 1]
    assert sum(dhondt_allocate([3, 2, 1], 10)) == 10
    # an unsplittable item never takes more than one seat
    assert dhondt_allocate([100, 1], 3, splittable=[False, True]) == [1, 2]
    assert dhondt_allocate([5.0], 0) == [0]


def test_segments_minimise_the_maximum_load():
    seg = get_segments([1, 1, 1, 1, 1, 1], 3)
    assert seg == [0, 0, 1, 1, 2, 2]
    seg = get_segments([5, 1, 1, 1, 1, 1], 2)
    assert seg == [0, 1, 1, 1, 1, 1]
    seg = get_segments([1, 2, 3, 4, 5, 6, 7, 8, 9], 3)
    loads = [sum(c for c, s in zip(range(1, 10), seg) if s == j) for j in range(3)]
    assert max(loads) == 17                                   # {1..5}=15 {6,7}=13 {8,9}=17
    assert seg == sorted(seg)                                 # contiguous
    # existing load on a device pushes work to the others
    seg = get_segments([1, 1, 1, 1], 2, base_loads=[3, 0])
    assert seg.count(1) > seg.count(0)
    assert get_segments([], 3) == []


class _Toy(nn.Module):
    def __init__(self, widths):
        super().__init__()
        self.blocks = nn.Sequential(*[nn.Linear(w, w) for w in widths])
        self.head = nn.Linear(widths[-1], 4)

    def forward(self, x):
        return self.head(self.blocks(x))


def _loads(assign, model):
    loads = {}
    for m, dev in assign.items():
        loads[dev] = loads.get(dev, 0) + sum(p.numel() for p in m.parameters(recurse=False))
    return loads


def test_partition_balances_parameters():
    torch.manual_seed(0)
    model = _Toy([64] * 8)
    assign = ModulePartitioner(model, 2, memory_weight=1.0).partition()
    assert assign[model] == 0                                 # the root lives on the first stage
    for m in model.modules():
        assert m in assign
    loads = _loads(assign, model)
    assert set(loads) == {0, 1}
    assert max(loads.values()) / sum(loads.values()) < 0.7
    devs = [assign[b] for b in model.blocks]
    assert devs == sorted(devs)                               # sequential children stay contiguous


def test_partition_four_way_and_manual_pin():
    model = _Toy([32] * 12)
    assign = ModulePartitioner(model, 4, memory_weight=1.0).partition()
    assert set(assign[b] for b in model.blocks) == {0, 1, 2, 3}
    pinned = model.blocks[0]
    assign = ModulePartitioner(model, 2, memory_weight=1.0, manual=lambda m: 1 if m is pinned else None).partition()
    assert assign[pinned] == 1


def test_shared_parameters_are_colocated():
    class Tied(nn.Module):
        def __init__(self):
            super().__init__()
            self.emb = nn.Embedding(50, 16)
            self.mid = nn.Sequential(*[nn.Linear(16, 16) for _ in range(6)])
            self.out = nn.Linear(16, 50, bias=False)
            self.out.weight = self.emb.weight

        def forward(self, x):
            return self.out(self.mid(self.emb(x)))

    model = Tied()
    assign = ModulePartitioner(model, 2, memory_weight=1.0).partition()
    assert assign[model.emb] == assign[model.out]


def test_untraced_time_estimate_ignores_embedding_tables():
    """Without a trace (delayed parameter initialisation) the time half of the cost comes from the GEMM weights: an
    embedding table is a lookup and must not weigh like layers, and a replicated output projection counts 1/tp."""
    import torch.nn as nn

    from smdistributed_modelparallel_b200.pipeline.partition import ModulePartitioner, estimated_work

    class LM(nn.Module):
        def __init__(self):
            super().__init__()
            self.emb = nn.Embedding(4096, 64)                      # 262k parameters, no GEMM
            self.layers = nn.Sequential(*[nn.Linear(64, 64, bias=False) for _ in range(16)])   # 4k each
            self.head = nn.Linear(64, 128, bias=False)             # 8k = two layers' worth

    m = LM()
    assert estimated_work(m.emb) == 0.0 and estimated_work(m.head, tp_size=4) == 64 * 128 / 4
    # time-weighted: 16 + 2 units split 9 / 9 -> the stage holding the embedding still gets ~9 layers
    assign = ModulePartitioner(m, 2, memory_weight=0.0).partition()
    n0 = sum(1 for l in m.layers if assign[l] == 0)
    assert 8 <= n0 <= 10, n0
    # memory-weighted (the reference's fallback): the embedding table pushes almost every layer to the other stage
    assign = ModulePartitioner(m, 2, memory_weight=1.0).partition()
    assert sum(1 for l in m.layers if assign[l] == assign[m.emb]) <= 2


def test_contiguous_false_interleaves_siblings_for_balance():
    """Costs 5,5,1,1 over two devices: the best contiguous cut is 5 | 7; interleaving gives 6 | 6."""
    import torch.nn as nn

    from smdistributed_modelparallel_b200.pipeline.partition import ModulePartitioner

    def build():
        return nn.Sequential(nn.Linear(50, 100, bias=False), nn.Linear(50, 100, bias=False),
                             nn.Linear(10, 100, bias=False), nn.Linear(10, 100, bias=False))

    def loads(assign, net):
        out = [0, 0]
        for m in net:
            out[assign[m]] += m.weight.numel()
        return out

    net = build()
    a = ModulePartitioner(net, 2, memory_weight=1.0).partition()
    assert [a[m] for m in net] == [0, 1, 1, 1]
    assert max(loads(a, net)) == 7000
    net = build()
    b = ModulePartitioner(net, 2, memory_weight=1.0, contiguous=False).partition()
    assert loads(b, net) == [6000, 6000]
