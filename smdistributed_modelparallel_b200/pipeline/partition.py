"""Automatic assignment of the module tree to pipeline ranks.

Same problem statement and cost model as the reference's ``ModulePartitioner`` (reference
smdistributed/modelparallel/torch/module_partition.py:182-905): the unit of placement is an
``nn.Module`` (any node of the tree, not a layer list); cost of a node =
``memory_weight * memory + (1 - memory_weight) * time`` over its subtree, normalised; the tree is walked
breadth first - a node owns a set of devices, keeps the first for itself and divides the set among its
children; a child that receives one device takes its whole subtree there; children too large for one
device receive several (d'Hondt seat allocation, reference :788-835) and are divided recursively;
runs of small children are cut into contiguous segments that minimise the most loaded device
(min-max segmentation, reference :837-905).  Modules that share a parameter always land together, and
manual ``smp.partition`` / ``smp.set_partition`` assignments are respected.

Independent implementation: plain recursive functions over a light ``Node`` tree, an O(n^2 k) dynamic
programme for the segmentation that accounts for load already placed on each device, and a greedy
highest-quotient loop for d'Hondt.
"""

from __future__ import annotations

from dataclasses import dataclass, field

import torch.nn as nn

from ..backend.exceptions import ModelTooSmallError
from ..backend.logger import get_logger


@dataclass
class Node:
    module: nn.Module
    name: str
    own_cost: float = 0.0           # cost of tensors owned directly by this module
    cost: float = 0.0               # subtree cost
    children: list = field(default_factory=list)
    pinned: int | None = None       # manual assignment


def dhondt_allocate(costs, seats: int, splittable=None):
    """Seats per item by repeatedly giving a seat to the highest cost / (seats + 1) quotient.  Items that
    cannot be split further (``splittable[i]`` False) stop at one seat."""
    n = len(costs)
    alloc = [0] * n
    for _ in range(seats):
        best, best_q = None, -1.0
        for i, c in enumerate(costs):
            if splittable is not None and not splittable[i] and alloc[i] >= 1:
                continue
            q = c / (alloc[i] + 1)
            if q > best_q:
                best, best_q = i, q
        if best is None:
            break
        alloc[best] += 1
    return alloc


def get_segments(costs, k: int, base_loads=None):
    """Cut ``costs`` into ``k`` contiguous (possibly empty) segments minimising
    max_j(base_loads[j] + sum(segment j)).  Returns the segment index of every item."""
    n = len(costs)
    base = list(base_loads) if base_loads is not None else [0.0] * k
    prefix = [0.0]
    for c in costs:
        prefix.append(prefix[-1] + c)
    INF = float("inf")
    # best[j][i]: minimal max-load using the first j segments for the first i items
    best = [[INF] * (n + 1) for _ in range(k + 1)]
    cut = [[0] * (n + 1) for _ in range(k + 1)]
    best[0][0] = 0.0
    for j in range(1, k + 1):
        for i in range(n + 1):
            for s in range(i + 1):
                if best[j - 1][s] == INF:
                    continue
                load = max(best[j - 1][s], base[j - 1] + prefix[i] - prefix[s])
                if load < best[j][i] - 1e-15:
                    best[j][i] = load
                    cut[j][i] = s
    seg = [0] * n
    i = n
    for j in range(k, 0, -1):
        s = cut[j][i]
        for t in range(s, i):
            seg[t] = j - 1
        i = s
    return seg


def estimated_work(m: nn.Module, tp_size: int = 1) -> float:
    """Stand-in for a module's own execution time when no trace exists (delayed parameter initialisation, or
    tracing switched off): the weights its GEMMs multiply, per tensor-parallel rank.  A tensor-parallel shard counts as
    it is; a replicated weight is applied to this rank's 1/t of the tokens, so it counts 1/t; embedding tables are
    looked up, not multiplied, and cost nothing.  (The reference always traces, torch/patches/tracing.py:41-86; its
    fallback is parameter volume, which puts a 200 M-row embedding on a par with four sharded transformer layers.)"""
    if isinstance(m, nn.Embedding) or type(m).__name__ in ("DistributedEmbedding",):
        return 0.0
    total = 0.0
    for p in m.parameters(recurse=False):
        if p.dim() < 2:
            continue
        sharded = getattr(p, "_smp_dist_axis", None) is not None
        total += float(p.numel()) / (1.0 if sharded else float(max(1, tp_size)))
    return total


class ModulePartitioner:
    def __init__(self, root: nn.Module, num_partitions: int, memory_weight: float = 0.8, module_times=None,
                 module_memory=None, manual=None, name_of=None, tp_size: int = 1, contiguous: bool = True):
        self.root_module = root
        self.contiguous = bool(contiguous)
        self.k = num_partitions
        self.tp_size = int(tp_size)
        self.w = float(memory_weight)
        self.times = module_times or {}
        self.memory = module_memory or {}
        self.manual = manual or (lambda m: None)
        self.name_of = name_of or (lambda m: type(m).__name__)
        self.load = [0.0] * num_partitions

    # ---- tree + costs
    def _build(self, m, seen):
        node = Node(m, self.name_of(m), pinned=self.manual(m))
        own_params = sum(p.numel() * p.element_size() for p in m.parameters(recurse=False))
        own_params += sum(b.numel() * b.element_size() for b in m.buffers(recurse=False))
        node.own_mem = float(own_params) + float(self.memory.get(m, 0.0))
        for c in m.children():
            if id(c) in seen:          # a reused module is placed where it was first met
                continue
            seen.add(id(c))
            node.children.append(self._build(c, seen))
        node.sub_mem = node.own_mem + sum(c.sub_mem for c in node.children)
        node.sub_work = estimated_work(m, self.tp_size) + sum(c.sub_work for c in node.children)
        if m in self.times:
            node.sub_time = float(self.times[m])
        else:
            node.sub_time = None
        return node

    def _fill_costs(self, node, total_mem, total_time, total_work=0.0):
        for c in node.children:
            self._fill_costs(c, total_mem, total_time, total_work)
        mem = node.sub_mem / total_mem if total_mem > 0 else 0.0
        if total_time and node.sub_time is not None:
            t = node.sub_time / total_time
        elif total_time:
            t = sum((c.sub_time or 0.0) for c in node.children) / total_time
        elif total_work > 0:
            t = node.sub_work / total_work      # no trace: time estimated from the GEMM weights (estimated_work)
        else:
            t = mem
        node.cost = self.w * mem + (1.0 - self.w) * t
        node.own_cost = max(0.0, node.cost - sum(c.cost for c in node.children))

    # ---- placement
    def _assign_subtree(self, node, dev, out):
        out[node.module] = dev
        for c in node.children:
            self._assign_subtree(c, c.pinned if c.pinned is not None else dev, out)

    def _place(self, node, devices, out):
        dev0 = node.pinned if node.pinned is not None else devices[0]
        out[node.module] = dev0
        self.load[dev0] += node.own_cost
        if len(devices) == 1 or not node.children:
            for c in node.children:
                self._assign_subtree(c, c.pinned if c.pinned is not None else dev0, out)
            self.load[dev0] += sum(c.cost for c in node.children)
            return
        kids = node.children
        free = [c for c in kids if c.pinned is None]
        for c in kids:
            if c.pinned is not None:
                self._assign_subtree(c, c.pinned, out)
                self.load[c.pinned] += c.cost
        if not free:
            return
        D = len(devices)
        total = sum(c.cost for c in free)
        share = total / D if D else total
        big = [bool(c.children) and c.cost > share * 1.0001 for c in free]
        if not any(big) and not self.contiguous:
            # ``contiguous: false`` (the reference reads the key only in its TensorFlow partitioner,
            # tensorflow/auto.py:189-224): siblings may interleave over the devices - largest first onto the least
            # loaded device, which balances better than any cut of the sequence at the price of more stage crossings.
            for c in sorted(free, key=lambda c: -c.cost):
                d = min(devices, key=lambda x: self.load[x])
                self._assign_subtree(c, d, out)
                self.load[d] += c.cost
            return
        if not any(big):
            seg = get_segments([c.cost for c in free], D, [self.load[d] for d in devices])
            for c, s in zip(free, seg):
                self._assign_subtree(c, devices[s], out)
                self.load[devices[s]] += c.cost
            return
        seats = dhondt_allocate([c.cost for c in free], D, splittable=[bool(c.children) for c in free])
        ptr = 0
        plan = []
        for c, s in zip(free, seats):
            if s >= 1:
                plan.append((c, devices[ptr:ptr + s]))
                ptr += s
            else:
                plan.append((c, None))
        # seat-less children ride with the less loaded of their neighbours' edge devices
        for i, (c, devs) in enumerate(plan):
            if devs is not None:
                continue
            prev_dev = next((plan[j][1][-1] for j in range(i - 1, -1, -1) if plan[j][1]), None)
            next_dev = next((plan[j][1][0] for j in range(i + 1, len(plan)) if plan[j][1]), None)
            cands = [d for d in (prev_dev, next_dev) if d is not None] or [devices[0]]
            d = min(cands, key=lambda x: self.load[x])
            self._assign_subtree(c, d, out)
            self.load[d] += c.cost
        # recurse into multi-device children smallest first, so their neighbours' load is known
        for c, devs in sorted([p for p in plan if p[1] is not None], key=lambda p: p[0].cost):
            if len(devs) == 1:
                self._assign_subtree(c, devs[0], out)
                self.load[devs[0]] += c.cost
            else:
                self._place(c, devs, out)

    def partition(self) -> dict:
        """module -> pipeline rank for every module in the tree."""
        seen = {id(self.root_module)}
        root = self._build(self.root_module, seen)
        total_time = root.sub_time if root.sub_time else 0.0
        self._fill_costs(root, root.sub_mem, total_time, root.sub_work)
        out = {}
        self._place(root, list(range(self.k)), out)
        self._colocate_shared_parameters(out)
        used = set(out.values())
        if len(used) < self.k:
            get_logger().warning(f"auto-partition used {len(used)} of {self.k} pipeline ranks: the model has too few "
                                 "separable modules")
        return out

    def _colocate_shared_parameters(self, out):
        owner = {}
        for m, dev in list(out.items()):
            for p in m.parameters(recurse=False):
                if id(p) in owner and owner[id(p)] != dev:
                    self._move_subtree(m, owner[id(p)], out)
                else:
                    owner.setdefault(id(p), dev)

    def _move_subtree(self, m, dev, out):
        for sub in m.modules():
            out[sub] = dev


def partition_model(model_root: nn.Module, module_manager, cfg, module_times=None, module_memory=None) -> dict:
    """Name -> pipeline rank for every named module, honouring manual assignments and ``auto_partition``."""
    pp = cfg.pipeline_parallel_degree
    if not cfg.auto_partition:
        default = cfg.default_partition

        def walk(m, inherited, out):
            mine = module_manager.manual_partition_of(m)
            dev = mine if mine is not None else inherited
            out[m] = dev
            for c in m.children():
                if c not in out:
                    walk(c, dev, out)

        assignment = {}
        walk(model_root, default, assignment)
    else:
        part = ModulePartitioner(model_root, pp, cfg.memory_weight, module_times, module_memory,
                                 manual=module_manager.manual_partition_of, name_of=module_manager.get_module_name,
                                 tp_size=cfg.tensor_parallel_degree, contiguous=cfg.contiguous)
        assignment = part.partition()
    if len(list(model_root.modules())) < pp:
        raise ModelTooSmallError()
    return {module_manager.get_module_name(m): d for m, d in assignment.items() if module_manager.has_module(m)}
