"""Rank algebra for pipeline (P) x reduced-data (D) x tensor (T) parallel placement.

A placement strategy is a permutation of the letters P, D, T: the right-most letter varies fastest
over global ranks (``cluster`` = DPT, ``spread`` = TPD).  The public method names and results match
the reference's ``Ranker`` (reference smdistributed/modelparallel/backend/core.py:26-162; expected
tables in reference test/backend/test_ranker.py:9-74).  Implementation: ranks are mixed-radix numbers
whose digit order is the strategy string; every query is digit extraction / recomposition.
"""

from __future__ import annotations

from itertools import product

_ALIASES = {"cluster": "DPT", "spread": "TPD"}


class Ranker:
    def __init__(self, placement_strategy: str, rdp_size: int, pp_size: int, tp_size: int):
        order = _ALIASES.get(placement_strategy, placement_strategy)
        if sorted(order) != ["D", "P", "T"]:
            raise ValueError(f"placement strategy must be a permutation of 'PDT', got {placement_strategy!r}")
        self.order = order
        self.extent = {"P": pp_size, "D": rdp_size, "T": tp_size}
        self.size = pp_size * rdp_size * tp_size
        # stride of each axis: product of the extents to its right
        self.stride = {}
        acc = 1
        for axis in reversed(order):
            self.stride[axis] = acc
            acc *= self.extent[axis]

    # ---- digit helpers
    def coord(self, rank: int, axis: str) -> int:
        return (rank // self.stride[axis]) % self.extent[axis]

    def coords(self, rank: int) -> dict:
        return {a: self.coord(rank, a) for a in "PDT"}

    def compose(self, c: dict) -> int:
        return sum(c[a] * self.stride[a] for a in "PDT")

    def translate(self, pp_rank: int, tp_rank: int, rdp_rank: int) -> int:
        return self.compose({"P": pp_rank, "T": tp_rank, "D": rdp_rank})

    def _line(self, rank: int, axis: str) -> list[int]:
        base = rank - self.coord(rank, axis) * self.stride[axis]
        return [base + i * self.stride[axis] for i in range(self.extent[axis])]

    def _plane(self, rank: int, a: str, b: str) -> list[int]:
        """Ranks sharing every coordinate except axes a and b; outer loop over the slower axis."""
        slow, fast = (a, b) if self.order.index(a) < self.order.index(b) else (b, a)
        c = self.coords(rank)
        out = []
        for i, j in product(range(self.extent[slow]), range(self.extent[fast])):
            c2 = dict(c)
            c2[slow], c2[fast] = i, j
            out.append(self.compose(c2))
        return out

    def _plane_rank(self, rank: int, a: str, b: str) -> int:
        slow, fast = (a, b) if self.order.index(a) < self.order.index(b) else (b, a)
        return self.coord(rank, slow) * self.extent[fast] + self.coord(rank, fast)

    # ---- reference-compatible API
    def get_pp_rank(self, rank): return self.coord(rank, "P")
    def get_tp_rank(self, rank): return self.coord(rank, "T")
    def get_rdp_rank(self, rank): return self.coord(rank, "D")
    def get_dp_rank(self, rank): return self._plane_rank(rank, "T", "D")
    def get_mp_rank(self, rank): return self._plane_rank(rank, "T", "P")
    def get_pp_group(self, rank): return self._line(rank, "P")
    def get_tp_group(self, rank): return self._line(rank, "T")
    def get_rdp_group(self, rank): return self._line(rank, "D")
    def get_dp_group(self, rank): return self._plane(rank, "T", "D")
    def get_mp_group(self, rank): return self._plane(rank, "T", "P")

    def _split_plane_rank(self, plane_rank: int, a: str, b: str, want: str) -> int:
        slow, fast = (a, b) if self.order.index(a) < self.order.index(b) else (b, a)
        return plane_rank // self.extent[fast] if want == slow else plane_rank % self.extent[fast]

    def get_rdp_rank_from_dp_rank(self, dp_rank): return self._split_plane_rank(dp_rank, "T", "D", "D")
    def get_tp_rank_from_dp_rank(self, dp_rank): return self._split_plane_rank(dp_rank, "T", "D", "T")
    def get_pp_rank_from_mp_rank(self, mp_rank): return self._split_plane_rank(mp_rank, "T", "P", "P")
    def get_tp_rank_from_mp_rank(self, mp_rank): return self._split_plane_rank(mp_rank, "T", "P", "T")

    # ---- enumerations used to create process groups identically on every rank
    def all_groups(self, kind: str) -> list[list[int]]:
        getter = {"pp": self.get_pp_group, "tp": self.get_tp_group, "rdp": self.get_rdp_group,
                  "dp": self.get_dp_group, "mp": self.get_mp_group}[kind]
        seen, out = set(), []
        for r in range(self.size):
            g = getter(r)
            key = tuple(sorted(g))
            if key not in seen:
                seen.add(key)
                out.append(g)
        return out
