"""Addressing and bookkeeping of pipeline tensor transfers.

``LinkManager`` hands out link ids: a link names one reusable transfer slot of a given byte size between this
pipeline rank and a peer.  Ids are recycled per step (``reset``) and a request is served by the smallest free
link that is large enough, so a steady-state step touches the same few links every time - the per-link counters
(``stats``) are what the timeline and the memory report show for pipeline traffic.  Every pipeline rank draws
from its own range of 65 536 ids, so ids never collide between ranks (reference torch/links.py:21-84).

``HandleManager`` maps a transfer handle to (direction, tensors): the tensors of an asynchronous send or receive
must stay referenced until the transfer has completed (reference torch/handle_manager.py:1-24).  ``PipelineP2P``
(runtime/server.py) keeps its in-flight NCCL / gloo work objects here.
"""

from __future__ import annotations

import bisect
import itertools

IDS_PER_RANK = 65536


class LinkManager:
    def __init__(self, pp_rank: int = 0):
        self.pp_rank = int(pp_rank)
        self._next = 0
        self._sizes: list = []          # sorted sizes for which a link exists
        self._free: dict = {}           # size -> [link ids not in use this step]
        self._busy: dict = {}           # link id -> size
        self.stats: dict = {}           # link id -> [uses, bytes]

    def _new_id(self):
        self._next = (self._next + 1) % IDS_PER_RANK
        return self._next + IDS_PER_RANK * self.pp_rank

    def get_link(self, nbytes: int) -> int:
        """The smallest free link of at least ``nbytes``; a new one of exactly that size if none is free."""
        nbytes = int(nbytes)
        i = bisect.bisect_left(self._sizes, nbytes)
        link = None
        while i < len(self._sizes):
            free = self._free.get(self._sizes[i])
            if free:
                link, size = free.pop(), self._sizes[i]
                break
            i += 1
        if link is None:
            link, size = self._new_id(), nbytes
            if nbytes not in self._free:
                bisect.insort(self._sizes, nbytes)
                self._free[nbytes] = []
        self._busy[link] = size
        st = self.stats.setdefault(link, [0, 0])
        st[0] += 1
        st[1] += nbytes
        return link

    def link_size(self, link: int) -> int:
        return self._busy[link]

    def in_use(self) -> int:
        return len(self._busy)

    def wasted_bytes(self) -> int:
        """Bytes of links that exist but were not needed in the step that just ended."""
        return sum(size * len(ids) for size, ids in self._free.items())

    def reset(self):
        """End of step: every link becomes available again."""
        for link, size in self._busy.items():
            self._free[size].append(link)
        self._busy.clear()


class HandleManager:
    SEND, RECV = "send", "recv"

    def __init__(self):
        self._ids = itertools.count(1)
        self._map: dict = {}

    def reset(self):
        self._map.clear()

    def register_handle(self, work, direction, tensors, link=None) -> int:
        h = next(self._ids)
        self._map[h] = (work, direction, tensors, link)
        return h

    def is_valid_handle(self, h) -> bool:
        return h in self._map

    def get_handle_direction(self, h):
        return self._map[h][1]

    def get_handle_output(self, h):
        return self._map[h][2]

    def get_handle_link(self, h):
        return self._map[h][3]

    def poll(self, h) -> bool:
        return self._map[h][0].is_completed()

    def wait(self, h):
        self._map[h][0].wait()

    def clear_handle(self, h):
        del self._map[h]

    def handles(self, direction=None):
        return [h for h, (_, d, _, _) in self._map.items() if direction is None or d == direction]

    def reap(self, direction=None):
        """Drop the handles whose transfer has completed; returns how many are still in flight."""
        for h in self.handles(direction):
            if self.poll(h):
                del self._map[h]
        return len(self._map)
