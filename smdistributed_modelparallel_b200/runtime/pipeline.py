"""Microbatch schedules used by the step leader.

Public behaviour follows the reference (reference smdistributed/modelparallel/torch/pipeline.py:9-145;
state sequences pinned by reference test/torch/test_pipeline.py:11-206):

* ``simple``      - every forward first, then backwards in index order;
* ``interleaved`` - a microbatch that is ready for backward always wins over starting a new forward
                    (1F1B-like in steady state);
* ``_only_forward`` - forwards only (evaluation / tests).

The cap on simultaneously active microbatches (``active_microbatches``) is applied by the executor
when it asks for the next microbatch (``in_flight()`` helps it), as in the reference where the server
queue enforces it (reference torch/server_queue.py:641-675).
"""

from __future__ import annotations

from enum import Enum


class MbStatus(Enum):
    READY_FOR_FWD = 0
    READY_FOR_BWD = 1
    DONE = 2
    FWD = 3
    BWD = 4


class PTPipeline:
    def __init__(self, num_microbatches=None, active_microbatches=None):
        self._n = num_microbatches
        self.active_microbatches = active_microbatches
        self.init_step()

    def _num(self):
        if self._n is not None:
            return self._n
        from ..state_mod import state

        return state.num_microbatches()

    def init_step(self):
        self.status = [MbStatus.READY_FOR_FWD] * self._num()
        self.mb_pointer = None

    # ---- transitions
    def mark_done(self, mb):
        self.status[mb] = MbStatus.DONE

    def mark_ready_for_backward(self, mb):
        self.status[mb] = MbStatus.READY_FOR_BWD

    def promote_status(self, mb):
        cur = self.status[mb]
        if cur is MbStatus.READY_FOR_FWD:
            self.status[mb] = MbStatus.FWD
        elif cur is MbStatus.READY_FOR_BWD:
            self.status[mb] = MbStatus.BWD

    # ---- queries
    def get_status(self, mb):
        return self.status[mb]

    def has_more_ticks(self):
        return any(s is not MbStatus.DONE for s in self.status)

    def in_flight(self):
        """Microbatches that hold activations: forward started, backward not finished."""
        return sum(s in (MbStatus.FWD, MbStatus.READY_FOR_BWD, MbStatus.BWD) for s in self.status)

    def _first(self, wanted, after=None):
        """Lowest microbatch index in one of the wanted states, scanning cyclically from after+1."""
        n = len(self.status)
        start = 0 if after is None else after + 1
        for k in range(n):
            mb = (start + k) % n
            if self.status[mb] in wanted:
                return mb
        return None

    def get_next_mb_with_status(self, status, after=None):
        wanted = tuple(status) if isinstance(status, (list, tuple)) else (status,)
        return self._first(wanted, after)

    def _pick(self):
        raise NotImplementedError

    def get_next_microbatch(self):
        mb = self._pick()
        if mb is not None:
            self.mb_pointer = mb
        return mb


class OnlyForwardPipeline(PTPipeline):
    def _pick(self):
        return self._first((MbStatus.READY_FOR_FWD,), self.mb_pointer)


class SimplePipeline(PTPipeline):
    def _pick(self):
        mb = self._first((MbStatus.READY_FOR_FWD,), self.mb_pointer)
        if mb is None and self._first((MbStatus.FWD,)) is None:
            mb = self._first((MbStatus.READY_FOR_BWD,))
        return mb


class InterleavedPipeline(PTPipeline):
    def _pick(self):
        mb = self._first((MbStatus.READY_FOR_BWD,))
        if mb is None:
            mb = self._first((MbStatus.READY_FOR_FWD,))
        return mb
