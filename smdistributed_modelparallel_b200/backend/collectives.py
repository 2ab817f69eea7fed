"""Python-object collectives over the native message bus.

API parity with the reference's ``CollectiveCommunicator`` (reference
smdistributed/modelparallel/backend/collectives.py:15-348): six groups (WORLD, PP, TP, RDP, DP, MP),
blocking ``broadcast / send / recv_from / allgather`` and group barriers for user code, and the
asynchronous transaction-id flavoured calls used by the runtime.  User-API traffic and internal
traffic use disjoint transaction ids (``2*id + is_user_api``), as in the reference.

Differences: messages ride the TCP bus of ``libsmpb200_core.so`` (no MPI); all-gather is one fan-out
broadcast plus N keyed receives; a waiting receive aborts with an error when a peer reports failure.
"""

from __future__ import annotations

import ctypes
import pickle
from enum import Enum
from typing import NamedTuple

from . import native_core
from .exceptions import CommGroupConfigError, InvalidTransactionIDError, SMPRuntimeError


class CommGroup(Enum):
    WORLD = 0
    PP_GROUP = 1
    TP_GROUP = 2
    RDP_GROUP = 3
    DP_GROUP = 4
    MP_GROUP = 5

    def get_rank_type(self):
        return RankType(self.value)


class RankType(Enum):
    WORLD_RANK = 0
    PP_RANK = 1
    TP_RANK = 2
    RDP_RANK = 3
    DP_RANK = 4
    MP_RANK = 5

    def get_comm_group(self):
        return CommGroup(self.value)


class TransactionIdentifier(NamedTuple):
    id: int
    is_user_api: bool

    def get(self) -> int:
        return 2 * self.id + int(self.is_user_api)


# reserved user-API transaction ids per primitive (internal ids come from ModelParallelState)
_TX_BCAST, _TX_SENDRECV, _TX_ALLGATHER, _TX_BARRIER = 1, 2, 3, 4


class CollectiveCommunicator:
    def __init__(self, core):
        self.core = core
        self.lib = native_core.load()

    # ------------------------------------------------------------------ group helpers
    def _get_global_ranks(self, comm_group, given_rank=None):
        c = self.core
        if comm_group == CommGroup.WORLD:
            return list(range(c.size()))
        if comm_group == CommGroup.PP_GROUP:
            return c.get_pp_group()
        if comm_group == CommGroup.TP_GROUP:
            return c.get_tp_group()
        if comm_group == CommGroup.RDP_GROUP:
            return c.get_rdp_group()
        if comm_group == CommGroup.DP_GROUP:
            return c.get_dp_group()
        if comm_group == CommGroup.MP_GROUP:
            return c.get_mp_group()
        raise CommGroupConfigError(f"unknown group {comm_group}")

    def _get_group_size(self, group):
        return len(self._get_global_ranks(group))

    def _to_global(self, rank, rank_type):
        c = self.core
        if rank_type == RankType.WORLD_RANK:
            return rank
        return {RankType.PP_RANK: c.pp_rank_to_rank, RankType.TP_RANK: c.tp_rank_to_rank,
                RankType.RDP_RANK: c.rdp_rank_to_rank, RankType.DP_RANK: c.dp_rank_to_rank,
                RankType.MP_RANK: c.mp_rank_to_rank}[rank_type](rank)

    def _group_rank(self, group):
        c = self.core
        return {CommGroup.WORLD: c.rank, CommGroup.PP_GROUP: c.pp_rank, CommGroup.TP_GROUP: c.tp_rank,
                CommGroup.RDP_GROUP: c.rdp_rank, CommGroup.DP_GROUP: c.dp_rank, CommGroup.MP_GROUP: c.mp_rank}[group]()

    # ------------------------------------------------------------------ blocking user API
    def broadcast(self, obj, group):
        """Send ``obj`` to every other member of ``group`` (receivers call ``recv_from``)."""
        me = self.core.rank()
        dests = [r for r in self._get_global_ranks(group) if r != me]
        tx = TransactionIdentifier(_TX_SENDRECV, True)
        self.async_bcast_global(obj, dests, tx)
        self.wait_transaction(tx)

    def send(self, obj, dest_rank, rank_type):
        tx = TransactionIdentifier(_TX_SENDRECV, True)
        self.async_send(obj, dest_rank, tx, rank_type)
        self.wait_transaction(tx)

    def recv_from(self, src_rank, rank_type):
        tx = TransactionIdentifier(_TX_SENDRECV, True)
        return self.wait_recv(src_rank, tx, rank_type)

    def allgather(self, obj, group):
        ranks = self._get_global_ranks(group)
        me = self.core.rank()
        tx = TransactionIdentifier(_TX_ALLGATHER, True)
        self.async_bcast_global(obj, [r for r in ranks if r != me], tx)
        out = [obj if r == me else self.recv_from_global(r, tx) for r in ranks]
        self.wait_transaction(tx)
        return out

    def gather(self, obj, group, dst_group_rank=0):
        ranks = self._get_global_ranks(group)
        me = self.core.rank()
        root = ranks[dst_group_rank]
        tx = TransactionIdentifier(_TX_ALLGATHER + 8, True)
        if me == root:
            return [obj if r == me else self.recv_from_global(r, tx) for r in ranks]
        self.async_send_global(obj, root, tx)
        self.wait_transaction(tx)
        return None

    def barrier(self, group=CommGroup.WORLD):
        ranks = self._get_global_ranks(group)
        me = self.core.rank()
        tx = TransactionIdentifier(_TX_BARRIER + 16 * group.value, True)
        self.async_bcast_global(1, [r for r in ranks if r != me], tx)
        for r in ranks:
            if r != me:
                self.recv_from_global(r, tx)
        self.wait_transaction(tx)

    def pp_barrier(self): self.barrier(CommGroup.PP_GROUP)
    def dp_barrier(self): self.barrier(CommGroup.DP_GROUP)
    def tp_barrier(self): self.barrier(CommGroup.TP_GROUP)
    def rdp_barrier(self): self.barrier(CommGroup.RDP_GROUP)
    def mp_barrier(self): self.barrier(CommGroup.MP_GROUP)

    # ------------------------------------------------------------------ asynchronous primitives
    def async_bcast(self, obj, transaction_id, group, server=False):
        me = self.core.rank()
        self.async_bcast_global(obj, [r for r in self._get_global_ranks(group) if r != me], transaction_id, server)

    def async_bcast_global(self, obj, dest_ranks, transaction_id, server=False):
        if not dest_ranks:
            return
        buf = self._serialize_object(obj)
        arr = (ctypes.c_int * len(dest_ranks))(*dest_ranks)
        rc = self.lib.smpc_bus_bcast(buf, len(buf), len(dest_ranks), arr, self._tx(transaction_id), int(server))
        if rc != 0:
            raise SMPRuntimeError("message bus broadcast failed")

    def async_send(self, obj, rank, transaction_id, rank_type, do_pickle=True, server=False):
        self.async_send_global(obj, self._to_global(rank, rank_type), transaction_id, do_pickle, server)

    def async_send_global(self, obj, rank, transaction_id, do_pickle=True, server=False):
        buf = self._serialize_object(obj, do_pickle)
        rc = self.lib.smpc_bus_send(buf, len(buf), int(rank), self._tx(transaction_id), int(server))
        if rc != 0:
            raise SMPRuntimeError(f"message bus send to rank {rank} failed")

    def recv_from_global(self, rank, transaction_id, depickle=True, server=False):
        return self.wait_recv_from_global(rank, transaction_id, depickle, server)

    def async_recv_from(self, rank, transaction_id, rank_type, server=False):
        # receives are always "posted": the bus thread drains sockets into the mailbox
        return None

    def async_recv_from_global(self, rank, transaction_id, server=False):
        return None

    def wait_recv(self, rank, transaction_id, rank_type, depickle=True, server=False):
        return self.wait_recv_from_global(self._to_global(rank, rank_type), transaction_id, depickle, server)

    def wait_recv_from_global(self, rank, transaction_id, depickle=True, server=False):
        tx = self._tx(transaction_id)
        n = self.lib.smpc_bus_wait(int(rank), tx, -1)
        if n == -2:
            raise SMPRuntimeError(f"rank {self.lib.smpc_bus_peer_failed()} reported failure; aborting receive")
        if n < 0:
            raise SMPRuntimeError("message bus receive failed")
        buf = ctypes.create_string_buffer(max(int(n), 1))
        got = self.lib.smpc_bus_retrieve(int(rank), tx, buf, n)
        if got != n:
            raise SMPRuntimeError("message bus retrieve failed")
        return self._deserialize_object(buf.raw[:n], depickle)

    def poll_recv_from_global(self, rank, transaction_id, server=False):
        return self.lib.smpc_bus_poll(int(rank), self._tx(transaction_id)) >= 0

    def wait_all_transactions(self, server=False):
        self.lib.smpc_bus_wait_sends(-1)

    def wait_transaction(self, transaction_id, server=False):
        self.lib.smpc_bus_wait_sends(self._tx(transaction_id))

    # ---- server channel (FIFO of control messages for the pipeline engine)
    def server_send(self, obj, rank, txid=0):
        buf = self._serialize_object(obj)
        if self.lib.smpc_bus_send(buf, len(buf), int(rank), int(txid), 1) != 0:
            raise SMPRuntimeError(f"server-channel send to rank {rank} failed")

    def server_has_message(self) -> bool:
        return self.lib.smpc_bus_server_peek(None, None, 0) >= 0

    def server_next_message(self, timeout_ms=-1):
        """(src, txid, obj) of the next server-channel message, or None on timeout."""
        src, tx = ctypes.c_int(), ctypes.c_longlong()
        n = self.lib.smpc_bus_server_peek(ctypes.byref(src), ctypes.byref(tx), int(timeout_ms))
        if n == -2:
            raise SMPRuntimeError(f"rank {self.lib.smpc_bus_peer_failed()} reported failure; aborting")
        if n < 0:
            return None
        buf = ctypes.create_string_buffer(max(int(n), 1))
        if self.lib.smpc_bus_server_pop(buf, n) != n:
            raise SMPRuntimeError("server-channel pop failed")
        return src.value, tx.value, pickle.loads(buf.raw[:n])

    # ------------------------------------------------------------------ helpers
    def _tx(self, transaction_id) -> int:
        if isinstance(transaction_id, TransactionIdentifier):
            return transaction_id.get()
        if isinstance(transaction_id, int) and transaction_id >= 0:
            return transaction_id
        raise InvalidTransactionIDError(f"bad transaction id {transaction_id!r}")

    def _validate_transaction_id(self, transaction_id, server):
        self._tx(transaction_id)

    def _deserialize_object(self, buf, depickle=True):
        return pickle.loads(buf) if depickle else buf

    def _serialize_object(self, obj, do_pickle=True):
        return pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL) if do_pickle else bytes(obj)

    def pickle_obj(self, obj):
        return pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
