"""Host control plane of a multi-rank job over the C++ parameter-server transport.

The reference brings a job up from ``TF_CONFIG`` alone: every task starts one ``tf.train.Server`` on
the address the cluster lists for it and talks to the others over that transport
(/root/reference/trainer/task.py:110-127).  The GPU job needs very little from the host -- exchanging
the slab descriptors once, a few start-up barriers, the chief's checkpoint trigger, the done-queues of
``auto_stop_ps`` -- and gets it from the same C++ transport (``csrc/native/ps.cc``: named variables
and blocking queues) instead of a second communication library.

``rank_layout`` turns a cluster into ranks: ``master`` is rank 0 (the chief), ``worker`` i is rank
1 + i.  The control store lives on ps task 0 when the cluster lists ps tasks, else in the master's own
server.  :class:`TransportGroup` offers the three collectives the start-up code uses, with
``torch.distributed``'s call shapes (``barrier``, ``all_gather_object``, ``broadcast_object_list``), so
the same call sites run under ``torchrun`` (gloo) or under ``TF_CONFIG``.
"""
from __future__ import annotations

import pickle
import time
from typing import List, Optional, Tuple

import numpy as np

from ..cluster import ClusterSpec
from ..session import PSClient


def rank_layout(cluster: ClusterSpec, task_type: str, task_index: int) -> Tuple[Optional[int], int, str]:
    """(rank or None for a ps task, world, address of the control store)."""
    jobs = cluster.jobs
    n_master = cluster.num_tasks("master") if "master" in jobs else 0
    n_worker = cluster.num_tasks("worker") if "worker" in jobs else 0
    world = n_master + n_worker
    if world < 1:
        raise ValueError("TF_CONFIG cluster lists neither a master nor workers")
    if "ps" in jobs and cluster.num_tasks("ps") > 0:
        store = cluster.task_address("ps", 0)
    elif n_master:
        store = cluster.task_address("master", 0)
    else:
        store = cluster.task_address("worker", 0)
    if task_type == "master":
        rank = task_index
    elif task_type == "worker":
        rank = n_master + task_index
    else:
        rank = None
    return rank, world, store


def put_object(client: PSClient, name: str, obj) -> None:
    """Store a picklable object as a named variable of the control store."""
    blob = pickle.dumps(obj, protocol=4)
    pad = (-len(blob)) % 4
    arr = np.frombuffer(blob + b"\0" * pad, dtype=np.int32)
    client.init_var(name, np.concatenate([np.array([len(blob)], np.int32), arr]), overwrite=True)


def get_object(client: PSClient, name: str, default=None):
    """Inverse of :func:`put_object`; ``default`` when the variable does not exist (yet)."""
    if client.var_info(name) is None:
        return default
    arr = client.pull(name)
    n = int(arr[0])
    return pickle.loads(arr[1:].tobytes()[:n])


class TransportGroup:
    """Collectives of ``world`` ranks through one control store.  Every rank must make the same calls in
    the same order.  A peer that does not show up within ``timeout_s`` raises ``ConnectionError`` (a dead
    rank surfaces instead of hanging the job)."""

    def __init__(self, address: str, rank: int, world: int, timeout_s: float = 120.0, connect_retry_secs: float = 60.0,
                 tag: str = "ctl"):
        self.rank, self.world, self.tag, self.timeout_s = rank, world, tag, timeout_s
        self.address = address
        self._c = PSClient(address, timeout_ms=int(max(1.0, timeout_s) * 1000), connect_retry_secs=connect_retry_secs)
        self._seq = 0

    def close(self) -> None:
        self._c.close()

    # -- helpers -------------------------------------------------------------------------------
    def _put(self, name: str, obj) -> None:
        put_object(self._c, name, obj)

    def _get(self, name: str):
        return get_object(self._c, name)

    def _take(self, queue: str) -> None:
        deadline = time.time() + self.timeout_s
        while True:
            left = deadline - time.time()
            if left <= 0:
                raise ConnectionError("rank %d: no token on %r within %.0f s (a peer is missing or dead)"
                                      % (self.rank, queue, self.timeout_s))
            if self._c.dequeue(queue, timeout_ms=int(min(left, 1.0) * 1000) + 1) is not None:
                return

    # -- collectives (torch.distributed call shapes) ------------------------------------------------
    def barrier(self) -> None:
        """Rank 0 collects one token per peer, then releases them.  The release queue alternates between two
        names: a rank that is already in barrier k+1 must not take a release token of barrier k that a
        slower peer is still waiting for; by barrier k+2 every token of barrier k has been taken."""
        if self.world == 1:
            return
        self._nbar = getattr(self, "_nbar", 0) + 1
        q_in, q_out = "%s/barrier/in" % self.tag, "%s/barrier/out%d" % (self.tag, self._nbar & 1)
        if self.rank == 0:
            for _ in range(self.world - 1):
                self._take(q_in)
            for _ in range(self.world - 1):
                self._c.enqueue(q_out, 1)
        else:
            self._c.enqueue(q_in, 1)
            self._take(q_out)

    def all_gather_object(self, out: List, obj) -> None:
        self._seq += 1
        self._put("%s/ag/%d" % (self.tag, self.rank), (self._seq, obj))
        self.barrier()
        for r in range(self.world):
            seq, val = self._get("%s/ag/%d" % (self.tag, r))
            if seq != self._seq:
                raise RuntimeError("rank %d is at collective %d, this rank at %d" % (r, seq, self._seq))
            out[r] = val
        self.barrier()            # nobody overwrites its entry before everyone has read it

    def broadcast_object_list(self, box: List, src: int = 0) -> None:
        self._seq += 1
        name = "%s/bc" % self.tag
        if self.rank == src:
            self._put(name, (self._seq, list(box)))
        self.barrier()
        if self.rank != src:
            seq, vals = self._get(name)
            if seq != self._seq:
                raise RuntimeError("broadcast out of step: %d vs %d" % (seq, self._seq))
            box[:] = vals
        self.barrier()

    # -- named counters / queues for the job logic -----------------------------------------------------
    def client(self) -> PSClient:
        return self._c
