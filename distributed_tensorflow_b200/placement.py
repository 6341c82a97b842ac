"""Variable placement over parameter-server tasks.

Two schemes:

* :class:`ReplicaDeviceSetter` -- what ``tf.train.replica_device_setter(cluster=...)`` does in the
  reference (/root/reference/trainer/task.py:124-127): each *whole* variable goes to
  ``/job:ps/task:k`` with ``k`` advancing round-robin in variable creation order; optimizer slots are
  colocated with their primary.  With the reference graph and two ps tasks: ``weight -> ps0``,
  ``bias -> ps1``, ``global_step -> ps0``.
* :class:`RowShard` / :class:`RangeShard` -- intra-variable partitioning used on the GPU box, where
  every rank is a ps shard living in its own HBM: embedding row ``r`` is owned by rank ``r % R``
  (local row ``r // R``), dense parameters are range-partitioned over a flat index space.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Tuple

from .cluster import ClusterSpec


class ReplicaDeviceSetter:
    def __init__(self, cluster: ClusterSpec = None, ps_tasks: int = None, worker_device: str = "/job:worker",
                 ps_device: str = "/job:ps"):
        if ps_tasks is None:
            ps_tasks = cluster.num_tasks("ps") if cluster is not None and "ps" in cluster.jobs else 0
        self.ps_tasks, self.worker_device, self.ps_device = int(ps_tasks), worker_device, ps_device
        self._next = 0
        self._placed: Dict[str, int] = {}

    def place(self, var_name: str) -> str:
        """Device string for a new variable (round-robin); repeat calls return the same answer."""
        if self.ps_tasks == 0:
            return self.worker_device
        if var_name not in self._placed:
            self._placed[var_name] = self._next
            self._next = (self._next + 1) % self.ps_tasks
        return "%s/task:%d" % (self.ps_device, self._placed[var_name])

    def task_of(self, var_name: str) -> int:
        """ps task index of a placed variable; slot names ``<var>/<Slot>`` follow their primary."""
        if var_name in self._placed:
            return self._placed[var_name]
        base = var_name.rsplit("/", 1)[0]
        if base in self._placed:
            return self._placed[base]
        raise KeyError(var_name)

    def assignment(self) -> Dict[str, int]:
        return dict(self._placed)


def replica_device_setter(cluster: ClusterSpec = None, worker_device: str = "/job:worker", ps_tasks: int = None,
                          ps_device: str = "/job:ps") -> ReplicaDeviceSetter:
    return ReplicaDeviceSetter(cluster=cluster, ps_tasks=ps_tasks, worker_device=worker_device, ps_device=ps_device)


@dataclass(frozen=True)
class RowShard:
    """Row ``r`` of a [rows, dim] table lives on rank ``r % world`` at local row ``r // world``.
    Used by the GPU trainer (cross-checked against the C++ plan's layout at construction), by the serving
    loader that re-assembles exported shards, and by the multi-rank verify step."""
    rows: int
    world: int

    def owner(self, row: int) -> int:
        return row % self.world

    def local_row(self, row: int) -> int:
        return row // self.world

    def local_rows(self, rank: int) -> int:
        return (self.rows - rank + self.world - 1) // self.world

    def max_local_rows(self) -> int:
        return (self.rows + self.world - 1) // self.world

    def global_row(self, rank: int, local: int) -> int:
        return local * self.world + rank


@dataclass(frozen=True)
class RangeShard:
    """A flat index space [0, n) cut into ``world`` contiguous ranges aligned to ``align`` elements."""
    n: int
    world: int
    align: int = 256

    def _chunk(self) -> int:
        per = (self.n + self.world - 1) // self.world
        return (per + self.align - 1) // self.align * self.align

    def bounds(self, rank: int) -> Tuple[int, int]:
        c = self._chunk()
        lo = min(self.n, rank * c)
        return lo, min(self.n, lo + c)

    def all_bounds(self) -> List[Tuple[int, int]]:
        return [self.bounds(r) for r in range(self.world)]

    def owner(self, index: int) -> int:
        return min(self.world - 1, index // self._chunk())
