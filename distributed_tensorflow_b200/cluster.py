"""Cluster description, ``TF_CONFIG`` parsing and the per-task server.

Mirrors what the reference does at /root/reference/trainer/task.py:110-122: read the ``TF_CONFIG``
JSON (``{"cluster": {"ps": [...], "worker": [...], "master": [...]}, "task": {"type", "index"}}``),
build a ``ClusterSpec``, start a ``Server`` for this task, and either ``join()`` (ps) or train
(worker / master).  The server is the C++ TCP parameter-server transport
(``csrc/native/ps.cc``); on a GPU box the same object is the control plane while the data plane is
NVLink symmetric memory (see ``parallel/symm.py``).
"""
from __future__ import annotations

import json
import os
from typing import Dict, List, Mapping, Optional, Sequence, Union

from . import _native


class ClusterSpec:
    """``job name -> ["host:port", ...]`` (tasks indexed by position)."""

    def __init__(self, cluster: Union["ClusterSpec", Mapping[str, Union[Sequence[str], Mapping[int, str]]]]):
        if isinstance(cluster, ClusterSpec):
            self._jobs: Dict[str, Dict[int, str]] = {j: dict(t) for j, t in cluster._jobs.items()}
            return
        self._jobs = {}
        for job, tasks in cluster.items():
            if isinstance(tasks, Mapping):
                self._jobs[str(job)] = {int(i): str(a) for i, a in tasks.items()}
            else:
                self._jobs[str(job)] = {i: str(a) for i, a in enumerate(tasks)}

    @property
    def jobs(self) -> List[str]:
        return sorted(self._jobs)

    def num_tasks(self, job_name: str) -> int:
        if job_name not in self._jobs:
            raise ValueError("No such job in cluster: %r" % job_name)
        return len(self._jobs[job_name])

    def task_indices(self, job_name: str) -> List[int]:
        if job_name not in self._jobs:
            raise ValueError("No such job in cluster: %r" % job_name)
        return sorted(self._jobs[job_name])

    def task_address(self, job_name: str, task_index: int) -> str:
        try:
            return self._jobs[job_name][task_index]
        except KeyError:
            raise ValueError("No task with index %r in job %r" % (task_index, job_name)) from None

    def job_tasks(self, job_name: str) -> List[str]:
        if job_name not in self._jobs:
            raise ValueError("No such job in cluster: %r" % job_name)
        t = self._jobs[job_name]
        return [t[i] for i in sorted(t)]

    def as_dict(self) -> Dict[str, List[str]]:
        return {j: self.job_tasks(j) for j in self.jobs}

    def __bool__(self) -> bool:
        return bool(self._jobs)

    def __eq__(self, other) -> bool:
        return isinstance(other, ClusterSpec) and self._jobs == other._jobs

    def __repr__(self) -> str:
        return "ClusterSpec(%r)" % (self.as_dict(),)


class TaskConfig:
    """Parsed ``TF_CONFIG``: the cluster plus this process's (type, index)."""

    def __init__(self, cluster: ClusterSpec, task_type: str, task_index: int, raw: dict):
        self.cluster, self.task_type, self.task_index, self.raw = cluster, task_type, task_index, raw

    @property
    def is_chief(self) -> bool:
        return self.task_type == "master"


def parse_tf_config(value: Optional[str] = None) -> Optional[TaskConfig]:
    """Returns None when ``TF_CONFIG`` is unset/empty (standalone mode, reference task.py:59)."""
    if value is None:
        value = os.environ.get("TF_CONFIG", "")
    if value == "":
        return None
    env = json.loads(value)
    task = env.get("task", None)
    if task is None:
        raise ValueError("TF_CONFIG has no 'task' entry")
    cluster = ClusterSpec(env["cluster"])
    return TaskConfig(cluster, str(task["type"]), int(task["index"]), env)


def split_address(addr: str):
    host, _, port = addr.rpartition(":")
    return host, int(port)


class Server:
    """One per task, started on construction, like ``tf.train.Server`` (reference task.py:118-119).

    Binds the ``host:port`` that the cluster lists for ``(job_name, task_index)`` and serves the
    parameter-server protocol.  ``target`` is what sessions connect to; ``join()`` blocks until the
    server is shut down (for a plain ps that is forever, reference task.py:121-122).
    """

    def __init__(self, cluster: Union[ClusterSpec, Mapping], job_name: str, task_index: int = 0, start: bool = True):
        self.cluster = ClusterSpec(cluster)
        self.job_name, self.task_index = job_name, int(task_index)
        self._addr = self.cluster.task_address(job_name, self.task_index)
        self._handle = None
        if start:
            self.start()

    def start(self) -> None:
        if self._handle is not None:
            return
        host, port = split_address(self._addr)
        h = _native.lib().dtfn_server_start(host.encode(), port)
        if not h:
            raise _native.NativeError("cannot start server on %s: %s" % (self._addr, _native.last_error()))
        self._handle = h
        self.port = int(_native.lib().dtfn_server_port(h))

    @property
    def target(self) -> str:
        return "dtf://localhost:%d" % self.port

    def join(self) -> None:
        _native.lib().dtfn_server_join(self._handle)

    def num_variables(self) -> int:
        return int(_native.lib().dtfn_server_num_variables(self._handle))

    def inject_drop(self, n: int) -> None:
        """Test hook: the next ``n`` requests get no answer (connection dropped)."""
        _native.lib().dtfn_server_inject_drop(self._handle, n)

    def stop(self) -> None:
        if self._handle is not None:
            _native.lib().dtfn_server_stop(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self.stop()
        except Exception:
            pass
