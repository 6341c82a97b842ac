"""Sessions: where variables live and how one training step reaches them.

* :class:`LocalSession` -- standalone mode (reference trainer/task.py:77-103): variables in this
  process, optimizer applied in place.
* :class:`PSSession` -- distributed mode (reference trainer/task.py:123-143, 226-246): variables are
  placed round-robin on the ps tasks; each ``train_step`` computes gradients locally from the last
  pulled (possibly stale) values and sends them to the owning ps tasks, which apply them immediately
  without locks or cross-worker barriers (async / Hogwild, reference task.py:138).  Push, global-step
  increment and pull of the fresh values travel in ONE round trip per ps task.

Both expose the store interface the Saver uses (``snapshot`` / ``assign``).
"""
from __future__ import annotations

import ctypes as C
import time
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import _native
from .checkpoint.bundle import DTYPE_TO_TF, TF_TO_DTYPE
from .cluster import ClusterSpec, split_address
from .optim import Optimizer
from .placement import ReplicaDeviceSetter


class LocalSession:
    def __init__(self, model, optimizer: Optional[Optimizer]):
        self.model, self.optimizer = model, optimizer
        self.store: Dict[str, np.ndarray] = {}

    # -- lifecycle ---------------------------------------------------------------------------
    def run_init(self) -> None:
        self.store.update(self.model.initial_values())
        if self.optimizer is not None:
            for n in self.model.trainable_names():
                self.optimizer.ensure_slots(self.store, n)

    def is_ready(self) -> bool:
        return all(v.name in self.store for v in self.model.variables)

    def end_restore(self) -> None:
        """After a restore: variables (and slots) the checkpoint did not hold get their initial value."""
        for k, v in self.model.initial_values().items():
            self.store.setdefault(k, v)
        if self.optimizer is not None:
            for n in self.model.trainable_names():
                self.optimizer.ensure_slots(self.store, n)

    def close(self) -> None:
        pass

    # -- training ----------------------------------------------------------------------------
    def train_step(self, x, y) -> None:
        _, grads = self.model.loss_and_grads(self.store, x, y)
        self.optimizer.apply(self.store, grads)
        gs = self.model.global_step_name
        self.store[gs] = np.array(self.store[gs] + 1, dtype=self.store[gs].dtype)

    def evaluate(self, x, y):
        """(loss, global_step) on a batch with the current values, no update (reference task.py:92)."""
        return self.model.loss(self.store, x, y), int(self.store[self.model.global_step_name])

    def value(self, name: str) -> np.ndarray:
        return self.store[name]

    def global_step(self) -> int:
        return int(self.store[self.model.global_step_name])

    # -- store interface ---------------------------------------------------------------------
    def snapshot(self) -> Dict[str, np.ndarray]:
        return {k: np.array(v) for k, v in self.store.items()}

    def assign(self, name: str, value: np.ndarray) -> None:
        self.store[name] = np.array(value)


class PSClient:
    """Connection to one parameter-server task."""

    def __init__(self, address: str, timeout_ms: int = 30000, connect_retry_secs: float = 0.0):
        self.address = address
        host, port = split_address(address)
        self._L = _native.lib()
        deadline = time.time() + connect_retry_secs
        while True:
            self._h = self._L.dtfn_client_connect(host.encode(), port, timeout_ms)
            if self._h or time.time() >= deadline:
                break
            time.sleep(0.1)
        if not self._h:
            raise ConnectionError("cannot reach ps task at %s: %s" % (address, _native.last_error()))

    def close(self) -> None:
        if self._h:
            self._L.dtfn_client_close(self._h)
            self._h = None

    def _check(self, rc: int, what: str) -> int:
        if rc < 0:
            msg = "%s on %s failed (%d): %s" % (what, self.address, rc, _native.last_error())
            raise (ConnectionError if rc <= -100 else _native.NativeError)(msg)
        return rc

    def ping(self) -> None:
        self._check(self._L.dtfn_client_ping(self._h), "ping")

    def init_var(self, name: str, value: np.ndarray, overwrite: bool = False) -> bool:
        a = np.asarray(value)
        if not a.flags.c_contiguous:
            a = a.copy()
        dims = (C.c_longlong * max(1, a.ndim))(*a.shape)
        rc = self._L.dtfn_client_init_var(self._h, name.encode(), DTYPE_TO_TF[a.dtype], a.ndim, dims,
                                          a.ctypes.data_as(C.c_void_p), a.nbytes, int(overwrite))
        return self._check(rc, "init_var %s" % name) == 0

    def var_info(self, name: str):
        dt, nb = C.c_int(), C.c_longlong()
        dims = (C.c_longlong * 8)()
        nd = self._L.dtfn_client_var_info(self._h, name.encode(), C.byref(dt), dims, C.byref(nb))
        if nd == -10:
            return None
        self._check(nd, "var_info %s" % name)
        return TF_TO_DTYPE[dt.value], tuple(dims[i] for i in range(nd)), nb.value

    def pull(self, name: str) -> np.ndarray:
        info = self.var_info(name)
        if info is None:
            raise KeyError("variable %r is not on %s" % (name, self.address))
        dtype, shape, nbytes = info
        out = np.empty(shape, dtype=dtype)
        n = self._L.dtfn_client_pull(self._h, name.encode(), out.ctypes.data_as(C.c_void_p), out.nbytes)
        self._check(int(n), "pull %s" % name)
        return out

    def list(self) -> List[str]:
        n = int(self._check(int(self._L.dtfn_client_list(self._h, None, 0)), "list"))
        buf = C.create_string_buffer(n + 1)
        self._check(int(self._L.dtfn_client_list(self._h, buf, n + 1)), "list")
        s = buf.value.decode()
        return s.split("\n") if s else []

    def push_pull(self, names: Sequence[str], grads: Sequence[np.ndarray], opt_id: int, lr: float,
                  step_name: Optional[str], step_inc: int = 1, want_values: bool = True):
        k = len(names)
        g = [np.ascontiguousarray(np.asarray(x, np.float32).reshape(-1)) for x in grads]
        outs = [np.empty_like(x) for x in g]
        fp = C.POINTER(C.c_float)
        c_names = (C.c_char_p * max(1, k))(*[n.encode() for n in names])
        c_grads = (fp * max(1, k))(*[x.ctypes.data_as(fp) for x in g])
        c_lens = (C.c_size_t * max(1, k))(*[x.size for x in g])
        c_outs = (fp * max(1, k))(*[x.ctypes.data_as(fp) for x in outs]) if want_values else None
        step = C.c_longlong(-1)
        rc = self._L.dtfn_client_push_pull(self._h, k, c_names, opt_id, float(lr), c_grads, c_lens,
                                           step_name.encode() if step_name else None, step_inc, c_outs, C.byref(step))
        self._check(rc, "push_pull")
        return (outs if want_values else None), int(step.value)

    def push_apply(self, name: str, grad: np.ndarray, opt_id: int, lr: float) -> None:
        g = np.ascontiguousarray(np.asarray(grad, np.float32).reshape(-1))
        rc = self._L.dtfn_client_push_apply(self._h, name.encode(), opt_id, float(lr),
                                            g.ctypes.data_as(C.POINTER(C.c_float)), g.size)
        self._check(rc, "push_apply %s" % name)

    def assign_add(self, name: str, delta: int) -> int:
        out = C.c_longlong()
        self._check(self._L.dtfn_client_assign_add_i64(self._h, name.encode(), int(delta), C.byref(out)), "assign_add")
        return int(out.value)

    def queue_create(self, name: str, capacity: int) -> None:
        self._check(self._L.dtfn_client_queue_create(self._h, name.encode(), capacity), "queue_create")

    def enqueue(self, name: str, value: int) -> None:
        self._check(self._L.dtfn_client_enqueue(self._h, name.encode(), int(value)), "enqueue")

    def dequeue(self, name: str, timeout_ms: int = 0) -> Optional[int]:
        v = C.c_int()
        rc = self._L.dtfn_client_dequeue(self._h, name.encode(), C.byref(v), int(timeout_ms))
        if rc == -2:
            return None
        self._check(rc, "dequeue %s" % name)
        return int(v.value)

    def save(self, prefix: str, exclude: Sequence[str] = ()) -> None:
        self._check(self._L.dtfn_client_save(self._h, prefix.encode(), "\n".join(exclude).encode()), "save")

    def shutdown_server(self) -> None:
        self._check(self._L.dtfn_client_shutdown_server(self._h), "shutdown")


class PSSession:
    def __init__(self, model, optimizer: Optional[Optimizer], cluster: ClusterSpec, timeout_ms: int = 30000,
                 connect_retry_secs: float = 60.0):
        self.model, self.optimizer, self.cluster = model, optimizer, ClusterSpec(cluster)
        self.setter = ReplicaDeviceSetter(cluster=self.cluster)
        for v in model.variables:            # placement follows variable creation order
            self.setter.place(v.name)
        self.clients = [PSClient(a, timeout_ms, connect_retry_secs) for a in self.cluster.job_tasks("ps")]
        self.cache: Dict[str, np.ndarray] = {}
        self._by_task: Dict[int, List[str]] = {}
        for n in model.trainable_names():
            self._by_task.setdefault(self.setter.task_of(n), []).append(n)
        self._step_task = self.setter.task_of(model.global_step_name)
        self._last_step = 0

    def close(self) -> None:
        for c in self.clients:
            c.close()

    # -- lifecycle ---------------------------------------------------------------------------
    READY_MARKER = "_dtf_ready"   # ps-side flag written last by the chief; never part of a checkpoint

    def run_init(self, overwrite: bool = True) -> None:
        """``sess.run(init_op)``: (re)create every model variable on its ps task (chief only), then
        publish readiness.  ``overwrite=False`` only fills in variables that do not exist yet."""
        for v in self.model.variables:
            self.clients[self.setter.task_of(v.name)].init_var(v.name, v.initial, overwrite=overwrite)
        self._set_ready(1)
        self.refresh()

    def _set_ready(self, value: int) -> None:
        for c in self.clients:
            c.init_var(self.READY_MARKER, np.array(value, np.int32), overwrite=True)

    def begin_restore(self) -> None:
        """Chief, before assigning checkpoint tensors: withdraw readiness so that workers that poll
        ``is_ready`` do not start on half-restored state.  Live variables are left alone (TF's
        Supervisor restores without running init_op)."""
        self._set_ready(0)

    def end_restore(self) -> None:
        """Variables the checkpoint did not hold get their initial value; readiness is published last."""
        self.run_init(overwrite=False)

    def is_ready(self) -> bool:
        """``report_uninitialized_variables`` is empty: every model variable exists on its ps and the
        chief has finished initialising or restoring (ready marker set on every ps task)."""
        try:
            for c in self.clients:
                if c.var_info(self.READY_MARKER) is None or int(c.pull(self.READY_MARKER)) != 1:
                    return False
            for v in self.model.variables:
                if self.clients[self.setter.task_of(v.name)].var_info(v.name) is None:
                    return False
            return True
        except ConnectionError:
            return False

    def refresh(self) -> None:
        for v in self.model.variables:
            self.cache[v.name] = self.clients[self.setter.task_of(v.name)].pull(v.name)

    # -- training ----------------------------------------------------------------------------
    def train_step(self, x, y) -> None:
        if not self.cache:
            self.refresh()
        _, grads = self.model.loss_and_grads(self.cache, x, y)
        tasks = sorted(set(self._by_task) | {self._step_task})
        for t in tasks:
            names = self._by_task.get(t, [])
            step_name = self.model.global_step_name if t == self._step_task else None
            vals, step = self.clients[t].push_pull(names, [grads[n] for n in names], self.optimizer.opt_id,
                                                   self.optimizer.learning_rate, step_name)
            for n, v in zip(names, vals):
                self.cache[n] = v.reshape(self.cache[n].shape)
            if step_name:
                self._last_step = step

    def evaluate(self, x, y):
        self.refresh()
        self._last_step = int(self.cache[self.model.global_step_name])
        return self.model.loss(self.cache, x, y), self._last_step

    def value(self, name: str) -> np.ndarray:
        return self.clients[self.setter.task_of(name)].pull(name)

    def global_step(self) -> int:
        return int(self.value(self.model.global_step_name))

    # -- store interface ---------------------------------------------------------------------
    def snapshot(self) -> Dict[str, np.ndarray]:
        out: Dict[str, np.ndarray] = {}
        for c in self.clients:
            for n in c.list():
                if n not in out and n != self.READY_MARKER:
                    out[n] = c.pull(n)
        return out

    def assign(self, name: str, value: np.ndarray) -> None:
        try:
            t = self.setter.task_of(name)
        except KeyError:
            t = 0
        if name in ("beta1_power", "beta2_power"):   # every ps keeps its own copy of the Adam powers
            for c in self.clients:
                c.init_var(name, value, overwrite=True)
            return
        self.clients[t].init_var(name, value, overwrite=True)

    def save_shards(self, tmp_prefix: str, var_list=None) -> List[str]:
        """Each ps task writes its own variables (ps-side SaveV2); returns the shard prefixes."""
        seen: set = set()
        prefixes = []
        for i, c in enumerate(self.clients):
            names = c.list()
            exclude = [n for n in names if n in seen or n == self.READY_MARKER or (var_list is not None and n not in var_list)]
            seen.update(names)
            p = "%s-%05d" % (tmp_prefix, i)
            c.save(p, exclude)
            prefixes.append(p)
        return prefixes
