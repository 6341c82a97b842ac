"""``tf.train.Supervisor`` / ``Coordinator`` equivalents.

What the reference uses (/root/reference/trainer/task.py:215-226)::

    sv = tf.train.Supervisor(is_chief=(task_type == "master"), logdir=FLAGS.checkpoint_path,
                             init_op=init_op, summary_op=None, saver=saver,
                             global_step=global_step, save_model_secs=60)
    with sv.managed_session(server.target) as sess: ...

Semantics kept:

* chief: open a session, **restore the latest checkpoint in logdir if there is one, else run
  init_op**; then start the standard services -- a checkpoint thread that saves immediately and
  then every ``save_model_secs`` (no forced save at exit), a step-counter thread that writes a
  ``global_step/sec`` summary to the logdir event file every ``save_summaries_secs`` (120) and,
  when ``summary_op`` is given, a summary thread on the same period.
* non-chief: poll until the model is ready (all variables initialised), every
  ``recovery_wait_secs`` (30), for at most ``max_wait_secs`` (7200).
* ``managed_session`` stops the services and closes the session on exit and re-raises anything
  except :class:`OutOfRangeError`.
"""
from __future__ import annotations

import contextlib
import os
import threading
import time
from typing import Callable, List, Optional

from .checkpoint.saver import Saver, latest_checkpoint
from .summary import FileWriter, Summary


class OutOfRangeError(Exception):
    pass


class Coordinator:
    def __init__(self):
        self._stop = threading.Event()
        self._lock = threading.Lock()
        self._exc: Optional[BaseException] = None
        self._threads: List[threading.Thread] = []

    def register_thread(self, t: threading.Thread) -> None:
        with self._lock:
            self._threads.append(t)

    def request_stop(self, ex: BaseException = None) -> None:
        with self._lock:
            if ex is not None and self._exc is None:
                self._exc = ex
        self._stop.set()

    def should_stop(self) -> bool:
        return self._stop.is_set()

    def has_exception(self) -> bool:
        """True when a stop was requested because of an error (not a normal end of training)."""
        with self._lock:
            return self._exc is not None

    def wait_for_stop(self, timeout: float = None) -> bool:
        return self._stop.wait(timeout)

    def clear_stop(self) -> None:
        with self._lock:
            self._exc = None
        self._stop.clear()

    def join(self, threads=None, stop_grace_period_secs: float = 120) -> None:
        with self._lock:
            ts = list(self._threads if threads is None else threads)
        deadline = time.time() + stop_grace_period_secs
        for t in ts:
            t.join(max(0.0, deadline - time.time()))
        alive = [t.name for t in ts if t.is_alive()]
        with self._lock:
            exc, self._exc = self._exc, None
            self._threads = [t for t in self._threads if t.is_alive()]
        if exc is not None:
            raise exc
        if alive:
            raise RuntimeError("Coordinator stopped with threads still running: %s" % " ".join(alive))


class LooperThread(threading.Thread):
    """Calls ``run_loop`` every ``timer_interval_secs`` until the coordinator stops; the first call
    happens immediately."""

    def __init__(self, coord: Coordinator, timer_interval_secs: float, target: Callable = None, name: str = None):
        super().__init__(name=name, daemon=True)
        self._coord, self._interval, self._target = coord, timer_interval_secs, target
        coord.register_thread(self)

    def run_loop(self) -> None:
        if self._target:
            self._target()

    def run(self) -> None:
        try:
            while not self._coord.should_stop():
                t0 = time.time()
                self.run_loop()
                if self._coord.wait_for_stop(max(0.0, self._interval - (time.time() - t0))):
                    break
        except BaseException as e:  # surfaces through Coordinator.join
            self._coord.request_stop(e)


class Supervisor:
    USE_DEFAULT = object()

    def __init__(self, is_chief: bool = True, logdir: str = None, init_op=USE_DEFAULT, summary_op=None,
                 saver: Saver = USE_DEFAULT, global_step=USE_DEFAULT, save_model_secs: float = 600,
                 save_summaries_secs: float = 120, recovery_wait_secs: float = 30, stop_grace_period_secs: float = 120,
                 checkpoint_basename: str = "model.ckpt", max_wait_secs: float = 7200,
                 session_factory: Callable[[str], object] = None, init_fn: Callable = None):
        self.is_chief, self.logdir = is_chief, logdir
        self._init_op, self._summary_op = init_op, summary_op
        self.saver = Saver() if saver is Supervisor.USE_DEFAULT else saver
        self._global_step = global_step
        self.save_model_secs, self.save_summaries_secs = save_model_secs, save_summaries_secs
        self.recovery_wait_secs, self.stop_grace_period_secs = recovery_wait_secs, stop_grace_period_secs
        self.max_wait_secs = max_wait_secs
        self.session_factory, self._init_fn = session_factory, init_fn
        self.coord = Coordinator()
        self.save_path = os.path.join(logdir, checkpoint_basename) if (logdir and is_chief) else None
        self.summary_writer: Optional[FileWriter] = None
        self._started: List[threading.Thread] = []
        self.restored_from: Optional[str] = None
        if self.is_chief and self.logdir:
            os.makedirs(self.logdir, exist_ok=True)

    # ------------------------------------------------------------------ session bring-up
    def prepare_or_wait_for_session(self, master: str = "", start_standard_services: bool = True):
        self.coord.clear_stop()
        if self.session_factory is None:
            raise ValueError("Supervisor needs a session_factory")
        if self.is_chief:
            sess = self.session_factory(master)
            ckpt = latest_checkpoint(self.logdir) if (self.logdir and self.saver is not None) else None
            if ckpt is not None:
                # like TF's SessionManager: restore WITHOUT running init_op -- assign() creates what is
                # missing, live variables of a running job are not reset, and readiness is only
                # published once the restore is complete
                if hasattr(sess, "begin_restore"):
                    sess.begin_restore()
                self.saver.recover_last_checkpoints(self.logdir)
                self.saver.restore(sess, ckpt)
                if hasattr(sess, "end_restore"):
                    sess.end_restore()
                else:
                    sess.run_init() if not sess.is_ready() else None
                if hasattr(sess, "refresh"):
                    sess.refresh()
                self.restored_from = ckpt
            else:
                if self._init_op is Supervisor.USE_DEFAULT or self._init_op is None:
                    sess.run_init()
                else:
                    self._init_op(sess)
                if self._init_fn:
                    self._init_fn(sess)
            if not sess.is_ready():
                raise RuntimeError("Init operations did not make model ready")
            if self.logdir:
                self._write_graph(sess)
            if start_standard_services:
                self.start_standard_services(sess)
            return sess
        # non-chief: wait until the chief has initialised (or restored) the variables
        deadline = time.time() + self.max_wait_secs
        sess = None
        while True:
            try:
                if sess is None:
                    sess = self.session_factory(master)
                if sess.is_ready():
                    if hasattr(sess, "refresh"):
                        sess.refresh()
                    return sess
            except ConnectionError:
                sess = None
            if time.time() >= deadline:
                raise TimeoutError("Session was not ready after waiting %d secs." % self.max_wait_secs)
            time.sleep(self.recovery_wait_secs)

    def _write_graph(self, sess) -> None:
        """``graph.pbtxt`` in ``logdir``: the text form of the model's GraphDef (the same graph the event
        file and ``saved_model.pb`` carry), as TF's Supervisor writes it (reference task.py:215-223)."""
        text = ""
        try:
            from tensorboard.compat.proto import graph_pb2
            from .export.graph_def import build_graph_def, graph_def_text
            g = graph_pb2.GraphDef()
            tensors = None
            try:
                tensors = sess.snapshot()
            except Exception:  # noqa: BLE001  (a session without snapshot(): model variables only)
                tensors = None
            build_graph_def(g, sess.model, tensors)
            g.versions.producer = 134
            text = graph_def_text(g)
        except Exception:  # noqa: BLE001  (no protobuf runtime: fall back to a bare variable list)
            try:
                names = [v.name for v in sess.model.variables]
            except Exception:  # noqa: BLE001
                names = []
            text = "".join('node {\n  name: "%s"\n  op: "VariableV2"\n}\n' % n for n in names)
        with open(os.path.join(self.logdir, "graph.pbtxt"), "w") as f:
            f.write(text)

    # ------------------------------------------------------------------ standard services
    def start_standard_services(self, sess) -> List[threading.Thread]:
        if not self.is_chief:
            raise RuntimeError("Only the chief supervisor can start standard services.")
        threads = []
        if self.logdir and self.save_summaries_secs and self._global_step is not None:
            if self.summary_writer is None:
                self.summary_writer = FileWriter(self.logdir)
            threads.append(_StepCounterThread(self, sess))
        if self.logdir and self.save_summaries_secs and self._summary_op is not None:
            if self.summary_writer is None:
                self.summary_writer = FileWriter(self.logdir)
            threads.append(_SummaryThread(self, sess))
        if self.saver is not None and self.save_model_secs and self.save_path:
            threads.append(_CheckpointThread(self, sess))
        for t in threads:
            t.start()
        self._started.extend(threads)
        return threads

    def stop(self, threads=None, close_summary_writer: bool = True) -> None:
        self.coord.request_stop()
        try:
            self.coord.join(threads, stop_grace_period_secs=self.stop_grace_period_secs)
        finally:
            if close_summary_writer and self.summary_writer is not None:
                self.summary_writer.close()
                self.summary_writer = None
            self._started = []

    def request_stop(self, ex: BaseException = None) -> None:
        self.coord.request_stop(ex)

    def should_stop(self) -> bool:
        return self.coord.should_stop()

    @contextlib.contextmanager
    def managed_session(self, master: str = "", start_standard_services: bool = True,
                        close_summary_writer: bool = True):
        sess = None
        try:
            sess = self.prepare_or_wait_for_session(master, start_standard_services)
            yield sess
        except BaseException as e:
            self.request_stop(e)
        finally:
            try:
                # raises the pending exception (the body's or a service thread's) unless OutOfRange
                try:
                    self.stop(close_summary_writer=close_summary_writer)
                except OutOfRangeError:
                    pass
            finally:
                if sess is not None:
                    try:
                        sess.close()
                    except Exception:
                        pass


class _CheckpointThread(LooperThread):
    """SVTimerCheckpointThread: first save right away, then every ``save_model_secs``."""

    def __init__(self, sv: Supervisor, sess):
        super().__init__(sv.coord, sv.save_model_secs, name="SVTimerCheckpointThread")
        self._sv, self._sess = sv, sess

    def run_loop(self) -> None:
        step = self._sess.global_step() if self._sv._global_step is not None else None
        self._sv.saver.save(self._sess, self._sv.save_path, global_step=step)


class _SummaryThread(LooperThread):
    """SVSummaryThread: every ``save_summaries_secs`` evaluate ``summary_op`` and write it at the current global
    step.  ``summary_op`` is a callable ``(sess) -> Summary`` (or an iterable of ``(tag, value)``); the reference's
    task_supervisor.py passes one (:150), task.py passes None (:219-220), which leaves this thread out."""

    def __init__(self, sv: Supervisor, sess):
        super().__init__(sv.coord, sv.save_summaries_secs, name="SVSummaryThread")
        self._sv, self._sess = sv, sess

    def run_loop(self) -> None:
        value = self._sv._summary_op(self._sess)
        if value is None or self._sv.summary_writer is None:
            return
        if not isinstance(value, Summary):
            value = Summary(value)
        step = self._sess.global_step() if self._sv._global_step is not None else 0
        self._sv.summary_writer.add_summary(value, step)
        self._sv.summary_writer.flush()


class _StepCounterThread(LooperThread):
    """SVStepCounterThread: ``global_step/sec`` summary every ``save_summaries_secs``."""

    def __init__(self, sv: Supervisor, sess):
        super().__init__(sv.coord, sv.save_summaries_secs, name="SVStepCounterThread")
        self._sv, self._sess = sv, sess
        self._last_time, self._last_step = 0.0, None

    def run_loop(self) -> None:
        step = self._sess.global_step()
        now = time.time()
        if self._last_step is not None and now > self._last_time and self._sv.summary_writer is not None:
            rate = (step - self._last_step) / (now - self._last_time)
            self._sv.summary_writer.add_summary(Summary([("global_step/sec", rate)]), step)
            self._sv.summary_writer.flush()
        self._last_time, self._last_step = now, step
