"""MonitoredTrainingSession + session-run hooks (the hook-driven alternative to the Supervisor).

The reference builds ``hooks = [tf.train.StopAtStepHook(last_step=1000000)]`` and keeps a
``tf.train.MonitoredTrainingSession(master, is_chief, checkpoint_dir, hooks)`` training loop in a
string literal (/root/reference/trainer/task.py:178-213): ``while not sess.should_stop(): ...
sess.run(train_op, ...)``.  This module provides that API over the framework's sessions
(:class:`~distributed_tensorflow_b200.session.LocalSession`, ``PSSession`` or any object with
``run_init / is_ready / train_step / global_step / close``):

* the chief initialises the variables or restores the newest checkpoint of ``checkpoint_dir``;
  other tasks wait until the chief has done so (same rule as the Supervisor);
* hooks see ``begin`` -> ``after_create_session`` -> (``before_run`` / ``after_run``)* -> ``end``;
* :class:`StopAtStepHook` requests a stop once ``global_step`` reaches ``last_step`` (or ``num_steps``
  after the step at which the session was created);
* :class:`CheckpointSaverHook` (added for the chief when ``checkpoint_dir`` is given) saves every
  ``save_checkpoint_secs`` seconds or ``save_checkpoint_steps`` steps and once more at the end;
* :class:`StepCounterHook` / :class:`SummarySaverHook` write ``global_step/sec`` and user scalars
  to an event file in ``checkpoint_dir``.
"""
from __future__ import annotations

import os
import time
from typing import Callable, Dict, List, Optional, Sequence

from .checkpoint import Saver, latest_checkpoint
from .summary import FileWriter, Summary


class SessionRunHook:
    def begin(self) -> None: ...
    def after_create_session(self, session, coord) -> None: ...
    def before_run(self, run_context) -> None: ...
    def after_run(self, run_context, run_values) -> None: ...
    def end(self, session) -> None: ...


class RunContext:
    def __init__(self, session):
        self.session = session
        self.stop_requested = False

    def request_stop(self) -> None:
        self.stop_requested = True


class StopAtStepHook(SessionRunHook):
    """Requests stop at a global step: exactly one of ``num_steps`` / ``last_step``."""

    def __init__(self, num_steps: Optional[int] = None, last_step: Optional[int] = None):
        if (num_steps is None) == (last_step is None):
            raise ValueError("One of num_steps or last_step must be specified.")
        self._num_steps, self._last_step = num_steps, last_step

    def after_create_session(self, session, coord) -> None:
        if self._last_step is None:
            self._last_step = session.global_step() + self._num_steps

    def after_run(self, run_context, run_values) -> None:
        if run_context.session.global_step() >= self._last_step:
            run_context.request_stop()


class CheckpointSaverHook(SessionRunHook):
    def __init__(self, checkpoint_dir: str, save_secs: Optional[float] = None, save_steps: Optional[int] = None,
                 saver: Optional[Saver] = None, checkpoint_basename: str = "model.ckpt"):
        if save_secs is None and save_steps is None:
            raise ValueError("either save_secs or save_steps should be set")
        self._dir, self._secs, self._steps = checkpoint_dir, save_secs, save_steps
        self.saver = saver or Saver()
        self._path = os.path.join(checkpoint_dir, checkpoint_basename)
        self._last_time, self._last_step = None, None

    def begin(self) -> None:
        os.makedirs(self._dir, exist_ok=True)

    def _save(self, session) -> None:
        step = session.global_step()
        self.saver.save(session, self._path, global_step=step)
        self._last_time, self._last_step = time.time(), step

    def after_create_session(self, session, coord) -> None:
        self._save(session)                      # TF saves once right after the session is created

    def after_run(self, run_context, run_values) -> None:
        step = run_context.session.global_step()
        due = (self._secs is not None and time.time() - self._last_time >= self._secs) or \
              (self._steps is not None and step - self._last_step >= self._steps)
        if due:
            self._save(run_context.session)

    def end(self, session) -> None:
        if session.global_step() != self._last_step:
            self._save(session)


class StepCounterHook(SessionRunHook):
    def __init__(self, every_n_steps: int = 100, output_dir: Optional[str] = None, writer: Optional[FileWriter] = None):
        self._every, self._dir, self._writer = every_n_steps, output_dir, writer
        self._t0, self._s0 = None, None

    def after_create_session(self, session, coord) -> None:
        if self._writer is None and self._dir:
            self._writer = FileWriter(self._dir)
        self._t0, self._s0 = time.time(), session.global_step()

    def after_run(self, run_context, run_values) -> None:
        step = run_context.session.global_step()
        if step - self._s0 >= self._every:
            now = time.time()
            if self._writer is not None and now > self._t0:
                self._writer.add_summary(Summary([("global_step/sec", (step - self._s0) / (now - self._t0))]), step)
            self._t0, self._s0 = now, step

    def end(self, session) -> None:
        if self._writer is not None:
            self._writer.flush()


class SummarySaverHook(SessionRunHook):
    """Every ``save_steps`` steps writes the scalars returned by ``scalars_fn(session)``."""

    def __init__(self, scalars_fn: Callable[[object], Dict[str, float]], save_steps: int = 100,
                 output_dir: Optional[str] = None, writer: Optional[FileWriter] = None):
        self._fn, self._every, self._dir, self._writer = scalars_fn, save_steps, output_dir, writer
        self._next = None

    def after_create_session(self, session, coord) -> None:
        if self._writer is None and self._dir:
            self._writer = FileWriter(self._dir)
        self._next = session.global_step()

    def after_run(self, run_context, run_values) -> None:
        step = run_context.session.global_step()
        if step >= self._next and self._writer is not None:
            self._writer.add_summary(Summary(sorted(self._fn(run_context.session).items())), step)
            self._next = step + self._every

    def end(self, session) -> None:
        if self._writer is not None:
            self._writer.flush()


class MonitoredSession:
    """The object ``MonitoredTrainingSession`` returns: ``run`` = one training step through the hooks."""

    def __init__(self, session, hooks: Sequence[SessionRunHook]):
        self._sess, self._hooks = session, list(hooks)
        self._stop = False
        self._closed = False

    @property
    def raw_session(self):
        return self._sess

    def should_stop(self) -> bool:
        return self._stop or self._closed

    def run(self, x, y):
        """``sess.run(train_op, feed_dict={X: x, Y: y})``: one optimizer step; returns the new global step."""
        if self._closed:
            raise RuntimeError("Run called even after should_stop requested.")
        ctx = RunContext(self._sess)
        for h in self._hooks:
            h.before_run(ctx)
        self._sess.train_step(x, y)
        step = self._sess.global_step()
        for h in self._hooks:
            h.after_run(ctx, step)
        self._stop = self._stop or ctx.stop_requested
        return step

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        try:
            for h in self._hooks:
                h.end(self._sess)
        finally:
            self._sess.close()

    def __enter__(self):
        return self

    def __exit__(self, exc_type, exc, tb):
        self.close()
        return False

    def __getattr__(self, name):            # evaluate / value / global_step / snapshot ... pass through
        return getattr(self._sess, name)


def MonitoredTrainingSession(master: str = "", is_chief: bool = True, checkpoint_dir: Optional[str] = None,
                             hooks: Optional[Sequence[SessionRunHook]] = None,
                             chief_only_hooks: Optional[Sequence[SessionRunHook]] = None,
                             save_checkpoint_secs: Optional[float] = 600, save_checkpoint_steps: Optional[int] = None,
                             log_step_count_steps: Optional[int] = 100, saver: Optional[Saver] = None,
                             session_factory: Callable[[str], object] = None, max_wait_secs: float = 7200,
                             recovery_wait_secs: float = 1.0) -> MonitoredSession:
    """See the module docstring.  ``session_factory(master)`` builds the underlying session."""
    if session_factory is None:
        raise ValueError("MonitoredTrainingSession needs a session_factory")
    all_hooks: List[SessionRunHook] = list(hooks or [])
    if is_chief:
        all_hooks += list(chief_only_hooks or [])
        if checkpoint_dir:
            if log_step_count_steps:
                all_hooks.append(StepCounterHook(log_step_count_steps, output_dir=checkpoint_dir))
            if save_checkpoint_secs or save_checkpoint_steps:
                all_hooks.append(CheckpointSaverHook(checkpoint_dir, save_secs=save_checkpoint_secs,
                                                     save_steps=save_checkpoint_steps, saver=saver))
    for h in all_hooks:
        h.begin()
    if is_chief:
        sess = session_factory(master)
        ckpt = latest_checkpoint(checkpoint_dir) if checkpoint_dir else None
        if ckpt is None:
            sess.run_init()
        else:
            # restore without init: live ps state is not reset and readiness is published last
            if hasattr(sess, "begin_restore"):
                sess.begin_restore()
            sv = saver or next((h.saver for h in all_hooks if isinstance(h, CheckpointSaverHook)), Saver())
            sv.recover_last_checkpoints(checkpoint_dir)
            sv.restore(sess, ckpt)
            if hasattr(sess, "end_restore"):
                sess.end_restore()
            elif not sess.is_ready():
                sess.run_init()
            if hasattr(sess, "refresh"):
                sess.refresh()
        if not sess.is_ready():
            raise RuntimeError("Init operations did not make model ready")
    else:
        deadline = time.time() + max_wait_secs
        sess = None
        while True:
            try:
                if sess is None:
                    sess = session_factory(master)
                if sess.is_ready():
                    if hasattr(sess, "refresh"):
                        sess.refresh()
                    break
            except ConnectionError:
                sess = None
            if time.time() >= deadline:
                raise TimeoutError("Session was not ready after waiting %d secs." % max_wait_secs)
            time.sleep(recovery_wait_secs)
    for h in all_hooks:
        h.after_create_session(sess, None)
    return MonitoredSession(sess, all_hooks)
