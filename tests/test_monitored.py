"""MonitoredTrainingSession + hooks (the hook-driven loop the reference keeps next to its Supervisor
loop, /root/reference/trainer/task.py:178-213): StopAtStepHook, checkpoint hook, restore on restart,
non-chief waits for the chief."""
import os
import threading
import time

import numpy as np
import pytest

from distributed_tensorflow_b200.checkpoint.saver import latest_checkpoint
from distributed_tensorflow_b200.cluster import ClusterSpec, Server
from distributed_tensorflow_b200.models.linear import LinearModel
from distributed_tensorflow_b200.monitored import (CheckpointSaverHook, MonitoredTrainingSession, SessionRunHook,
                                                    StopAtStepHook, SummarySaverHook)
from distributed_tensorflow_b200.optim import get_optimizer
from distributed_tensorflow_b200.session import LocalSession, PSSession
from distributed_tensorflow_b200.summary import read_events


def _factory(opt="adagrad", lr=0.1):
    model = LinearModel()
    return lambda target: LocalSession(model, get_optimizer(opt, lr))


X, Y = np.array([[0.5]]), np.array([[11.0]])


def test_stop_at_step_hook_ends_the_while_loop_and_final_checkpoint_is_written(tmp_path):
    d = str(tmp_path / "ckpt")
    calls = []

    class Probe(SessionRunHook):
        def begin(self): calls.append("begin")
        def after_create_session(self, session, coord): calls.append("created")
        def before_run(self, ctx): calls.append("before")
        def after_run(self, ctx, values): calls.append("after")
        def end(self, session): calls.append("end")

    n = 0
    with MonitoredTrainingSession(master="", is_chief=True, checkpoint_dir=d, hooks=[StopAtStepHook(last_step=25), Probe()],
                                  save_checkpoint_steps=10, save_checkpoint_secs=None, log_step_count_steps=5,
                                  session_factory=_factory()) as sess:
        while not sess.should_stop():
            sess.run(X, Y)
            n += 1
    assert n == 25 and calls[0] == "begin" and calls[1] == "created" and calls[-1] == "end"
    assert calls.count("before") == 25 and calls.count("after") == 25
    for step in (0, 10, 20, 25):                      # immediate, every 10 steps, and once more at the end
        assert os.path.exists(os.path.join(d, "model.ckpt-%d.index" % step)), step
    assert latest_checkpoint(d).endswith("model.ckpt-25")
    ev = [f for f in os.listdir(d) if f.startswith("events.out.tfevents")]
    tags = {t for e in read_events(os.path.join(d, ev[0])) for t in e.get("summary", {})}
    assert "global_step/sec" in tags
    with pytest.raises(RuntimeError, match="should_stop"):
        sess.run(X, Y)


def test_restart_restores_and_num_steps_counts_from_the_restored_step(tmp_path):
    d = str(tmp_path / "ckpt")
    with MonitoredTrainingSession(is_chief=True, checkpoint_dir=d, hooks=[StopAtStepHook(last_step=8)],
                                  save_checkpoint_secs=1000, session_factory=_factory()) as sess:
        while not sess.should_stop():
            sess.run(X, Y)
        w = float(sess.value("weight"))
        acc = float(sess.value("weight/Adagrad"))
    with MonitoredTrainingSession(is_chief=True, checkpoint_dir=d, hooks=[StopAtStepHook(num_steps=4)],
                                  save_checkpoint_secs=1000, session_factory=_factory()) as sess:
        assert sess.global_step() == 8
        assert float(sess.value("weight")) == pytest.approx(w) and float(sess.value("weight/Adagrad")) == pytest.approx(acc)
        while not sess.should_stop():
            sess.run(X, Y)
        assert sess.global_step() == 12
    assert latest_checkpoint(d).endswith("model.ckpt-12")


def test_hook_argument_checks_and_summary_hook(tmp_path):
    with pytest.raises(ValueError):
        StopAtStepHook()
    with pytest.raises(ValueError):
        StopAtStepHook(num_steps=1, last_step=2)
    with pytest.raises(ValueError):
        CheckpointSaverHook(str(tmp_path))
    d = str(tmp_path / "tb")
    hook = SummarySaverHook(lambda s: {"loss": float(s.evaluate(X, Y)[0])}, save_steps=3, output_dir=d)
    with MonitoredTrainingSession(is_chief=True, hooks=[StopAtStepHook(last_step=7), hook], session_factory=_factory()) as sess:
        while not sess.should_stop():
            sess.run(X, Y)
    ev = [f for f in os.listdir(d) if f.startswith("events.out.tfevents")]
    steps = [e["step"] for e in read_events(os.path.join(d, ev[0])) if "loss" in e.get("summary", {})]
    assert steps == [1, 4, 7]


def test_non_chief_waits_for_the_chief_over_the_parameter_server(tmp_path):
    cluster = ClusterSpec({"ps": ["127.0.0.1:0"], "master": ["127.0.0.1:0"], "worker": ["127.0.0.1:0"]})
    ps = Server(cluster, "ps", 0)
    spec = ClusterSpec({"ps": ["127.0.0.1:%d" % ps.port], "master": ["127.0.0.1:0"], "worker": ["127.0.0.1:0"]})
    try:
        mk = lambda target: PSSession(LinearModel(), get_optimizer("sgd", 0.01), spec)
        got = {}

        def worker():
            with MonitoredTrainingSession(is_chief=False, hooks=[StopAtStepHook(num_steps=5)], session_factory=mk,
                                          recovery_wait_secs=0.05, max_wait_secs=20) as s:
                got["ready_at"] = time.time()
                while not s.should_stop():
                    s.run(X, Y)
                got["step"] = s.global_step()

        t = threading.Thread(target=worker)
        t.start()
        time.sleep(0.4)
        assert "ready_at" not in got                          # still waiting: nobody has initialised the variables
        t_chief = time.time()
        with MonitoredTrainingSession(is_chief=True, checkpoint_dir=str(tmp_path / "c"), hooks=[StopAtStepHook(num_steps=3)],
                                      save_checkpoint_secs=1000, session_factory=mk) as s:
            while not s.should_stop():
                s.run(X, Y)
            t.join(20)
            assert not t.is_alive() and got["ready_at"] >= t_chief
            # one shared counter on the ps: each task stops once IT sees start + num_steps, whoever took the steps
            assert s.global_step() >= 3 and got["step"] >= 5
    finally:
        ps.stop()
