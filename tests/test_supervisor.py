"""Supervisor semantics (reference trainer/task.py:215-226): chief restore-or-init,
non-chief wait, immediate first checkpoint then periodic, no forced save at exit, exceptions
re-raised by managed_session."""
import os
import threading
import time

import numpy as np
import pytest

from distributed_tensorflow_b200.checkpoint.saver import Saver, latest_checkpoint
from distributed_tensorflow_b200.cluster import ClusterSpec, Server
from distributed_tensorflow_b200.models.linear import LinearModel
from distributed_tensorflow_b200.optim import get_optimizer
from distributed_tensorflow_b200.session import LocalSession, PSSession
from distributed_tensorflow_b200.summary import read_events
from distributed_tensorflow_b200.supervisor import Coordinator, LooperThread, OutOfRangeError, Supervisor


def _local_factory():
    model = LinearModel()
    return lambda target: LocalSession(model, get_optimizer("adagrad", 0.1))


def test_chief_inits_saves_immediately_then_periodically_and_not_at_exit(tmp_path):
    d = str(tmp_path / "ckpt")
    sv = Supervisor(is_chief=True, logdir=d, saver=Saver(), global_step="global_step", save_model_secs=0.3,
                    save_summaries_secs=0.2, session_factory=_local_factory())
    with sv.managed_session("") as sess:
        assert sess.is_ready() and sess.global_step() == 0
        time.sleep(0.15)
        assert os.path.exists(os.path.join(d, "model.ckpt-0.index"))            # first save is immediate
        for _ in range(7):
            sess.train_step(np.array([[0.5]]), np.array([[11.0]]))
        time.sleep(0.45)
        assert os.path.exists(os.path.join(d, "model.ckpt-7.index"))            # timer tick
        for _ in range(3):
            sess.train_step(np.array([[0.5]]), np.array([[11.0]]))
    assert not os.path.exists(os.path.join(d, "model.ckpt-10.index"))           # no final save
    assert os.path.exists(os.path.join(d, "graph.pbtxt"))
    assert sv.restored_from is None
    ev = [f for f in os.listdir(d) if f.startswith("events.out.tfevents")]
    tags = {t for e in read_events(os.path.join(d, ev[0])) for t in e.get("summary", {})}
    assert tags == {"global_step/sec"}


def test_chief_restart_restores_latest_checkpoint_including_slots(tmp_path):
    d = str(tmp_path / "ckpt")
    sv = Supervisor(is_chief=True, logdir=d, saver=Saver(), global_step="global_step", save_model_secs=1000,
                    session_factory=_local_factory())
    with sv.managed_session("") as sess:
        while not os.path.exists(os.path.join(d, "checkpoint")):               # let the immediate save land
            time.sleep(0.01)
        for _ in range(5):
            sess.train_step(np.array([[0.5]]), np.array([[11.0]]))
        sv.saver.save(sess, sv.save_path, global_step=sess.global_step())
        w, acc = float(sess.value("weight")), float(sess.value("weight/Adagrad"))
    sv2 = Supervisor(is_chief=True, logdir=d, saver=Saver(), global_step="global_step", save_model_secs=1000,
                     session_factory=_local_factory())
    with sv2.managed_session("") as sess:
        assert sv2.restored_from == os.path.join(d, "model.ckpt-5")
        assert sess.global_step() == 5
        assert float(sess.value("weight")) == w and float(sess.value("weight/Adagrad")) == acc


def test_non_chief_waits_until_chief_has_initialised():
    srv = Server(ClusterSpec({"ps": ["127.0.0.1:0"]}), "ps", 0)
    cluster = ClusterSpec({"ps": ["127.0.0.1:%d" % srv.port]})
    model = LinearModel()
    factory = lambda target: PSSession(model, get_optimizer("sgd", 0.01), cluster, connect_retry_secs=1)
    try:
        worker = Supervisor(is_chief=False, logdir=None, saver=None, global_step="global_step",
                            recovery_wait_secs=0.05, session_factory=factory)
        got = {}

        def run_worker():
            with worker.managed_session("") as s:
                got["step"] = s.global_step()
        t = threading.Thread(target=run_worker)
        t.start()
        time.sleep(0.3)
        assert t.is_alive() and "step" not in got                                # still polling
        chief = Supervisor(is_chief=True, logdir=None, saver=None, global_step="global_step",
                           session_factory=factory)
        with chief.managed_session("") as s:
            t.join(5)
            assert got == {"step": 0}
        with pytest.raises(RuntimeError):
            worker.start_standard_services(None)
    finally:
        srv.stop()


def test_non_chief_gives_up_after_max_wait():
    srv = Server(ClusterSpec({"ps": ["127.0.0.1:0"]}), "ps", 0)
    cluster = ClusterSpec({"ps": ["127.0.0.1:%d" % srv.port]})
    try:
        sv = Supervisor(is_chief=False, recovery_wait_secs=0.05, max_wait_secs=0.2, saver=None,
                        session_factory=lambda t: PSSession(LinearModel(), None, cluster, connect_retry_secs=1))
        with pytest.raises(TimeoutError):
            with sv.managed_session(""):
                pass
    finally:
        srv.stop()


def test_managed_session_reraises_but_swallows_out_of_range(tmp_path):
    sv = Supervisor(is_chief=True, logdir=str(tmp_path), saver=Saver(), global_step="global_step",
                    save_model_secs=1000, session_factory=_local_factory())
    with pytest.raises(ZeroDivisionError):
        with sv.managed_session(""):
            1 / 0
    assert sv.should_stop()
    with sv.managed_session(""):
        raise OutOfRangeError()


def test_coordinator_and_looper():
    coord = Coordinator()
    ticks = []
    t = LooperThread(coord, 0.02, target=lambda: ticks.append(time.time()))
    t.start()
    time.sleep(0.15)
    coord.request_stop()
    coord.join()
    assert len(ticks) >= 3 and not t.is_alive()
    coord.clear_stop()
    bad = LooperThread(coord, 0.01, target=lambda: 1 / 0)
    bad.start()
    assert coord.wait_for_stop(2)
    with pytest.raises(ZeroDivisionError):
        coord.join()


def test_chief_restart_restores_without_resetting_live_ps_state_and_publishes_readiness_last(tmp_path):
    """A chief that comes back while the job is running must not run init_op over live variables
    (TF's Supervisor restores instead), must keep optimizer slots consistent with the checkpoint, and
    workers polling is_ready() must not see 'ready' before the restore has finished."""
    import numpy as np
    from distributed_tensorflow_b200.cluster import ClusterSpec, Server
    from distributed_tensorflow_b200.models.linear import LinearModel
    from distributed_tensorflow_b200.optim import get_optimizer
    from distributed_tensorflow_b200.session import PSSession
    cluster = ClusterSpec({"ps": ["127.0.0.1:0"], "master": ["127.0.0.1:0"]})
    srv = Server(cluster, "ps", 0)
    cluster = ClusterSpec({"ps": ["127.0.0.1:%d" % srv.port], "master": ["127.0.0.1:0"]})
    logdir = str(tmp_path / "ckpt")
    factory = lambda master: PSSession(LinearModel(), get_optimizer("adagrad", 0.1), cluster, connect_retry_secs=5)
    sv = Supervisor(is_chief=True, logdir=logdir, save_model_secs=0, session_factory=factory, recovery_wait_secs=0.05)
    sess = sv.prepare_or_wait_for_session("", start_standard_services=False)
    for _ in range(5):
        sess.train_step(0.5, 11.0)
    sv.saver.save(sess, sv.save_path, global_step=sess.global_step())
    saved_w = float(sess.value("weight"))
    for _ in range(3):                                     # the job keeps training after the checkpoint
        sess.train_step(0.5, 11.0)
    seen = []
    orig_assign = PSSession.assign

    def spying_assign(self, name, value):                  # what a polling worker would see mid-restore
        seen.append(factory("").is_ready())
        return orig_assign(self, name, value)

    PSSession.assign = spying_assign
    try:
        sv2 = Supervisor(is_chief=True, logdir=logdir, save_model_secs=0, session_factory=factory)
        sess2 = sv2.prepare_or_wait_for_session("", start_standard_services=False)
    finally:
        PSSession.assign = orig_assign
    assert sv2.restored_from is not None and seen and not any(seen)      # not ready while tensors were being assigned
    assert sess2.is_ready()
    assert float(sess2.value("weight")) == saved_w and sess2.global_step() == 5
    assert "_dtf_ready" not in sess2.snapshot()
    worker = factory("")
    assert worker.is_ready()
    worker.close(), sess.close(), sess2.close()
    srv.stop()


def test_summary_thread_writes_summary_op_at_the_global_step(tmp_path):
    """SVSummaryThread (reference task_supervisor.py:150 passes summary_op to the Supervisor): the chief evaluates
    it every save_summaries_secs and writes it to the logdir event file at the current global step."""
    import glob
    import time
    import numpy as np
    from distributed_tensorflow_b200.models.linear import LinearModel
    from distributed_tensorflow_b200.optim import get_optimizer
    from distributed_tensorflow_b200.session import LocalSession
    from distributed_tensorflow_b200.summary import read_events
    logdir = str(tmp_path / "log")
    sess0 = LocalSession(LinearModel(), get_optimizer("sgd", 0.01))
    sv = Supervisor(is_chief=True, logdir=logdir, save_model_secs=0, save_summaries_secs=0.05,
                    summary_op=lambda s: [("loss", float(s.evaluate(np.array([[1.0]]), np.array([[12.0]]))[0]))],
                    global_step="global_step", session_factory=lambda t: sess0)
    with sv.managed_session("") as sess:
        for _ in range(5):
            sess.train_step(np.array([[1.0]]), np.array([[12.0]]))
        time.sleep(0.3)
    events = [e for f in glob.glob(logdir + "/events.out.tfevents.*") for e in read_events(f)]
    losses = [(e["step"], v) for e in events for tag, v in e.get("summary", {}).items() if tag == "loss"]
    assert losses and losses[-1][0] == 5 and 0 < losses[-1][1] < 144.0
