"""Shared driver behind the three entry points ``trainer/task.py``, ``auto_stop_ps/task.py`` and
``trainer/task_supervisor.py``.

It reproduces the observable behaviour of the reference scripts -- role switch on ``TF_CONFIG``
(/root/reference/trainer/task.py:59,109-123), the printed progress lines (:41,79,82,96,101,103,
244,251), summary tags and gating (:72-73,89-96,238-246), Supervisor-managed session with a 60 s
checkpoint timer (:215-226), master-only export (:254-258), swallowed exceptions in distributed mode
(:260-261) and the ps auto-stop queues (auto_stop_ps/task.py:117-150,175-178,270-272) -- on top of
this package's own runtime instead of TensorFlow.
"""
from __future__ import annotations

import datetime
import logging
import os
import sys

import numpy as np

from .checkpoint.saver import Saver
from .cluster import Server, parse_tf_config
from .export import export_saved_model, export_session_bundle
from .models.linear import LinearModel, make_toy_data
from .optim import get_optimizer
from .session import LocalSession, PSClient, PSSession
from .summary import FileWriter, Summary
from .supervisor import Supervisor


def done_queue_name(ps_index: int) -> str:
    return "worker_done_queue_%d" % ps_index


class Variant:
    """Flag-name differences between trainer.task / auto_stop_ps.task and trainer.task_supervisor."""

    def __init__(self, flags, legacy: bool = False, auto_stop: bool = False):
        self.flags, self.legacy, self.auto_stop = flags, legacy, auto_stop

    @property
    def checkpoint_dir(self):
        return self.flags.checkpoint_dir if self.legacy else self.flags.checkpoint_path

    @property
    def tensorboard_dir(self):
        return self.flags.tensorboard_dir if self.legacy else self.flags.output_path

    def export(self, sess, announce: bool = True) -> None:
        f = self.flags
        if self.legacy:
            # session_bundle path of task_supervisor.py:185-200 / 206-219
            print("Exporting trained model to {}".format(f.export_dir))
            export_session_bundle(sess, f.export_dir, f.export_version)
            print("Done exporting!")
            return
        # saved_model() of task.py:264-291: failures are logged, never raised
        logging.info("Export the saved model to {}".format(f.saved_model_path))
        try:
            export_saved_model(sess, f.saved_model_path, f.model_version)
        except Exception as e:  # noqa: BLE001
            logging.error("Fail to export saved model, exception: {}".format(e))


def _flag(flags, name, default):
    return getattr(flags, name) if name in flags else default


def run(flags, legacy: bool = False, auto_stop: bool = False) -> None:
    variant = Variant(flags, legacy=legacy, auto_stop=auto_stop)
    if _flag(flags, "model", "linear") == "wide_deep":
        # GPU job: master / worker tasks train (every rank is also a ps shard in HBM); a task started with
        # the ps role hosts the control-plane store and leaves like any other ps (forever / auto-stop)
        cfg = parse_tf_config()
        if cfg is not None and cfg.task_type == "ps":
            server = Server(cfg.cluster, job_name=cfg.task_type, task_index=cfg.task_index)
            _run_ps_task(variant, cfg, server, int(_flag(flags, "ps_timeout_ms", 30000)))
            return
        from .models.wide_deep_trainer import run_wide_deep
        run_wide_deep(flags, auto_stop=auto_stop)
        return

    train_x, train_y = make_toy_data()
    learning_rate = flags.learning_rate
    start_training_time = datetime.datetime.now()

    print("Use the optimizer: {}".format(flags.optimizer))
    try:
        optimizer = get_optimizer(flags.optimizer, learning_rate)
    except KeyError:
        print("Unknow optimizer: {}, exit now".format(flags.optimizer))
        sys.exit(1)

    model = LinearModel()
    cfg = parse_tf_config()
    if cfg is None:
        _run_standalone(variant, model, optimizer, train_x, train_y, start_training_time)
    else:
        _run_distributed(variant, cfg, model, optimizer, train_x, train_y, start_training_time)


def _summaries(values: dict, tags) -> Summary:
    return Summary((t, values["loss"]) for t in tags)


def _run_standalone(variant, model, optimizer, train_x, train_y, t0) -> None:
    flags = variant.flags
    sess = LocalSession(model, optimizer)
    sess.run_init()
    tags = ["loss"] if variant.legacy else ["loss", "training/hptuning/metric"]
    print("Save tensorboard files into: {}".format(variant.tensorboard_dir))
    writer = FileWriter(variant.tensorboard_dir, graph=_graph_bytes(model))
    print("Run training with epoch number: {}".format(flags.max_epochs))
    for i in range(flags.max_epochs):
        for (x, y) in zip(train_x, train_y):
            sess.train_step(np.array([[x]]), np.array([[y]]))
        if i % flags.checkpoint_period == 0:
            loss_value, step = sess.evaluate(np.array([[train_x[0]]]), np.array([[train_y[0]]]))
            writer.add_summary(_summaries({"loss": loss_value}, tags), step)
            print("Epoch: {}, loss: {}".format(i, loss_value))
    writer.close()
    print("[{}] End of standalone training.".format(datetime.datetime.now() - t0))
    print("Get the model, w: {}, b: {}".format(sess.value("weight"), sess.value("bias")))
    # The reference leaves its standalone export commented out (task.py:106) while the legacy
    # variant does export (task_supervisor.py:107); here standalone exports unless switched off.
    if variant.legacy or _flag(flags, "export_standalone", True):
        variant.export(sess)


def _graph_bytes(model) -> bytes:
    try:
        from tensorboard.compat.proto import graph_pb2
        from .export.saved_model import _toy_graph_def
        g = graph_pb2.GraphDef()
        _toy_graph_def(g, model)
        return g.SerializeToString()
    except Exception:  # noqa: BLE001
        return b""


def _run_ps_task(variant, cfg, server, timeout_ms: int) -> None:
    """The ps role: plain = ``server.join()`` forever (reference trainer/task.py:121-122); auto-stop = wait
    for one token per worker and per master on this ps's queue, then exit (auto_stop_ps/task.py:137-150)."""
    cluster, task_index = cfg.cluster, cfg.task_index
    if not variant.auto_stop:
        server.join()
        return
    worker_number = cluster.num_tasks("worker") if "worker" in cluster.jobs else 0
    master_worker_number = worker_number + 1        # auto_stop_ps/task.py:117-121 assumes exactly one master
    me = PSClient(cluster.task_address("ps", task_index), timeout_ms=timeout_ms)
    me.queue_create(done_queue_name(task_index), master_worker_number)
    for i in range(master_worker_number):
        while me.dequeue(done_queue_name(task_index), timeout_ms=1000) is None:
            pass
        logging.info("{} workers are already done".format(i + 1))
    logging.info("PS exists after all workers done")
    me.close()
    server.stop()


def _run_distributed(variant, cfg, model, optimizer, train_x, train_y, t0) -> None:
    flags = variant.flags
    cluster, task_type, task_index = cfg.cluster, cfg.task_type, cfg.task_index
    server = Server(cluster, job_name=task_type, task_index=task_index)

    ps_number = cluster.num_tasks("ps")
    # auto_stop_ps/task.py:117-121 assumes exactly one master next to the workers
    worker_number = cluster.num_tasks("worker") if "worker" in cluster.jobs else 0
    master_worker_number = worker_number + 1
    timeout_ms = int(_flag(flags, "ps_timeout_ms", 30000))

    if task_type == "ps":
        _run_ps_task(variant, cfg, server, timeout_ms)
        return

    if task_type not in ("worker", "master"):
        return
    is_chief = task_type == "master"
    saver = Saver(sharded=True) if variant.legacy else Saver()
    recovery_wait = float(_flag(flags, "recovery_wait_secs", 30))
    # task.py passes summary_op=None (:219-220); task_supervisor.py hands the Supervisor its summary op (:150), which
    # in TF cannot be evaluated there (it depends on fed placeholders) -- here it is the loss on the first example
    summary_op = None
    if variant.legacy:
        summary_op = lambda s: [("loss", float(s.evaluate(np.array([[train_x[0]]]), np.array([[train_y[0]]]))[0]))]  # noqa: E731
    sv = Supervisor(is_chief=is_chief, logdir=variant.checkpoint_dir, summary_op=summary_op, saver=saver,
                    global_step=model.global_step_name, save_model_secs=float(_flag(flags, "save_model_secs", 60)),
                    recovery_wait_secs=recovery_wait,
                    session_factory=lambda target: PSSession(model, optimizer, cluster, timeout_ms=timeout_ms,
                                                             connect_retry_secs=float(_flag(flags, "connect_retry_secs", 60))))
    try:
        with sv.managed_session(server.target) as sess:
            print("Save tensorboard files into: {}".format(variant.tensorboard_dir))
            writer = FileWriter(variant.tensorboard_dir, graph=_graph_bytes(model),
                                filename_suffix=".%s%d" % (task_type, task_index))
            print("Run training with epoch number: {}".format(flags.max_epochs))
            for i in range(flags.max_epochs):
                for (x, y) in zip(train_x, train_y):
                    sess.train_step(np.array([[x]]), np.array([[y]]))
                if i % flags.checkpoint_period == 0:
                    loss_value, step = sess.evaluate(np.array([[train_x[0]]]), np.array([[train_y[0]]]))
                    print("Epoch: {}, loss: {}".format(i, loss_value))
                    if is_chief:
                        writer.add_summary(_summaries({"loss": loss_value}, ["loss"]), step)
            writer.close()
            print("[{}] End of distributed training.".format(datetime.datetime.now() - t0))

            # The reference's master tells the ps it is done *before* exporting
            # (auto_stop_ps/task.py:270-278), so a ps can vanish under the export; here every task
            # reports after its last read of ps state.
            if is_chief:
                variant.export(sess)
            if variant.auto_stop:
                logging.info("Notify ps that this worker is done")
                for i in range(ps_number):
                    sess.clients[i].enqueue(done_queue_name(i), 1)
    except Exception as e:  # noqa: BLE001  (reference task.py:260-261 swallows everything)
        if variant.legacy:
            print("Catch the exception, refer to https://github.com/tensorflow/tensorflow/issues/5110")
        print(e)
    finally:
        server.stop()
