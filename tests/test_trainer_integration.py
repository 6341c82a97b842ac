"""End-to-end runs of the entry points as subprocesses -- the reference's only "test" is its README
run-book (README.md:1-41): standalone, then ps + worker + master on loopback ports, then serving."""
import json
import os
import signal
import socket
import subprocess
import sys
import time

import numpy as np
import pytest

from distributed_tensorflow_b200.checkpoint.bundle import BundleReader
from distributed_tensorflow_b200.checkpoint.saver import latest_checkpoint
from distributed_tensorflow_b200.summary import read_events

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.timeout(180)


def _free_ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def _env(tf_config=None):
    env = dict(os.environ)
    env["PYTHONPATH"] = REPO + os.pathsep + env.get("PYTHONPATH", "")
    env["CUDA_VISIBLE_DEVICES"] = ""
    env.pop("TF_CONFIG", None)
    if tf_config is not None:
        env["TF_CONFIG"] = json.dumps(tf_config)
    return env


def _spawn(module, args, cwd, tf_config=None):
    return subprocess.Popen([sys.executable, "-m", module, *args], cwd=cwd, env=_env(tf_config),
                            stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)


def test_standalone_prints_trains_writes_events_and_exports(tmp_path):
    p = _spawn("trainer.task", ["--max_epochs=10"], str(tmp_path))
    out, _ = p.communicate(timeout=120)
    assert p.returncode == 0, out
    lines = out.strip().splitlines()
    assert lines[0] == "Use the optimizer: sgd"
    assert lines[1] == "Save tensorboard files into: ./tensorboard/"
    assert lines[2] == "Run training with epoch number: 10"
    assert [l.split(",")[0] for l in lines[3:13]] == ["Epoch: %d" % i for i in range(10)]
    assert lines[13].endswith("End of standalone training.") and lines[13].startswith("[0:00:")
    w, b = [float(x.split(":")[1]) for x in lines[14].replace("Get the model, ", "").split(",")]
    assert abs(w - 2) < 0.3 and abs(b - 10) < 0.3                              # ground truth w=2, b=10
    ev = os.listdir(tmp_path / "tensorboard")
    assert len(ev) == 1
    sums = [e for e in read_events(str(tmp_path / "tensorboard" / ev[0])) if "summary" in e]
    assert [e["step"] for e in sums] == [100 * (i + 1) for i in range(10)]     # batch 1, 100 steps / epoch
    assert set(sums[0]["summary"]) == {"loss", "training/hptuning/metric"}
    assert not (tmp_path / "checkpoint").exists()                              # standalone never checkpoints
    assert (tmp_path / "saved_model" / "1" / "saved_model.pb").exists()


def test_unknown_optimizer_exits_1(tmp_path):
    p = _spawn("trainer.task", ["--optimizer=momentum"], str(tmp_path))
    out, _ = p.communicate(timeout=60)
    assert p.returncode == 1
    assert out.strip().splitlines() == ["Use the optimizer: momentum", "Unknow optimizer: momentum, exit now"]


def _cluster(n_ps=1):
    ports = _free_ports(n_ps + 2)
    return {"ps": ["127.0.0.1:%d" % p for p in ports[:n_ps]], "worker": ["127.0.0.1:%d" % ports[n_ps]],
            "master": ["127.0.0.1:%d" % ports[n_ps + 1]]}


def test_readme_cluster_with_auto_stop_ps(tmp_path):
    cluster = _cluster(n_ps=2)
    args = ["--max_epochs=4", "--recovery_wait_secs=0.1", "--optimizer=adagrad", "--learning_rate=1.0"]
    cfg = lambda t, i=0: {"cluster": cluster, "task": {"type": t, "index": i}}
    ps = [_spawn("auto_stop_ps.task", args, str(tmp_path), cfg("ps", i)) for i in range(2)]
    worker = _spawn("auto_stop_ps.task", args, str(tmp_path), cfg("worker"))
    time.sleep(0.3)
    master = _spawn("auto_stop_ps.task", args, str(tmp_path), cfg("master"))
    mout, _ = master.communicate(timeout=120)
    wout, _ = worker.communicate(timeout=120)
    assert master.returncode == 0 and worker.returncode == 0, (mout, wout)
    for p in ps:                                                              # ps exits by itself, status 0
        pout, _ = p.communicate(timeout=30)
        assert p.returncode == 0, pout
    assert "End of distributed training." in mout and "End of distributed training." in wout
    assert mout.count("Epoch: ") == 4 and wout.count("Epoch: ") == 4
    ck = str(tmp_path / "checkpoint")
    assert latest_checkpoint(ck) is not None                                   # Supervisor's immediate save
    with BundleReader(str(tmp_path / "saved_model" / "1" / "variables" / "variables")) as r:
        # no data sharding: worker and master each run all 100 * max_epochs steps (reference A14); the
        # master exports when *it* is done, so the exported step lies between its own and the total
        assert 100 * 4 <= int(r.get_tensor("global_step")) <= 2 * 100 * 4
        assert abs(float(r.get_tensor("weight")) - 2) < 1.0 and abs(float(r.get_tensor("bias")) - 10) < 1.0
        assert {"weight/Adagrad", "bias/Adagrad"} <= set(r.keys())
    # every task opens an event writer, only the master records summaries (reference :228,:245-246)
    files = sorted(os.listdir(tmp_path / "tensorboard"))
    assert len(files) == 2
    by_role = {("master" if "master" in f else "worker"): [e for e in read_events(str(tmp_path / "tensorboard" / f))
                                                             if "summary" in e] for f in files}
    assert len(by_role["master"]) == 4 and by_role["worker"] == []


def test_plain_task_ps_stays_alive_and_master_resumes_from_checkpoint(tmp_path):
    cluster = _cluster()
    cfg = lambda t: {"cluster": cluster, "task": {"type": t, "index": 0}}
    args = ["--max_epochs=2", "--recovery_wait_secs=0.1", "--save_model_secs=0.2", "--export_standalone=false"]
    ps = _spawn("trainer.task", args, str(tmp_path), cfg("ps"))
    try:
        m1 = _spawn("trainer.task", args, str(tmp_path), cfg("master"))
        out1, _ = m1.communicate(timeout=120)
        assert m1.returncode == 0, out1
        time.sleep(0.2)
        assert ps.poll() is None                                               # plain ps: server.join() forever
        ck = latest_checkpoint(str(tmp_path / "checkpoint"))
        with BundleReader(ck) as r:
            step1 = int(r.get_tensor("global_step"))
        # wipe the ps state by restarting it: the restarted chief must restore, not re-initialise
        ps.send_signal(signal.SIGTERM)
        ps.wait(timeout=10)
        ps = _spawn("trainer.task", args, str(tmp_path), cfg("ps"))
        m2 = _spawn("trainer.task", args + ["--model_version=2"], str(tmp_path), cfg("master"))
        out2, _ = m2.communicate(timeout=120)
        assert m2.returncode == 0, out2
        with BundleReader(str(tmp_path / "saved_model" / "2" / "variables" / "variables")) as r:
            assert int(r.get_tensor("global_step")) == step1 + 200             # continued, not restarted
    finally:
        ps.kill()
        ps.wait()


def test_task_supervisor_variant_exports_session_bundle(tmp_path):
    export_dir = str(tmp_path / "linear_model")
    p = _spawn("trainer.task_supervisor", ["--max_epochs=3", "--export_dir=" + export_dir, "--export_version=4"],
               str(tmp_path))
    out, _ = p.communicate(timeout=120)
    assert p.returncode == 0, out
    assert "Exporting trained model to " + export_dir in out and "Done exporting!" in out
    assert os.path.exists(os.path.join(export_dir, "00000004", "export.meta"))
    ev = os.listdir(tmp_path / "tensorboard")
    tags = {t for e in read_events(str(tmp_path / "tensorboard" / ev[0])) for t in e.get("summary", {})}
    assert tags == {"loss"}                                                    # no hptuning tag in this variant
