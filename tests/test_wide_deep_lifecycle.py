"""The GPU job's lifecycle through the reference's interface (trainer/task.py:110-127, 215-261;
auto_stop_ps/task.py:127-150, 270-272), on the CPU: three real processes started with nothing but
TF_CONFIG -- ps, master, worker -- where the sm_100a trainer is replaced by a CPU stand-in
(tests/helpers/wd_lifecycle_task.py).  What runs for real: role dispatch, one Server per task, the
control-plane collectives over the C++ transport, Supervisor restore-or-init with non-chief waiting,
timer checkpoints served by every rank without stopping training, the merged checkpoint-V2, master-only
export, ps auto-stop, and the non-zero exit of a survivor whose peer died."""
import json
import os
import socket
import subprocess
import sys
import time

import numpy as np
import pytest

from distributed_tensorflow_b200.checkpoint.bundle import BundleReader
from distributed_tensorflow_b200.checkpoint.saver import latest_checkpoint

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def _spawn(cluster, task_type, index, tmp, extra_env=None, args=()):
    env = dict(os.environ)
    env.pop("RANK", None), env.pop("WORLD_SIZE", None)
    env["TF_CONFIG"] = json.dumps({"cluster": cluster, "task": {"type": task_type, "index": index}})
    env["PYTHONPATH"] = REPO
    env.update(extra_env or {})
    cmd = [sys.executable, os.path.join(REPO, "tests", "helpers", "wd_lifecycle_task.py"), "--model=wide_deep", "--device=cpu",
           "--embedding_rows=101", "--batch_size=8", "--train_steps=40", "--max_epochs=4", "--save_model_secs=0.05",
           "--checkpoint_path=%s" % os.path.join(tmp, "ckpt"), "--output_path=%s" % os.path.join(tmp, "tb"),
           "--saved_model_path=%s" % os.path.join(tmp, "saved_model"), "--ps_timeout_ms=8000", *args]
    return subprocess.Popen(cmd, cwd=REPO, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)


def _cluster():
    p = _free_ports(3)
    return {"ps": ["127.0.0.1:%d" % p[0]], "master": ["127.0.0.1:%d" % p[1]], "worker": ["127.0.0.1:%d" % p[2]]}


def test_three_processes_from_tf_config_train_checkpoint_export_and_ps_auto_stops(tmp_path):
    cluster, tmp = _cluster(), str(tmp_path)
    procs = {"ps": _spawn(cluster, "ps", 0, tmp), "worker": _spawn(cluster, "worker", 0, tmp)}
    time.sleep(0.5)                                              # the worker must wait for the (late) chief
    procs["master"] = _spawn(cluster, "master", 0, tmp)
    out = {k: p.communicate(timeout=120)[0] for k, p in procs.items()}
    assert all(p.returncode == 0 for p in procs.values()), out
    for role in ("master", "worker"):
        assert "Use the optimizer: adagrad" in out[role] and "End of distributed training." in out[role], out[role]
        assert "Epoch: 0, loss: 0.5" in out[role]
    ck = latest_checkpoint(os.path.join(tmp, "ckpt"))
    assert ck is not None and os.path.exists(os.path.join(tmp, "ckpt", "graph.pbtxt"))
    with BundleReader(ck) as r:                                  # one merged V2 checkpoint, one data shard per rank
        assert r.num_shards == 2
        keys = r.keys()
        assert {"embedding/part_0", "embedding/part_1", "embedding/part_1/Adagrad", "global_step"} <= set(keys)
        step = int(r.get_tensor("global_step"))
        e0, e1 = r.get_tensor("embedding/part_0"), r.get_tensor("embedding/part_1")
        # the same bytes as one TF-style partitioned variable (slices along dimension 0)
        assert r.num_slices("embedding") == 2 and r.get_tensor("embedding").shape == (101, 128)
        assert np.array_equal(r.get_tensor("embedding")[:51], e0) and np.array_equal(r.get_tensor("embedding/Adagrad")[51:],
                                                                                     r.get_tensor("embedding/part_1/Adagrad"))
    assert 0 <= step <= 80 and e0.shape == (51, 128) and e1.shape == (50, 128)
    from tensorboard.compat.proto import meta_graph_pb2
    mg = meta_graph_pb2.MetaGraphDef()
    mg.ParseFromString(open(ck + ".meta", "rb").read())                      # Saver's .meta next to the checkpoint
    vars_ = {n.name: [d.size for d in n.attr["shape"].shape.dim] for n in mg.graph_def.node if n.op == "VariableV2"}
    assert vars_["embedding"] == [101, 128] and vars_["embedding/part_1"] == [50, 128] and mg.saver_def.restore_op_name == "save/restore_all"
    assert e0[0, 0] >= 1.0 and e1[0, 0] >= 2.0                   # rank-specific init, then +0.5 per step
    first = sorted(f for f in os.listdir(os.path.join(tmp, "ckpt")) if f.endswith(".index"))
    assert any(f == "model.ckpt-0.index" for f in first) or len(first) == 5      # immediate first save (kept unless rotated out)
    assert os.path.exists(os.path.join(tmp, "saved_model", "1", "saved_model.pb"))
    assert any("master" not in f and f.endswith(".rank0") for f in os.listdir(os.path.join(tmp, "tb")))

    # restart: the chief restores the newest checkpoint, the worker loads its own shard of it
    cluster2 = _cluster()
    procs = {"ps": _spawn(cluster2, "ps", 0, tmp), "master": _spawn(cluster2, "master", 0, tmp, args=("--model_version=2",)),
             "worker": _spawn(cluster2, "worker", 0, tmp, args=("--model_version=2",))}
    out = {k: p.communicate(timeout=120)[0] for k, p in procs.items()}
    assert all(p.returncode == 0 for p in procs.values()), out
    ck2 = latest_checkpoint(os.path.join(tmp, "ckpt"))
    with BundleReader(ck2) as r:
        step2 = int(r.get_tensor("global_step"))
        e1b = r.get_tensor("embedding/part_1")
    assert step2 > step and e1b[0, 0] > e1[0, 0]                 # continued from the restored state, not from init


def test_survivor_of_a_dead_rank_exits_non_zero_within_the_timeout(tmp_path):
    cluster, tmp = _cluster(), str(tmp_path)
    t0 = time.time()
    ps = _spawn(cluster, "ps", 0, tmp, extra_env={"FAKE_AUTO_STOP": "0"})
    master = _spawn(cluster, "master", 0, tmp, extra_env={"FAKE_PEER_DEAD_AT_STEP": "15", "FAKE_AUTO_STOP": "0"})
    worker = _spawn(cluster, "worker", 0, tmp, extra_env={"FAKE_DIE_AT_STEP": "12", "FAKE_AUTO_STOP": "0"})
    out_m = master.communicate(timeout=60)[0]
    worker.communicate(timeout=60)
    ps.kill()
    assert worker.returncode == 9
    assert master.returncode == 1 and "timed out waiting for rank 1" in out_m, out_m
    assert time.time() - t0 < 45
