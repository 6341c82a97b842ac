"""Flags, TF_CONFIG parsing, ClusterSpec and placement (reference trainer/task.py:17-31,110-127)."""
import json
import os

import pytest

from distributed_tensorflow_b200 import flags as F
from distributed_tensorflow_b200.cluster import ClusterSpec, parse_tf_config
from distributed_tensorflow_b200.placement import RangeShard, ReplicaDeviceSetter, RowShard, replica_device_setter

REFERENCE_EXAMPLE = ('{"cluster": {"ps": ["127.0.0.1:3001"], "worker": ["127.0.0.1:3002", "127.0.0.1:3003"], '
                     '"master": ["127.0.0.1:3004"]}, "task": {"index": 0, "type": "master"}}')


def _fv():
    fv = F.FlagValues()
    F.DEFINE_integer("max_epochs", 10, "", flag_values=fv)
    F.DEFINE_string("optimizer", "sgd", "", flag_values=fv)
    F.DEFINE_float("learning_rate", 0.01, "", flag_values=fv)
    F.DEFINE_boolean("sync_replicas", False, "", flag_values=fv)
    return fv


def test_defaults_and_both_syntaxes():
    fv = _fv()
    rest = fv(["--max_epochs=3", "--optimizer", "adam", "positional"])
    assert (fv.max_epochs, fv.optimizer, fv.learning_rate) == (3, "adam", 0.01)
    assert rest == ["positional"]


def test_unknown_flags_are_ignored():
    fv = _fv()
    fv(["--job-dir=gs://x", "--whatever", "--learning_rate=0.5"])
    assert fv.learning_rate == 0.5 and fv.max_epochs == 10


def test_boolean_forms_and_errors():
    fv = _fv()
    fv(["--sync_replicas"])
    assert fv.sync_replicas is True
    fv(["--nosync_replicas"])
    assert fv.sync_replicas is False
    fv(["--sync_replicas=true"])
    assert fv.sync_replicas is True
    with pytest.raises(ValueError):
        fv(["--max_epochs=abc"])
    with pytest.raises(AttributeError):
        _ = fv.not_a_flag


def test_lazy_parse_reads_sys_argv(monkeypatch):
    fv = _fv()
    monkeypatch.setattr("sys.argv", ["prog", "--max_epochs=7"])
    assert fv.max_epochs == 7


def test_tf_config_reference_example():
    cfg = parse_tf_config(REFERENCE_EXAMPLE)
    assert cfg.task_type == "master" and cfg.task_index == 0 and cfg.is_chief
    assert cfg.cluster.num_tasks("worker") == 2
    assert cfg.cluster.task_address("worker", 1) == "127.0.0.1:3003"
    assert cfg.cluster.as_dict() == json.loads(REFERENCE_EXAMPLE)["cluster"]
    assert cfg.cluster.jobs == ["master", "ps", "worker"]


def test_tf_config_empty_means_standalone(monkeypatch):
    monkeypatch.delenv("TF_CONFIG", raising=False)
    assert parse_tf_config() is None
    monkeypatch.setenv("TF_CONFIG", "")
    assert parse_tf_config() is None


def test_cluster_spec_errors_and_sparse():
    c = ClusterSpec({"ps": {0: "a:1", 2: "b:2"}})
    assert c.task_indices("ps") == [0, 2] and c.job_tasks("ps") == ["a:1", "b:2"]
    with pytest.raises(ValueError):
        c.num_tasks("worker")
    with pytest.raises(ValueError):
        c.task_address("ps", 1)


def test_round_robin_placement_matches_tf():
    # reference graph creation order with two ps tasks: weight->ps0, bias->ps1, global_step->ps0
    s = replica_device_setter(cluster=ClusterSpec({"ps": ["h:1", "h:2"], "worker": ["h:3"]}))
    assert s.place("weight") == "/job:ps/task:0"
    assert s.place("bias") == "/job:ps/task:1"
    assert s.place("global_step") == "/job:ps/task:0"
    assert s.place("bias") == "/job:ps/task:1"           # stable
    assert s.task_of("bias/Adagrad") == 1                # slots colocate with the primary
    assert ReplicaDeviceSetter(ps_tasks=0, worker_device="/job:worker/task:3").place("w") == "/job:worker/task:3"


def test_row_and_range_shards():
    rs = RowShard(rows=10, world=4)
    assert [rs.owner(r) for r in range(10)] == [0, 1, 2, 3, 0, 1, 2, 3, 0, 1]
    assert [rs.local_rows(k) for k in range(4)] == [3, 3, 2, 2]
    assert all(rs.global_row(rs.owner(r), rs.local_row(r)) == r for r in range(10))
    g = RangeShard(n=1000, world=3, align=256)
    b = g.all_bounds()
    assert b[0][0] == 0 and b[-1][1] == 1000 and all(b[i][1] == b[i + 1][0] for i in range(2))
    assert all(lo % 256 == 0 or lo == 1000 for lo, _ in b)
    assert [g.owner(i) for i in (0, 511, 512, 999)] == [0, 0, 1, 1]


def test_entry_points_keep_the_reference_flag_names_and_defaults():
    """SURVEY.md A2 / A2' / A2'': same nine (eight) flags, same defaults, max_epochs 10 / 20 / 100."""
    import json
    import subprocess
    import sys
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    common = {"checkpoint_period": 1, "learning_rate": 0.01, "optimizer": "sgd"}
    expect = {
        "trainer.task": dict(common, max_epochs=10, checkpoint_path="./checkpoint/", output_path="./tensorboard/",
                             model_path="./model/", saved_model_path="./saved_model/", model_version=1),
        "auto_stop_ps.task": dict(common, max_epochs=20, checkpoint_path="./checkpoint/", output_path="./tensorboard/",
                                  model_path="./model/", saved_model_path="./saved_model/", model_version=1),
        "trainer.task_supervisor": dict(common, max_epochs=100, checkpoint_dir="./checkpoint/", tensorboard_dir="./tensorboard/",
                                        export_version=1, export_dir="/tmp/linear_model/"),
    }
    for mod, want in expect.items():
        code = "import json, sys; sys.argv=['x']; import %s as m; print(json.dumps(m.FLAGS.flag_values_dict()))" % mod
        p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120, cwd=repo,
                           env=dict(os.environ, PYTHONPATH=repo))
        assert p.returncode == 0, p.stderr[-1500:]
        got = json.loads(p.stdout.strip().splitlines()[-1])
        for k, v in want.items():
            assert got[k] == v, (mod, k, got.get(k), v)
