"""``tf.train.Saver`` equivalent on the C++ bundle writer.

Behaviour kept from the reference's use of Saver (/root/reference/trainer/task.py:143,215-223 and
``Saver(sharded=True)`` at trainer/task_supervisor.py:145):

* ``save(store, "<dir>/model.ckpt", global_step=N)`` writes ``model.ckpt-N.index``,
  ``model.ckpt-N.data-0000i-of-0000S``, ``model.ckpt-N.meta`` and updates the text-proto
  ``checkpoint`` state file; only the newest ``max_to_keep`` (default 5) checkpoints survive.
* ``sharded=True`` writes one data shard per parameter-server task (each task saves its own
  variables, then the indices are merged -- MergeV2Checkpoints).
* ``restore(store, path)`` assigns every saved tensor that the store knows.

A *store* is anything with ``snapshot() -> {name: ndarray}`` and ``assign(name, ndarray)``; stores
that live on parameter servers may also offer ``save_shards(tmp_prefix) -> [prefix, ...]``.
"""
from __future__ import annotations

import os
import re
import threading
import time
import uuid
from typing import Dict, List, Optional

import numpy as np

from . import bundle as _bundle

_STATE_FILE = "checkpoint"


# ------------------------------------------------------------------------------ state file
class CheckpointState:
    def __init__(self, model_checkpoint_path: str = "", all_model_checkpoint_paths: List[str] = None):
        self.model_checkpoint_path = model_checkpoint_path
        self.all_model_checkpoint_paths = list(all_model_checkpoint_paths or [])


def _escape(s: str) -> str:
    return s.replace("\\", "\\\\").replace('"', '\\"')


def _unescape(s: str) -> str:
    return s.replace('\\"', '"').replace("\\\\", "\\")


def update_checkpoint_state(save_dir: str, model_checkpoint_path: str, all_model_checkpoint_paths: List[str] = None,
                            latest_filename: str = None) -> None:
    paths = list(all_model_checkpoint_paths or [])
    if not paths or paths[-1] != model_checkpoint_path:
        paths.append(model_checkpoint_path)

    def rel(p: str) -> str:  # TF stores paths relative to the state file when it can
        if os.path.isabs(p) and not os.path.isabs(save_dir):
            return p
        try:
            r = os.path.relpath(p, save_dir)
            return p if r.startswith("..") else r
        except ValueError:
            return p

    lines = ['model_checkpoint_path: "%s"' % _escape(rel(model_checkpoint_path))]
    lines += ['all_model_checkpoint_paths: "%s"' % _escape(rel(p)) for p in paths]
    fn = os.path.join(save_dir, latest_filename or _STATE_FILE)
    tmp = fn + ".tmp" + uuid.uuid4().hex
    with open(tmp, "w") as f:
        f.write("\n".join(lines) + "\n")
    os.replace(tmp, fn)


def get_checkpoint_state(checkpoint_dir: str, latest_filename: str = None) -> Optional[CheckpointState]:
    fn = os.path.join(checkpoint_dir, latest_filename or _STATE_FILE)
    if not os.path.exists(fn):
        return None
    st = CheckpointState()
    pat = re.compile(r'^\s*(model_checkpoint_path|all_model_checkpoint_paths)\s*:\s*"((?:[^"\\]|\\.)*)"\s*$')
    with open(fn) as f:
        for line in f:
            m = pat.match(line)
            if not m:
                continue
            p = _unescape(m.group(2))
            if not os.path.isabs(p):
                p = os.path.join(checkpoint_dir, p)
            if m.group(1) == "model_checkpoint_path":
                st.model_checkpoint_path = p
            else:
                st.all_model_checkpoint_paths.append(p)
    return st


def checkpoint_exists(prefix: str) -> bool:
    return os.path.exists(prefix + ".index")


def latest_checkpoint(checkpoint_dir: str, latest_filename: str = None) -> Optional[str]:
    st = get_checkpoint_state(checkpoint_dir, latest_filename)
    if st and st.model_checkpoint_path and checkpoint_exists(st.model_checkpoint_path):
        return st.model_checkpoint_path
    return None


# ------------------------------------------------------------------------------ meta graph
def _write_meta(path: str, var_names: List[str], sharded: bool, max_to_keep: int, model=None, tensors=None) -> None:
    """``<prefix>.meta``: the MetaGraphDef ``tf.train.Saver`` writes next to every checkpoint.  With a model
    description (the store's ``model``) it holds the model's graph in TF-1 layout with the saver sub-graph the
    saver_def names and the ``variables`` / ``trainable_variables`` collections (``export/graph_def.py``), i.e.
    what ``tf.train.import_meta_graph`` reads; without one, a bare variable list.  Restore never needs it."""
    try:
        from tensorboard.compat.proto import meta_graph_pb2, saver_pb2
    except Exception:  # tensorboard missing: the .meta file is optional
        return
    mg = meta_graph_pb2.MetaGraphDef()
    if model is not None:
        try:
            from ..export.graph_def import fill_meta_graph
            shapes = {k: (np.broadcast_to(np.zeros((), np.asarray(v).dtype), np.asarray(v).shape) if np.asarray(v).size > 16 else np.asarray(v))
                      for k, v in (tensors or {}).items()}
            fill_meta_graph(mg, model, shapes, sharded=sharded, max_to_keep=max_to_keep)
            with open(path, "wb") as f:
                f.write(mg.SerializeToString())
            return
        except Exception:  # noqa: BLE001  (fall back to the bare form below)
            mg = meta_graph_pb2.MetaGraphDef()
    mg.meta_info_def.tensorflow_version = "1.15.0"
    mg.meta_info_def.tags.append("train")
    sd = mg.saver_def
    sd.filename_tensor_name = "save/Const:0"
    sd.save_tensor_name = "save/control_dependency:0"
    sd.restore_op_name = "save/restore_all"
    sd.max_to_keep = max_to_keep
    sd.sharded = sharded
    sd.keep_checkpoint_every_n_hours = 10000.0
    sd.version = saver_pb2.SaverDef.V2
    for n in var_names:
        node = mg.graph_def.node.add()
        node.name = n
        node.op = "VariableV2"
    with open(path, "wb") as f:
        f.write(mg.SerializeToString())


# ------------------------------------------------------------------------------ Saver
class Saver:
    def __init__(self, var_list: List[str] = None, sharded: bool = False, max_to_keep: int = 5,
                 write_meta_graph: bool = True):
        self.var_list = list(var_list) if var_list is not None else None
        self.sharded, self.max_to_keep, self.write_meta_graph = sharded, max_to_keep, write_meta_graph
        self._last_checkpoints: List[str] = []
        self._lock = threading.Lock()   # the Supervisor's timer thread and the trainer may both save

    @property
    def last_checkpoints(self) -> List[str]:
        return list(self._last_checkpoints)

    def recover_last_checkpoints(self, checkpoint_dir: str) -> None:
        st = get_checkpoint_state(checkpoint_dir)
        if st:
            self._last_checkpoints = [p for p in st.all_model_checkpoint_paths if checkpoint_exists(p)]

    def _filter(self, tensors: Dict[str, np.ndarray]) -> Dict[str, np.ndarray]:
        if self.var_list is None:
            return tensors
        return {k: v for k, v in tensors.items() if k in self.var_list}

    def save(self, store, save_path: str, global_step: Optional[int] = None, write_state: bool = True) -> str:
        with self._lock:
            return self._save(store, save_path, global_step, write_state)

    def _save(self, store, save_path: str, global_step: Optional[int], write_state: bool) -> str:
        prefix = save_path if global_step is None else "%s-%d" % (save_path, int(global_step))
        save_dir = os.path.dirname(prefix) or "."
        os.makedirs(save_dir, exist_ok=True)
        names: List[str]
        if self.sharded and hasattr(store, "save_shards"):
            tmp_dir = prefix + "_temp_" + uuid.uuid4().hex
            os.makedirs(tmp_dir, exist_ok=True)
            shard_prefixes = store.save_shards(os.path.join(tmp_dir, "part"), self.var_list)
            _bundle.merge_bundles(shard_prefixes, prefix)
            try:
                os.rmdir(tmp_dir)
            except OSError:
                pass
            with _bundle.BundleReader(prefix) as r:
                names = r.keys()
                tensors = {k: r.get_tensor(k) for k in names if int(np.prod(r.get_variable_to_shape_map()[k])) <= 16}
        else:
            tensors = self._filter(store.snapshot())
            _bundle.write_bundle(prefix, tensors)
            names = sorted(tensors)
        if self.write_meta_graph:
            _write_meta(prefix + ".meta", names, self.sharded, self.max_to_keep, getattr(store, "model", None), tensors)
        if prefix in self._last_checkpoints:
            self._last_checkpoints.remove(prefix)
        self._last_checkpoints.append(prefix)
        if self.max_to_keep and self.max_to_keep > 0:
            while len(self._last_checkpoints) > self.max_to_keep:
                self._delete(self._last_checkpoints.pop(0))
        if write_state:
            update_checkpoint_state(save_dir, prefix, self._last_checkpoints)
        return prefix

    @staticmethod
    def _delete(prefix: str) -> None:
        d = os.path.dirname(prefix) or "."
        base = os.path.basename(prefix)
        for fn in os.listdir(d):
            if fn == base + ".index" or fn == base + ".meta" or fn.startswith(base + ".data-"):
                try:
                    os.remove(os.path.join(d, fn))
                except OSError:
                    pass

    def restore(self, store, save_path: str) -> List[str]:
        """Assigns every tensor of the checkpoint that the store accepts; returns their names."""
        if not checkpoint_exists(save_path):
            raise FileNotFoundError("no checkpoint at %r" % save_path)
        restored = []
        with _bundle.BundleReader(save_path) as r:
            for name in r.keys():
                if self.var_list is not None and name not in self.var_list:
                    continue
                store.assign(name, r.get_tensor(name))
                restored.append(name)
        return restored


def load_checkpoint(prefix_or_dir: str) -> _bundle.BundleReader:
    """``tf.train.load_checkpoint``: accepts a prefix or a directory holding a ``checkpoint`` file."""
    if os.path.isdir(prefix_or_dir):
        p = latest_checkpoint(prefix_or_dir)
        if p is None:
            raise FileNotFoundError("no checkpoint state in %r" % prefix_or_dir)
        prefix_or_dir = p
    return _bundle.BundleReader(prefix_or_dir)
