"""SavedModel-compatible export (``SavedModelBuilder``) and its loader.

Layout produced, as by the reference's ``saved_model()`` (/root/reference/trainer/task.py:264-291)::

    <saved_model_path>/<model_version>/saved_model.pb
    <saved_model_path>/<model_version>/variables/variables.index
    <saved_model_path>/<model_version>/variables/variables.data-00000-of-00001
    <saved_model_path>/<model_version>/assets.extra/dtf_model.json     (framework-native description)

``saved_model.pb`` is ``SavedModel{saved_model_schema_version = 1, meta_graphs = [MetaGraphDef]}``
with tag ``serve``, signature key ``serving_default``, method ``tensorflow/serving/predict``, inputs
``keys``/``features`` and outputs ``keys``/``prediction`` (task.py:164-173), a ``legacy_init_op``
collection (task.py:175-176), a saver_def and a device-free GraphDef of the model.  Like TF's
builder, construction **fails if the export directory already exists** (the reference then only
logs the error, task.py:290-291).
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional

import numpy as np

from ..checkpoint import bundle as _bundle

SERVING = "serve"
DEFAULT_SERVING_SIGNATURE_DEF_KEY = "serving_default"
PREDICT_METHOD_NAME = "tensorflow/serving/predict"
SAVED_MODEL_SCHEMA_VERSION = 1
_NP_TO_DT = {np.dtype(np.float32): 1, np.dtype(np.int32): 3, np.dtype(np.int64): 9}


def _varint(v: int) -> bytes:
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _pb():
    from tensorboard.compat.proto import (attr_value_pb2, graph_pb2, meta_graph_pb2, node_def_pb2, saver_pb2,
                                          tensor_shape_pb2, types_pb2)
    return meta_graph_pb2, graph_pb2, saver_pb2, tensor_shape_pb2, types_pb2


def _tensor_info(ti, name: str, dtype, shape) -> None:
    ti.name = name
    ti.dtype = _NP_TO_DT[np.dtype(dtype)]
    for d in shape:
        ti.tensor_shape.dim.add().size = d


def build_signature_def(model):
    """SignatureDef for ``model.signature_inputs/outputs`` with the predict method name."""
    meta_graph_pb2 = _pb()[0]
    sig = meta_graph_pb2.SignatureDef()
    for alias, (tname, dt, shape) in model.signature_inputs.items():
        _tensor_info(sig.inputs[alias], tname, dt, shape)
    for alias, (tname, dt, shape) in model.signature_outputs.items():
        _tensor_info(sig.outputs[alias], tname, dt, shape)
    sig.method_name = PREDICT_METHOD_NAME
    return sig


def _toy_graph_def(graph_def, model, tensors=None) -> dict:
    """Device-free GraphDef of the model in TF-1 layout (export/graph_def.py); returns the SaverDef names."""
    from .graph_def import build_graph_def
    return build_graph_def(graph_def, model, tensors)


class SavedModelBuilder:
    def __init__(self, export_dir):
        if isinstance(export_dir, bytes):
            export_dir = export_dir.decode()
        self.export_dir = export_dir
        if os.path.exists(export_dir):
            raise AssertionError("Export directory already exists. Please specify a different export directory: %s"
                                 % export_dir)
        os.makedirs(export_dir)
        self._meta_graph = None
        self._model_desc: Optional[dict] = None

    def add_meta_graph_and_variables(self, sess, tags, signature_def_map: Dict[str, object] = None,
                                     clear_devices: bool = True, legacy_init_op: Optional[str] = "legacy_init_op",
                                     model=None) -> None:
        """Snapshot every saveable variable of ``sess`` into ``variables/`` and record the MetaGraphDef."""
        meta_graph_pb2, graph_pb2, saver_pb2, _, _ = _pb()
        model = model if model is not None else sess.model
        var_dir = os.path.join(self.export_dir, "variables")
        os.makedirs(var_dir, exist_ok=True)
        tensors = sess.snapshot()  # all variables incl. global_step and optimizer slots, like TF's builder
        _bundle.write_bundle(os.path.join(var_dir, "variables"), tensors)

        mg = meta_graph_pb2.MetaGraphDef()
        mg.meta_info_def.tags.extend(tags)
        mg.meta_info_def.tensorflow_version = "1.15.0"
        mg.meta_info_def.tensorflow_git_version = "distributed_tensorflow_b200"
        saver_names = _toy_graph_def(mg.graph_def, model, tensors)
        mg.graph_def.versions.producer = 134
        sd = mg.saver_def
        sd.filename_tensor_name = saver_names["filename_tensor_name"]
        sd.save_tensor_name = saver_names["save_tensor_name"]
        sd.restore_op_name = saver_names["restore_op_name"]
        sd.max_to_keep = 5
        sd.sharded = True
        sd.keep_checkpoint_every_n_hours = 10000.0
        sd.version = saver_pb2.SaverDef.V2
        for key, sig in (signature_def_map or {}).items():
            mg.signature_def[key].CopyFrom(sig)
        if legacy_init_op:
            mg.collection_def["legacy_init_op"].node_list.value.append(legacy_init_op)
        self._meta_graph = mg
        self._model_desc = {
            "framework": "distributed_tensorflow_b200",
            "model": getattr(model, "name", type(model).__name__),
            "config": getattr(model, "export_config", lambda: {})(),
            "variables": sorted(tensors),
            "signature": {
                "inputs": {k: {"name": v[0], "dtype": np.dtype(v[1]).name, "shape": list(v[2])}
                           for k, v in model.signature_inputs.items()},
                "outputs": {k: {"name": v[0], "dtype": np.dtype(v[1]).name, "shape": list(v[2])}
                            for k, v in model.signature_outputs.items()},
            },
        }

    def save(self, as_text: bool = False) -> str:
        if self._meta_graph is None:
            raise AssertionError("add_meta_graph_and_variables() must be called before save()")
        mg_bytes = self._meta_graph.SerializeToString()
        # SavedModel { int64 saved_model_schema_version = 1; repeated MetaGraphDef meta_graphs = 2; }
        blob = b"\x08" + _varint(SAVED_MODEL_SCHEMA_VERSION) + b"\x12" + _varint(len(mg_bytes)) + mg_bytes
        path = os.path.join(self.export_dir, "saved_model.pb")
        with open(path, "wb") as f:
            f.write(blob)
        extra = os.path.join(self.export_dir, "assets.extra")
        os.makedirs(extra, exist_ok=True)
        with open(os.path.join(extra, "dtf_model.json"), "w") as f:
            json.dump(self._model_desc, f, indent=1, sort_keys=True)
        return path


def export_saved_model(sess, saved_model_path: str, model_version: int, model=None) -> str:
    """The reference's ``saved_model(sess, model_signature, legacy_init_op)`` (task.py:264-291)."""
    model = model if model is not None else sess.model
    export_path = os.path.join(saved_model_path, str(model_version))
    builder = SavedModelBuilder(export_path)
    builder.add_meta_graph_and_variables(
        sess, [SERVING], clear_devices=True,
        signature_def_map={DEFAULT_SERVING_SIGNATURE_DEF_KEY: build_signature_def(model)},
        legacy_init_op="legacy_init_op", model=model)
    builder.save()
    return export_path


def parse_saved_model_pb(path: str):
    """Returns (schema_version, [MetaGraphDef, ...])."""
    meta_graph_pb2 = _pb()[0]
    with open(path, "rb") as f:
        b = f.read()
    i, version, graphs = 0, 0, []

    def rd_varint(i):
        v = shift = 0
        while True:
            c = b[i]
            i += 1
            v |= (c & 0x7F) << shift
            shift += 7
            if not c & 0x80:
                return v, i
    while i < len(b):
        key, i = rd_varint(i)
        if key == 0x08:
            version, i = rd_varint(i)
        elif key == 0x12:
            n, i = rd_varint(i)
            mg = meta_graph_pb2.MetaGraphDef()
            mg.ParseFromString(b[i:i + n])
            graphs.append(mg)
            i += n
        else:
            raise ValueError("unexpected field key 0x%x in %s" % (key, path))
    return version, graphs


def load_saved_model(export_dir: str, tag: str = SERVING) -> dict:
    """Load ``{signature, variables, description}`` from one version directory."""
    version, graphs = parse_saved_model_pb(os.path.join(export_dir, "saved_model.pb"))
    mg = next((g for g in graphs if tag in g.meta_info_def.tags), None)
    if mg is None:
        raise ValueError("no MetaGraphDef tagged %r in %s" % (tag, export_dir))
    variables = _bundle.read_bundle(os.path.join(export_dir, "variables", "variables"))
    desc_path = os.path.join(export_dir, "assets.extra", "dtf_model.json")
    desc = None
    if os.path.exists(desc_path):
        with open(desc_path) as f:
            desc = json.load(f)
    return {"schema_version": version, "meta_graph": mg, "signatures": dict(mg.signature_def), "variables": variables,
            "description": desc}
