"""Legacy ``session_bundle`` export (the path the reference's Python-2 variant uses).

``exporter.Exporter(saver).init(graph_def, named_graph_signatures={"inputs": generic_signature(
{"keys", "X"}), "outputs": generic_signature({"keys", "predict"})}).export(dir, version, sess)``
(/root/reference/trainer/task_supervisor.py:185-200, 206-219; dead copy at trainer/task.py:294-307)
writes ``<export_dir>/<version as 8 digits>/export.meta`` plus a sharded checkpoint named
``export`` -- note the aliases differ from the SavedModel signature (``X``/``predict``).
"""
from __future__ import annotations

import json
import os

import numpy as np

from ..checkpoint import bundle as _bundle


def export_session_bundle(sess, export_dir: str, version: int, model=None) -> str:
    model = model if model is not None else sess.model
    path = os.path.join(export_dir, "%08d" % int(version))
    tmp = path + "-tmp"
    os.makedirs(tmp, exist_ok=True)
    tensors = sess.snapshot()
    _bundle.write_bundle(os.path.join(tmp, "export"), tensors)   # export.index + export.data-00000-of-00001
    # export.meta: MetaGraphDef whose "serving_signatures" collection carries the named signatures
    try:
        from tensorboard.compat.proto import meta_graph_pb2
        mg = meta_graph_pb2.MetaGraphDef()
        mg.meta_info_def.tags.append("serve")
        for v in sorted(tensors):
            n = mg.graph_def.node.add()
            n.name, n.op = v, "VariableV2"
        sigs = {"inputs": {a: model.signature_inputs[t][0] for a, t in model.legacy_inputs.items()},
                "outputs": {a: model.signature_outputs[t][0] for a, t in model.legacy_outputs.items()}}
        mg.collection_def["serving_signatures"].bytes_list.value.append(json.dumps(sigs, sort_keys=True).encode())
        with open(os.path.join(tmp, "export.meta"), "wb") as f:
            f.write(mg.SerializeToString())
    except ImportError:
        pass
    with open(os.path.join(tmp, "dtf_model.json"), "w") as f:
        json.dump({"model": getattr(model, "name", "model"),
                   "named_graph_signatures": {"inputs": sorted(model.legacy_inputs), "outputs": sorted(model.legacy_outputs)}},
                  f, indent=1)
    if os.path.exists(path):
        raise RuntimeError("Overwriting exports can cause corruption and are not allowed. Duplicate export dir: %s" % path)
    os.rename(tmp, path)
    return path
