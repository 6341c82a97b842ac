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


def _len_delim(field: int, payload: bytes) -> bytes:
    out = bytearray([(field << 3) | 2])
    n = len(payload)
    while n >= 0x80:
        out.append((n & 0x7F) | 0x80)
        n >>= 7
    out.append(n)
    return bytes(out) + payload


def encode_signatures(named: dict) -> bytes:
    """Serialized ``tensorflow.serving.Signatures`` (session_bundle/manifest.proto) holding
    ``named_signatures[name] = Signature{generic_signature{map[alias] = TensorBinding{tensor_name}}}``::

        Signatures       { Signature default_signature = 1; map<string, Signature> named_signatures = 2; }
        Signature        { oneof type { ...; GenericSignature generic_signature = 3; } }
        GenericSignature { map<string, TensorBinding> map = 1; }
        TensorBinding    { string tensor_name = 1; }

    This is what ``exporter.generic_signature`` / ``Exporter.init(named_graph_signatures=...)`` build in
    the reference (trainer/task_supervisor.py:187-199)."""
    out = b""
    for name in sorted(named):
        generic = b""
        for alias in sorted(named[name]):
            binding = _len_delim(1, named[name][alias].encode())
            generic += _len_delim(1, _len_delim(1, alias.encode()) + _len_delim(2, binding))
        signature = _len_delim(3, generic)
        out += _len_delim(2, _len_delim(1, name.encode()) + _len_delim(2, signature))
    return out


def decode_signatures(blob: bytes) -> dict:
    """Inverse of :func:`encode_signatures` (named generic signatures only)."""
    def fields(b):
        i = 0
        while i < len(b):
            key = b[i]; i += 1
            n = shift = 0
            while True:
                c = b[i]; i += 1
                n |= (c & 0x7F) << shift
                shift += 7
                if not c & 0x80:
                    break
            yield key >> 3, b[i:i + n]
            i += n
    named = {}
    for f, entry in fields(blob):
        if f != 2:
            continue
        kv = dict(fields(entry))
        sig = dict(fields(kv[2]))
        bindings = {}
        for _, m in fields(sig.get(3, b"")):
            mkv = dict(fields(m))
            bindings[mkv[1].decode()] = dict(fields(mkv[2]))[1].decode()
        named[kv[1].decode()] = bindings
    return named


SIGNATURES_TYPE_URL = "type.googleapis.com/tensorflow.serving.Signatures"
GRAPH_TYPE_URL = "type.googleapis.com/tensorflow.GraphDef"


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
        from .graph_def import build_graph_def
        names = build_graph_def(mg.graph_def, model, tensors)
        mg.graph_def.versions.producer = 134
        mg.saver_def.filename_tensor_name = names["filename_tensor_name"]
        mg.saver_def.save_tensor_name = names["save_tensor_name"]
        mg.saver_def.restore_op_name = names["restore_op_name"]
        mg.saver_def.sharded = True
        mg.saver_def.version = 2
        sigs = {"inputs": {a: model.signature_inputs[t][0] for a, t in model.legacy_inputs.items()},
                "outputs": {a: model.signature_outputs[t][0] for a, t in model.legacy_outputs.items()}}
        # collections as the legacy Exporter fills them: Any-wrapped protos under the serving_* keys
        any_sig = mg.collection_def["serving_signatures"].any_list.value.add()
        any_sig.type_url = SIGNATURES_TYPE_URL
        any_sig.value = encode_signatures(sigs)
        any_graph = mg.collection_def["serving_graph"].any_list.value.add()
        any_graph.type_url = GRAPH_TYPE_URL
        any_graph.value = mg.graph_def.SerializeToString()
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
