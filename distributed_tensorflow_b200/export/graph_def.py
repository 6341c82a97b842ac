"""GraphDef of an exported model, written the way TensorFlow 1.x lays the same graph out.

The reference hands its graph to three consumers: ``tf.summary.FileWriter(output_path, sess.graph)``
(/root/reference/trainer/task.py:80,228), the Supervisor's ``graph.pbtxt`` (task.py:215-223) and the
MetaGraphDef inside ``saved_model.pb`` (task.py:275-289), whose ``saver_def`` names three graph
elements that a loader feeds and runs: the filename tensor, the save tensor and the restore op.
This module builds that graph from a model description so that every name the SignatureDef and the
SaverDef mention exists, with the op types and attributes TensorFlow's own loader expects:

* variables: ``<v>/initial_value`` (Const) -> ``<v>`` (VariableV2) -> ``<v>/Assign`` / ``<v>/read``
* toy forward + loss (task.py:130-140): ``Mul, sub, sub_1, Square, Const, Sum``, ``Mul_1, add``,
  the two scalar summaries and ``Merge/MergeSummary``
* ``init`` (NoOp over every ``<v>/Assign``) and ``legacy_init_op``
* the sharded saver sub-graph of ``Saver(sharded=True)``, which is what SavedModelBuilder uses:
  ``save/Const, save/StringJoin, save/ShardedFilename, save/SaveV2, save/control_dependency,
  save/MergeV2Checkpoints, save/Identity`` and ``save/RestoreV2, save/Assign*, save/restore_shard,
  save/restore_all``

Models without a TensorFlow graph (wide-and-deep: its forward is a chain of sm_100a kernels) get
placeholders for their inputs, variables, the saver sub-graph and one ``DtfModelOutput`` node per
output; their executable description lives in ``assets.extra/dtf_model.json``.
"""
from __future__ import annotations

from typing import Dict, Iterable, Optional, Tuple

import numpy as np

DT_FLOAT, DT_INT32, DT_STRING, DT_INT64, DT_BFLOAT16 = 1, 3, 7, 9, 14
_NP_TO_DT = {np.dtype(np.float32): DT_FLOAT, np.dtype(np.int32): DT_INT32, np.dtype(np.int64): DT_INT64}


def _dt(dtype) -> int:
    return _NP_TO_DT.get(np.dtype(dtype), DT_FLOAT)


class _G:
    def __init__(self, graph_def):
        self.g = graph_def

    def node(self, name: str, op: str, inputs: Iterable[str] = ()):
        n = self.g.node.add()
        n.name, n.op = name, op
        n.input.extend(inputs)
        return n

    def const(self, name: str, value, dtype=None, colocate: Optional[str] = None, splat_shape: Optional[tuple] = None):
        """Const node.  ``splat_shape``: tensor of that shape filled with the scalar ``value`` (a TensorProto
        with fewer values than elements repeats its last value), so big variables cost no bytes here."""
        n = self.node(name, "Const")
        t = n.attr["value"].tensor
        if dtype == "string":
            vals = [value] if isinstance(value, (str, bytes)) else list(value)
            n.attr["dtype"].type = t.dtype = DT_STRING
            if not isinstance(value, (str, bytes)):
                t.tensor_shape.dim.add().size = len(vals)
            t.string_val.extend(v.encode() if isinstance(v, str) else v for v in vals)
        else:
            arr = np.asarray(value, dtype=dtype)
            n.attr["dtype"].type = t.dtype = _dt(arr.dtype)
            for d in (splat_shape if splat_shape is not None else arr.shape):
                t.tensor_shape.dim.add().size = d
            if arr.size <= 1 or splat_shape is not None:
                (t.float_val if arr.dtype == np.float32 else (t.int64_val if arr.dtype == np.int64 else t.int_val)).append(
                    arr.reshape(-1)[0].item() if arr.size else 0)
            else:
                t.tensor_content = arr.tobytes()
        if colocate:
            n.attr["_class"].list.s.append(("loc:@" + colocate).encode())
        return n

    def placeholder(self, name: str, dtype, shape):
        n = self.node(name, "Placeholder")
        n.attr["dtype"].type = _dt(dtype)
        sh = n.attr["shape"].shape
        for d in shape:
            sh.dim.add().size = d
        return n

    def unary(self, name: str, op: str, inputs, dtype, colocate: Optional[str] = None):
        n = self.node(name, op, inputs)
        n.attr["T"].type = _dt(dtype) if not isinstance(dtype, int) else dtype
        if colocate:
            n.attr["_class"].list.s.append(("loc:@" + colocate).encode())
        return n

    def variable(self, name: str, dtype, shape, initial=None):
        """Const initial value, VariableV2, Assign, read -- the four nodes ``tf.Variable`` creates."""
        if initial is not None and np.asarray(initial).size <= 16 and np.asarray(initial).shape == tuple(shape):
            self.const(name + "/initial_value", np.asarray(initial, dtype), dtype)
        else:
            self.const(name + "/initial_value", np.zeros((), dtype), dtype, splat_shape=tuple(shape))
        v = self.node(name, "VariableV2")
        v.attr["dtype"].type = _dt(dtype)
        sh = v.attr["shape"].shape
        for d in shape:
            sh.dim.add().size = d
        v.attr["container"].s = b""
        v.attr["shared_name"].s = b""
        a = self.unary(name + "/Assign", "Assign", [name, name + "/initial_value"], dtype, colocate=name)
        a.attr["use_locking"].b = True
        a.attr["validate_shape"].b = True
        self.unary(name + "/read", "Identity", [name], dtype, colocate=name)


def add_saver_subgraph(graph_def, variables: Dict[str, Tuple[object, tuple]]) -> Dict[str, str]:
    """Append the ``Saver(sharded=True)`` sub-graph over ``variables`` (name -> (dtype, shape), saved in
    name order like TF) and return the three SaverDef names."""
    g = _G(graph_def)
    names = sorted(variables)
    dts = [_dt(variables[n][0]) for n in names]
    g.const("save/Const", "model", "string")
    g.const("save/StringJoin/inputs_1", "_temp_dtf_b200/part", "string")
    sj = g.node("save/StringJoin", "StringJoin", ["save/Const", "save/StringJoin/inputs_1"])
    sj.attr["N"].i = 2
    sj.attr["separator"].s = b""
    g.const("save/num_shards", 1, np.int32)
    g.const("save/ShardedFilename/shard", 0, np.int32)
    g.node("save/ShardedFilename", "ShardedFilename", ["save/StringJoin", "save/ShardedFilename/shard", "save/num_shards"])
    g.const("save/SaveV2/tensor_names", names, "string")
    g.const("save/SaveV2/shape_and_slices", [""] * len(names), "string")
    sv = g.node("save/SaveV2", "SaveV2", ["save/ShardedFilename", "save/SaveV2/tensor_names", "save/SaveV2/shape_and_slices"] + names)
    sv.attr["dtypes"].list.type.extend(dts)
    g.unary("save/control_dependency", "Identity", ["save/ShardedFilename", "^save/SaveV2"], DT_STRING, colocate="save/ShardedFilename")
    pk = g.unary("save/MergeV2Checkpoints/checkpoint_prefixes", "Pack", ["save/ShardedFilename", "^save/control_dependency"], DT_STRING)
    pk.attr["N"].i = 1
    pk.attr["axis"].i = 0
    mg = g.node("save/MergeV2Checkpoints", "MergeV2Checkpoints", ["save/MergeV2Checkpoints/checkpoint_prefixes", "save/Const"])
    mg.attr["delete_old_dirs"].b = True
    g.unary("save/Identity", "Identity", ["save/Const", "^save/control_dependency", "^save/MergeV2Checkpoints"], DT_STRING)
    g.const("save/RestoreV2/tensor_names", names, "string")
    g.const("save/RestoreV2/shape_and_slices", [""] * len(names), "string")
    rs = g.node("save/RestoreV2", "RestoreV2", ["save/Const", "save/RestoreV2/tensor_names", "save/RestoreV2/shape_and_slices"])
    rs.attr["dtypes"].list.type.extend(dts)
    assigns = []
    for i, n in enumerate(names):
        an = "save/Assign" if i == 0 else "save/Assign_%d" % i
        a = g.unary(an, "Assign", [n, "save/RestoreV2" if i == 0 else "save/RestoreV2:%d" % i], variables[n][0], colocate=n)
        a.attr["use_locking"].b = True
        a.attr["validate_shape"].b = True
        assigns.append("^" + an)
    g.node("save/restore_shard", "NoOp", assigns)
    g.node("save/restore_all", "NoOp", ["^save/restore_shard"])
    return {"filename_tensor_name": "save/Const:0", "save_tensor_name": "save/Identity:0", "restore_op_name": "save/restore_all"}


def build_graph_def(graph_def, model, tensors: Optional[Dict[str, np.ndarray]] = None) -> Dict[str, str]:
    """Fill ``graph_def`` for ``model``; ``tensors`` (a variable snapshot) adds optimizer slots and
    other saveable variables that are not part of the model's own list.  Returns the SaverDef names."""
    g = _G(graph_def)
    variables: Dict[str, Tuple[object, tuple]] = {}
    if getattr(model, "name", "") == "linear":
        g.placeholder("Placeholder", np.int32, (-1, 1))
        g.unary("Identity", "Identity", ["Placeholder"], np.int32)
        g.placeholder("Placeholder_1", np.float32, (-1, 1))
        g.placeholder("Placeholder_2", np.float32, (-1, 1))
        for v in model.variables:
            g.variable(v.name, v.dtype, tuple(v.initial.shape), v.initial)
            variables[v.name] = (v.dtype, tuple(v.initial.shape))
        g.unary("Mul", "Mul", ["Placeholder_1", "weight/read"], np.float32)
        g.unary("sub", "Sub", ["Placeholder_2", "Mul"], np.float32)
        g.unary("sub_1", "Sub", ["sub", "bias/read"], np.float32)
        g.unary("Square", "Square", ["sub_1"], np.float32)
        g.const("Const", np.array([0, 1], np.int32))
        s = g.unary("Sum", "Sum", ["Square", "Const"], np.float32)
        s.attr["Tidx"].type = DT_INT32
        s.attr["keep_dims"].b = False
        g.unary("Mul_1", "Mul", ["Placeholder_1", "weight/read"], np.float32)
        g.unary("add", "Add", ["Mul_1", "bias/read"], np.float32)
        g.const("loss/tags", "loss", "string")
        g.unary("loss", "ScalarSummary", ["loss/tags", "Sum"], np.float32)
        m = g.node("Merge/MergeSummary", "MergeSummary", ["loss"])
        m.attr["N"].i = 1
    else:
        for alias, (tname, dt, shape) in model.signature_inputs.items():
            g.placeholder(tname.split(":")[0], dt, shape)
        for alias, (tname, dt, shape) in model.signature_outputs.items():
            g.unary(tname.split(":")[0], "DtfModelOutput", [], dt)
    for name in sorted(tensors or {}):
        if name not in variables:
            arr = np.asarray(tensors[name])
            dt = arr.dtype if arr.dtype in _NP_TO_DT else np.dtype(np.float32)
            g.variable(name, dt, tuple(arr.shape), arr)
            variables[name] = (dt, tuple(arr.shape))
    g.node("init", "NoOp", ["^%s/Assign" % n for n in variables])
    g.node("legacy_init_op", "NoOp")
    return add_saver_subgraph(graph_def, variables)


def graph_def_text(graph_def) -> str:
    """Text form, as ``tf.train.write_graph(..., 'graph.pbtxt')`` writes it."""
    from google.protobuf import text_format
    return text_format.MessageToString(graph_def)


def variable_def_bytes(name: str, trainable: bool) -> bytes:
    """Serialized ``VariableDef`` (tensorflow/core/framework/variable.proto), the items of the ``variables`` /
    ``trainable_variables`` collections that ``tf.train.import_meta_graph`` turns back into Variable objects:
    variable_name = 1, initializer_name = 2, snapshot_name = 3, initial_value_name = 6, trainable = 7."""
    def field(num: int, text: str) -> bytes:
        b = text.encode()
        assert len(b) < 128
        return bytes([(num << 3) | 2, len(b)]) + b
    out = field(1, name + ":0") + field(2, name + "/Assign") + field(3, name + "/read:0") + field(6, name + "/initial_value:0")
    if trainable:
        out += bytes([(7 << 3) | 0, 1])
    return out


def fill_meta_graph(mg, model, tensors: Optional[Dict[str, np.ndarray]], tags=("train",), sharded: bool = True,
                    max_to_keep: int = 5) -> None:
    """MetaGraphDef as ``Saver.save`` writes it next to a checkpoint (``<prefix>.meta``): the model's graph with the
    saver sub-graph, a saver_def that names nodes of that graph, and the variable collections."""
    from tensorboard.compat.proto import saver_pb2
    mg.meta_info_def.tensorflow_version = "1.15.0"
    mg.meta_info_def.tensorflow_git_version = "distributed_tensorflow_b200"
    mg.meta_info_def.tags.extend(tags)
    names = build_graph_def(mg.graph_def, model, tensors)
    mg.graph_def.versions.producer = 134
    sd = mg.saver_def
    sd.filename_tensor_name = names["filename_tensor_name"]
    sd.save_tensor_name = names["save_tensor_name"]
    sd.restore_op_name = names["restore_op_name"]
    sd.max_to_keep = max_to_keep
    sd.sharded = sharded
    sd.keep_checkpoint_every_n_hours = 10000.0
    sd.version = saver_pb2.SaverDef.V2
    trainable = set(getattr(model, "trainable_names", lambda: [])())
    var_nodes = [n.name for n in mg.graph_def.node if n.op == "VariableV2"]
    for n in var_nodes:
        mg.collection_def["variables"].bytes_list.value.append(variable_def_bytes(n, n in trainable))
        if n in trainable:
            mg.collection_def["trainable_variables"].bytes_list.value.append(variable_def_bytes(n, True))
