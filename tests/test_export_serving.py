"""SavedModel-compatible export and the JSON predict endpoint (reference trainer/task.py:164-176,
264-291; README.md:29-41)."""
import json
import os
import threading
import urllib.request

import numpy as np
import pytest

from distributed_tensorflow_b200.export import (SavedModelBuilder, export_saved_model, export_session_bundle,
                                                load_saved_model)
from distributed_tensorflow_b200.export.saved_model import parse_saved_model_pb
from distributed_tensorflow_b200.checkpoint.bundle import BundleReader
from distributed_tensorflow_b200.models.linear import LinearModel
from distributed_tensorflow_b200.optim import get_optimizer
from distributed_tensorflow_b200.serving import ModelServer, make_http_server
from distributed_tensorflow_b200.session import LocalSession


def _trained(opt="adagrad"):
    s = LocalSession(LinearModel(), get_optimizer(opt, 0.1))
    s.run_init()
    s.assign("weight", np.array(2.0, np.float32))
    s.assign("bias", np.array(10.0, np.float32))
    s.assign("global_step", np.array(1000, np.int32))
    return s


def test_saved_model_layout_signature_and_variables(tmp_path):
    base = str(tmp_path / "saved_model")
    path = export_saved_model(_trained(), base, 1)
    assert path == os.path.join(base, "1")
    assert sorted(os.listdir(path)) == ["assets.extra", "saved_model.pb", "variables"]
    assert sorted(os.listdir(os.path.join(path, "variables"))) == ["variables.data-00000-of-00001", "variables.index"]
    version, graphs = parse_saved_model_pb(os.path.join(path, "saved_model.pb"))
    assert version == 1 and len(graphs) == 1
    mg = graphs[0]
    assert list(mg.meta_info_def.tags) == ["serve"]
    sig = mg.signature_def["serving_default"]
    assert sig.method_name == "tensorflow/serving/predict"
    assert sorted(sig.inputs) == ["features", "keys"] and sorted(sig.outputs) == ["keys", "prediction"]
    assert sig.inputs["keys"].dtype == 3 and sig.inputs["features"].dtype == 1          # DT_INT32 / DT_FLOAT
    assert [d.size for d in sig.inputs["features"].tensor_shape.dim] == [-1, 1]
    assert sig.outputs["keys"].name == "Identity:0" and sig.outputs["prediction"].name == "add:0"
    assert list(mg.collection_def["legacy_init_op"].node_list.value) == ["legacy_init_op"]
    assert all(n.device == "" for n in mg.graph_def.node)                                 # clear_devices=True
    names = {n.name for n in mg.graph_def.node}
    assert {"Placeholder", "Identity", "Placeholder_1", "weight", "bias", "global_step", "add"} <= names
    with BundleReader(os.path.join(path, "variables", "variables")) as r:                # incl. step + slots
        assert r.keys() == ["bias", "bias/Adagrad", "global_step", "weight", "weight/Adagrad"]


def test_builder_refuses_existing_directory(tmp_path):
    base = str(tmp_path / "sm")
    export_saved_model(_trained(), base, 3)
    with pytest.raises(AssertionError, match="already exists"):
        SavedModelBuilder(os.path.join(base, "3"))


def test_session_bundle_export_uses_legacy_aliases(tmp_path):
    p = export_session_bundle(_trained(), str(tmp_path / "legacy"), 7)
    assert os.path.basename(p) == "00000007"
    assert {"export.index", "export.data-00000-of-00001", "export.meta"} <= set(os.listdir(p))
    sig = json.load(open(os.path.join(p, "dtf_model.json")))["named_graph_signatures"]
    assert sig == {"inputs": ["X", "keys"], "outputs": ["keys", "predict"]}
    with pytest.raises(RuntimeError):
        export_session_bundle(_trained(), str(tmp_path / "legacy"), 7)


def test_predict_endpoint_readme_request(tmp_path):
    base = str(tmp_path / "saved_model")
    export_saved_model(_trained(), base, 1)
    s2 = _trained()
    s2.assign("weight", np.array(3.0, np.float32))
    export_saved_model(s2, base, 12)                                          # highest version wins
    ms = ModelServer(base)
    assert ms.version == 12
    httpd = make_http_server(ms, "127.0.0.1", 0)
    port = httpd.server_address[1]
    t = threading.Thread(target=httpd.serve_forever, daemon=True)
    t.start()
    try:
        body = json.dumps({"keys": [[11.0], [2.0]], "features": [[1], [2]]}).encode()   # README.md:40 verbatim
        req = urllib.request.Request("http://127.0.0.1:%d" % port, data=body, headers={"Content-Type": "application/json"})
        out = json.loads(urllib.request.urlopen(req, timeout=5).read())
        assert out == {"keys": [[11], [2]], "prediction": [[13.0], [16.0]]}
        env = json.dumps({"data": {"keys": [[1]], "features": [[0.5]]}}).encode()
        out = json.loads(urllib.request.urlopen(urllib.request.Request("http://127.0.0.1:%d" % port, data=env), timeout=5).read())
        assert out == {"keys": [[1]], "prediction": [[11.5]]}
        bad = urllib.request.Request("http://127.0.0.1:%d" % port, data=json.dumps({"keys": [[1]]}).encode())
        with pytest.raises(urllib.error.HTTPError) as ei:
            urllib.request.urlopen(bad, timeout=5)
        assert ei.value.code == 400 and "features" in json.loads(ei.value.read())["error"]
        info = json.loads(urllib.request.urlopen("http://127.0.0.1:%d" % port, timeout=5).read())
        assert info["version"] == 12 and info["signature_inputs"] == ["features", "keys"]
    finally:
        httpd.shutdown()
        httpd.server_close()


def test_load_saved_model_roundtrip(tmp_path):
    base = str(tmp_path / "sm")
    export_saved_model(_trained("sgd"), base, 1)
    m = load_saved_model(os.path.join(base, "1"))
    assert float(m["variables"]["weight"]) == 2.0 and m["description"]["model"] == "linear"
    assert m["description"]["signature"]["outputs"]["prediction"]["name"] == "add:0"


def test_wide_deep_numpy_predictor_assembles_shards_and_serves():
    """CPU restatement of the serving forward on synthetic shards from two ranks."""
    from distributed_tensorflow_b200.models.wide_deep_serving import WideDeepPredictor, assemble
    rng = np.random.RandomState(0)
    cfg = dict(emb_rows=11, emb_dim=128, slots=2, dense_in=3, dense_pad=64, h1=64, h2=64, h3=256, world=2)
    K0 = 2 * 128 + 64
    al = lambda v: (v + 255) // 256 * 256
    P = al(al(al(al(al(al(al(64 * K0) + 64 * 64) + 256 * 64) + 64) + 64) + 256) + 256) + 8
    P_pad = (P + 511) // 512 * 512
    flat = (rng.randn(P_pad) * 0.05).astype(np.float32)
    emb = rng.randn(11, 128).astype(np.float32)
    wide = rng.randn(11).astype(np.float32)
    variables = {"embedding/part_0": emb[0::2], "embedding/part_1": emb[1::2], "wide_weights/part_0": wide[0::2],
                 "wide_weights/part_1": wide[1::2], "dense_flat/part_0": flat[: P_pad // 2], "dense_flat/part_1": flat[P_pad // 2:]}
    assert np.array_equal(assemble(variables, "embedding", 2), emb)
    p = WideDeepPredictor(variables, cfg, device="cpu")
    feats = np.array([[3, 10, 0.5, -1.0, 2.0], [0, 0, 0.0, 0.0, 0.0]], np.float32)
    out = p.predict({"keys": [[7.0], [8.0]], "features": feats})
    assert out["keys"].tolist() == [[7], [8]] and out["prediction"].shape == (2, 1)
    assert np.all((out["prediction"] > 0) & (out["prediction"] < 1))
    x = np.zeros(64, np.float32); x[:3] = feats[0, 2:]
    a = np.concatenate([emb[3], emb[10], x])
    h = np.maximum(p.w1 @ a + p.b1, 0); h = np.maximum(p.w2 @ h + p.b2, 0); h = np.maximum(p.w3 @ h + p.b3, 0)
    logit = h @ p.w4 + p.b4 + wide[3] + wide[10] + p.wide_b
    assert np.isclose(out["prediction"][0, 0], 1 / (1 + np.exp(-logit)), rtol=1e-4)
    with pytest.raises(ValueError):
        p.predict({"keys": [[1]], "features": [[1.0, 2.0]]})


def _tiny_wide_deep_variables(rows=11, world=1, seed=0):
    rng = np.random.RandomState(seed)
    cfg = dict(emb_rows=rows, emb_dim=128, slots=2, dense_in=3, dense_pad=64, h1=64, h2=64, h3=256, world=world)
    K0 = 2 * 128 + 64
    al = lambda v: (v + 255) // 256 * 256
    P = al(al(al(al(al(al(al(64 * K0) + 64 * 64) + 256 * 64) + 64) + 64) + 256) + 256) + 8
    flat = (rng.randn((P + 255) // 256 * 256) * 0.05).astype(np.float32)
    emb = rng.randn(rows, 128).astype(np.float32)
    wide = rng.randn(rows).astype(np.float32)
    return {"embedding/part_0": emb, "wide_weights/part_0": wide, "dense_flat/part_0": flat}, cfg


def test_wide_deep_ids_stay_exact_and_inexact_float32_ids_are_refused():
    """Ids may come inside ``features`` (float64 end to end) or as the integer input ``ids``; a float32
    array that can no longer tell neighbouring ids apart (>= 2**24) is an error, not a silent wrong row."""
    from distributed_tensorflow_b200.models.wide_deep_serving import WideDeepPredictor
    variables, cfg = _tiny_wide_deep_variables()
    p = WideDeepPredictor(variables, cfg, device="cpu")
    feats = np.array([[3, 10, 0.5, -1.0, 2.0]], np.float64)
    a = p.predict({"keys": [[1]], "features": feats})["prediction"]
    b = p.predict({"keys": [[1]], "features": feats[:, 2:], "ids": np.array([[3, 10]], np.int64)})["prediction"]
    assert np.array_equal(a, b)
    big = np.array([[2 ** 24 + 1, 10, 0.5, -1.0, 2.0]], np.float64)       # exact in float64: row (2**24 + 1) % 11
    c = p.predict({"keys": [[1]], "features": big})["prediction"]
    d = p.predict({"keys": [[1]], "features": big[:, 2:], "ids": np.array([[(2 ** 24 + 1) % 11, 10]])})["prediction"]
    assert np.array_equal(c, d)
    with pytest.raises(ValueError, match="2\\*\\*24"):
        p.predict({"keys": [[1]], "features": big.astype(np.float32)})
    with pytest.raises(ValueError, match="integers"):
        p.predict({"keys": [[1]], "features": np.array([[3.5, 10, 0, 0, 0]])})


def test_wide_deep_http_concurrent_posts_each_get_their_own_answer(tmp_path):
    """Eight clients post different batches at once (ThreadingHTTPServer); every response must equal the
    predictor's answer for that client's own inputs."""
    import http.client
    import threading
    from distributed_tensorflow_b200.export.saved_model import DEFAULT_SERVING_SIGNATURE_DEF_KEY, SERVING, build_signature_def
    from distributed_tensorflow_b200.models.wide_deep_serving import WideDeepPredictor
    from distributed_tensorflow_b200.models.wide_deep_trainer import WideDeepServingSpec
    variables, cfg = _tiny_wide_deep_variables(rows=64)
    base = str(tmp_path / "wd")

    class _Store:
        model = WideDeepServingSpec(cfg)

        def snapshot(self):
            return dict(variables)

    builder = SavedModelBuilder(os.path.join(base, "1"))
    builder.add_meta_graph_and_variables(_Store(), [SERVING], signature_def_map={
        DEFAULT_SERVING_SIGNATURE_DEF_KEY: build_signature_def(_Store.model)}, model=_Store.model)
    builder.save()
    ms = ModelServer(base, device="cpu")
    ref = WideDeepPredictor(variables, cfg, device="cpu")
    httpd = make_http_server(ms, "127.0.0.1", 0)
    port = httpd.server_address[1]
    t = threading.Thread(target=httpd.serve_forever, daemon=True)
    t.start()
    results, errors = {}, []

    def client(i):
        try:
            rng = np.random.RandomState(100 + i)
            n = 3 + i
            feats = np.concatenate([rng.randint(0, 64, (n, 2)).astype(np.float64), rng.randn(n, 3)], axis=1)
            body = json.dumps({"keys": [[i * 100 + j] for j in range(n)], "features": feats.tolist()})
            for _ in range(5):
                c = http.client.HTTPConnection("127.0.0.1", port, timeout=20)
                c.request("POST", "/", body, {"Content-Type": "application/json"})
                r = json.loads(c.getresponse().read())
                c.close()
                want = ref.predict({"keys": [[i * 100 + j] for j in range(n)], "features": feats})
                assert r["keys"] == want["keys"].tolist()
                assert np.allclose(np.array(r["prediction"]), want["prediction"], rtol=1e-5, atol=1e-6)
            results[i] = True
        except Exception as e:  # noqa: BLE001
            errors.append((i, repr(e)))

    threads = [threading.Thread(target=client, args=(i,)) for i in range(8)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    httpd.shutdown()
    httpd.server_close()
    assert not errors and len(results) == 8, errors


def _check_graph_is_closed(graph_def):
    """Every input of every node (data or ^control, with or without :k) names a node of the graph."""
    names = {n.name for n in graph_def.node}
    assert len(names) == len(graph_def.node), "duplicate node names"
    for n in graph_def.node:
        for i in n.input:
            src = i.lstrip("^").split(":")[0]
            assert src in names, "%s reads %s which is not in the graph" % (n.name, i)
    return {n.name: n for n in graph_def.node}


def test_saved_model_graph_has_a_saver_subgraph_tf_could_run(tmp_path):
    """The three names in saver_def exist in graph_def and are wired like TF's Saver(sharded=True):
    save/Const is a string Const with a value, SaveV2/RestoreV2 carry the variable names, every variable has
    a save/Assign fed by RestoreV2:i, restore_all depends on all of them, init covers every <v>/Assign."""
    path = export_saved_model(_trained(), str(tmp_path / "sm"), 1)
    mg = parse_saved_model_pb(os.path.join(path, "saved_model.pb"))[1][0]
    nodes = _check_graph_is_closed(mg.graph_def)
    sd = mg.saver_def
    fname, save, restore = sd.filename_tensor_name, sd.save_tensor_name, sd.restore_op_name
    assert fname.split(":")[0] in nodes and save.split(":")[0] in nodes and restore in nodes
    const = nodes[fname.split(":")[0]]
    assert const.op == "Const" and const.attr["dtype"].type == 7 and list(const.attr["value"].tensor.string_val) == [b"model"]
    saved = ["bias", "bias/Adagrad", "global_step", "weight", "weight/Adagrad"]
    assert [s.decode() for s in nodes["save/SaveV2/tensor_names"].attr["value"].tensor.string_val] == saved
    assert list(nodes["save/SaveV2"].input[3:]) == saved
    assert list(nodes["save/SaveV2"].attr["dtypes"].list.type) == [1, 1, 3, 1, 1]
    assert nodes["save/RestoreV2"].input[0] == "save/Const"
    assigns = [n for n in mg.graph_def.node if n.name.startswith("save/Assign")]
    assert sorted(a.input[0] for a in assigns) == saved
    assert {a.input[1].split(":")[0] for a in assigns} == {"save/RestoreV2"}
    assert sorted(i for i in nodes["save/restore_shard"].input) == sorted("^" + a.name for a in assigns)
    assert list(nodes[restore].input) == ["^save/restore_shard"]
    assert "^save/SaveV2" in nodes["save/control_dependency"].input
    assert nodes[save.split(":")[0]].op == "Identity" and "^save/MergeV2Checkpoints" in nodes[save.split(":")[0]].input
    assert sorted(nodes["init"].input) == sorted("^%s/Assign" % v for v in saved)
    # forward graph of reference task.py:137-139 and the tensors the signature names
    assert nodes["add"].op == "Add" and list(nodes["add"].input) == ["Mul_1", "bias/read"]
    assert nodes["Sum"].op == "Sum" and nodes["weight/initial_value"].attr["value"].tensor.float_val[0] == 0.0
    for ti in list(mg.signature_def["serving_default"].inputs.values()) + list(mg.signature_def["serving_default"].outputs.values()):
        assert ti.name.split(":")[0] in nodes


def test_session_bundle_carries_a_signatures_proto(tmp_path):
    """export.meta holds tensorflow.serving.Signatures (Any-wrapped, collection serving_signatures) with the
    named generic signatures inputs{keys, X} / outputs{keys, predict} of reference task_supervisor.py:187-199."""
    from tensorboard.compat.proto import meta_graph_pb2
    from distributed_tensorflow_b200.export.session_bundle import decode_signatures, encode_signatures
    p = export_session_bundle(_trained(), str(tmp_path / "legacy"), 3)
    mg = meta_graph_pb2.MetaGraphDef()
    mg.ParseFromString(open(os.path.join(p, "export.meta"), "rb").read())
    anys = mg.collection_def["serving_signatures"].any_list.value
    assert len(anys) == 1 and anys[0].type_url == "type.googleapis.com/tensorflow.serving.Signatures"
    named = decode_signatures(anys[0].value)
    assert named == {"inputs": {"keys": "Placeholder:0", "X": "Placeholder_1:0"},
                     "outputs": {"keys": "Identity:0", "predict": "add:0"}}
    # wire format spot check: Signatures.named_signatures is field 2, a map entry {1: key, 2: Signature{3: Generic}}
    blob = encode_signatures({"inputs": {"X": "a:0"}})
    assert blob == (bytes([0x12, 24, 0x0a, 6]) + b"inputs" + bytes([0x12, 14, 0x1a, 12, 0x0a, 10, 0x0a, 1]) + b"X"
                    + bytes([0x12, 5, 0x0a, 3]) + b"a:0")
    _check_graph_is_closed(mg.graph_def)
    assert mg.collection_def["serving_graph"].any_list.value[0].type_url == "type.googleapis.com/tensorflow.GraphDef"
    assert mg.saver_def.restore_op_name == "save/restore_all"


def test_keep_alive_requests_do_not_pay_the_delayed_ack(tmp_path):
    """Head and body of a response leave in one segment: on a persistent connection the second small write used
    to wait for the client's delayed ACK (measured 44 ms per request, any body size under ~1 MB)."""
    import http.client
    import time
    base = str(tmp_path / "sm")
    export_saved_model(_trained("sgd"), base, 1)
    httpd = make_http_server(ModelServer(base), "127.0.0.1", 0)
    threading.Thread(target=httpd.serve_forever, daemon=True).start()
    c = http.client.HTTPConnection("127.0.0.1", httpd.server_address[1])
    body = json.dumps({"keys": [[1.0]], "features": [[2.0]]})
    ts = []
    for _ in range(30):
        t0 = time.perf_counter()
        c.request("POST", "/", body, {"Content-Type": "application/json"})
        assert json.loads(c.getresponse().read())["keys"] == [[1]]
        ts.append(time.perf_counter() - t0)
    httpd.shutdown()
    httpd.server_close()
    assert sorted(ts)[len(ts) // 2] < 0.02, sorted(ts)
