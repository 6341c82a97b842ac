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
