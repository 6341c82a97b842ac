#!/usr/bin/env python
"""HTTP + JSON + validation cost of the wide-and-deep predict endpoint with the model taken out (a stub
predictor that only checks the ids and echoes a column), product request shape: 26 sparse ids + 13 dense
values per example, one keep-alive client.  Runs on the CPU; what remains of the end-to-end HTTP latency
after the ~0.1-0.25 ms GPU predictor (BASELINE.md serving table).  Writes profiles/serving_http_r2.json."""
import http.client
import json
import os
import sys
import tempfile
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def main():
    from distributed_tensorflow_b200.export.saved_model import (DEFAULT_SERVING_SIGNATURE_DEF_KEY, SERVING, SavedModelBuilder,
                                                                build_signature_def)
    import distributed_tensorflow_b200.models.wide_deep_serving as wds
    from distributed_tensorflow_b200.models.wide_deep_trainer import WideDeepServingSpec
    from distributed_tensorflow_b200.serving import ModelServer, make_http_server
    cfg = dict(emb_rows=10_000_000, emb_dim=128, slots=26, dense_in=13, dense_pad=64, h1=64, h2=64, h3=256, world=1)
    variables = {"embedding/part_0": np.zeros((4, 128), np.float32), "wide_weights/part_0": np.zeros(4, np.float32),
                 "dense_flat/part_0": np.zeros(8, np.float32)}
    tmp = tempfile.mkdtemp()

    class Store:
        model = WideDeepServingSpec(cfg)

        def snapshot(self):
            return dict(variables)
    b = SavedModelBuilder(os.path.join(tmp, "1"))
    b.add_meta_graph_and_variables(Store(), [SERVING], signature_def_map={
        DEFAULT_SERVING_SIGNATURE_DEF_KEY: build_signature_def(Store.model)}, model=Store.model)
    b.save()
    real = wds.WideDeepPredictor

    class Stub(real):
        def __init__(self, variables, config, device="auto", max_batch=4096):
            self.config, self.slots, self.dense_in, self.emb_rows = dict(config), 26, 13, config["emb_rows"]
            self.backend = "stub"

        def _predict_cpu(self, ids, dense):
            return dense[:, 0]
    wds.WideDeepPredictor = Stub
    ms = ModelServer(tmp, device="cpu")
    httpd = make_http_server(ms, "127.0.0.1", 0)
    port = httpd.server_address[1]
    threading.Thread(target=httpd.serve_forever, daemon=True).start()
    rng = np.random.RandomState(0)
    out = {}
    for n in (1, 16, 256, 1024, 4096):
        feats = np.concatenate([rng.randint(0, 10_000_000, (n, 26)).astype(np.float64), rng.randn(n, 13).round(4)], 1)
        body = json.dumps({"keys": [[i] for i in range(n)], "features": feats.tolist()})
        c = http.client.HTTPConnection("127.0.0.1", port)
        ts = []
        for _ in range(200 if n <= 256 else 40):
            t0 = time.perf_counter()
            c.request("POST", "/", body, {"Content-Type": "application/json"})
            r = c.getresponse().read()
            ts.append(time.perf_counter() - t0)
        assert len(json.loads(r)["prediction"]) == n
        ts.sort()
        out[str(n)] = {"p50_ms": round(ts[len(ts) // 2] * 1e3, 3), "p99_ms": round(ts[int(len(ts) * 0.99) - 1] * 1e3, 3),
                       "request_bytes": len(body), "qps_one_client": round(1.0 / ts[len(ts) // 2], 1)}
    httpd.shutdown()
    res = {"what": "HTTP + JSON + id validation only (stub predictor), 26 ids + 13 dense per example, keep-alive client, CPU",
           "batches": out}
    with open(os.path.join(REPO, "profiles", "serving_http_r2.json"), "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
