"""Serving latency / QPS sweep (BASELINE.json config 5): the JSON {keys, features} predict endpoint on
an exported model, batch 1 .. 4096.

Three layers are timed for the wide-and-deep export:
  kernel   : device time of the forward chain (CUDA events) -- fused gather + GEMMs + logit kernel
  predictor: WideDeepPredictor.predict on NumPy inputs (pinned H2D, kernels, D2H), wall clock
  http     : full POST round trip through the HTTP server incl. JSON encode/decode, wall clock
and the http layer for the reference's toy model (CPU).  Writes gpurun_out/serving.json.

    python bench/serving_bench.py [--rows 1000000] [--reps 200]
"""
import argparse
import json
import os
import sys
import tempfile
import threading
import time
import urllib.request

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def pct(xs, q):
    xs = sorted(xs)
    return xs[min(len(xs) - 1, int(q * len(xs)))]


def http_sweep(port, make_body, batches, reps):
    out = []
    for b in batches:
        body = json.dumps(make_body(b)).encode()
        req = lambda: urllib.request.urlopen(urllib.request.Request(
            "http://127.0.0.1:%d" % port, data=body, headers={"Content-Type": "application/json"}), timeout=30).read()
        for _ in range(5):
            req()
        n = max(10, reps // (1 + b // 256))
        ts = []
        for _ in range(n):
            t0 = time.perf_counter()
            req()
            ts.append((time.perf_counter() - t0) * 1e3)
        out.append({"batch": b, "p50_ms": pct(ts, 0.5), "p99_ms": pct(ts, 0.99), "qps": 1e3 / (sum(ts) / len(ts)),
                    "examples_per_s": b * 1e3 / (sum(ts) / len(ts))})
        print("http", out[-1], flush=True)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--reps", type=int, default=200)
    args = ap.parse_args()
    batches = [1, 4, 16, 64, 256, 1024, 4096]
    result = {}
    tmp = tempfile.mkdtemp()

    # ---- toy model (the reference's own serving scenario, CPU)
    from distributed_tensorflow_b200.export import export_saved_model
    from distributed_tensorflow_b200.models.linear import LinearModel
    from distributed_tensorflow_b200.optim import get_optimizer
    from distributed_tensorflow_b200.serving import ModelServer, make_http_server
    from distributed_tensorflow_b200.session import LocalSession
    s = LocalSession(LinearModel(), get_optimizer("sgd", 0.01))
    s.run_init()
    s.assign("weight", np.array(2.0, np.float32))
    s.assign("bias", np.array(10.0, np.float32))
    export_saved_model(s, os.path.join(tmp, "toy"), 1)
    ms = ModelServer(os.path.join(tmp, "toy"))
    httpd = make_http_server(ms, "127.0.0.1", 0)
    threading.Thread(target=httpd.serve_forever, daemon=True).start()
    result["toy_http"] = http_sweep(httpd.server_address[1],
                                    lambda b: {"keys": [[float(i)] for i in range(b)], "features": [[1.0]] * b}, batches, args.reps)
    httpd.shutdown()

    import torch
    if not torch.cuda.is_available():
        json.dump(result, open("gpurun_out/serving.json", "w"), indent=1)
        return
    # ---- wide-and-deep export served from the GPU
    from distributed_tensorflow_b200.models.wide_deep import WideDeepConfig, WideDeepTrainer
    from distributed_tensorflow_b200.models.wide_deep_trainer import _Dist, export_wide_deep
    cfg = WideDeepConfig(emb_rows=args.rows, batch=4096, optimizer="sgd")
    tr = WideDeepTrainer(cfg)
    tr.init_params()
    export_wide_deep(tr, _Dist(), os.path.join(tmp, "wd"), 1)
    # kernel-only latency on the trainer's plan
    g = torch.Generator().manual_seed(0)
    ids = torch.randint(0, cfg.emb_rows, (4096, cfg.slots), generator=g, dtype=torch.int32).cuda()
    dense = torch.randn(4096, cfg.dense_in, generator=g).cuda()
    kern = []
    for b in batches:
        for _ in range(5):
            tr.predict(ids, dense, n=b)
        ts = []
        for _ in range(args.reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); tr.predict(ids, dense, n=b); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        kern.append({"batch": b, "p50_ms": pct(ts, 0.5), "p99_ms": pct(ts, 0.99), "examples_per_s": b * 1e3 / pct(ts, 0.5)})
        print("kernel", kern[-1], flush=True)
    result["wide_deep_kernel"] = kern
    tr.close()
    ms = ModelServer(os.path.join(tmp, "wd"), device="cuda")
    feats = np.concatenate([ids.cpu().numpy().astype(np.float32), dense.cpu().numpy()], axis=1)
    keys = np.arange(4096, dtype=np.int32).reshape(-1, 1)
    pred = []
    for b in batches:
        inp = {"keys": keys[:b], "features": feats[:b]}
        for _ in range(5):
            ms._impl(inp)
        ts = []
        for _ in range(args.reps):
            t0 = time.perf_counter(); ms._impl(inp); ts.append((time.perf_counter() - t0) * 1e3)
        pred.append({"batch": b, "p50_ms": pct(ts, 0.5), "p99_ms": pct(ts, 0.99), "qps": 1e3 / pct(ts, 0.5),
                     "examples_per_s": b * 1e3 / pct(ts, 0.5)})
        print("predictor", pred[-1], flush=True)
    result["wide_deep_predictor"] = pred
    httpd = make_http_server(ms, "127.0.0.1", 0)
    threading.Thread(target=httpd.serve_forever, daemon=True).start()
    fl = feats.tolist()
    result["wide_deep_http"] = http_sweep(httpd.server_address[1],
                                          lambda b: {"keys": [[float(i)] for i in range(b)], "features": fl[:b]}, batches, args.reps)
    httpd.shutdown()
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(result, open("gpurun_out/serving.json", "w"), indent=1)


if __name__ == "__main__":
    main()
