"""CPU parameter-server transport numbers (SURVEY.md work-list row T1 and the toy-model rows):

* PUSH_PULL round trip of the toy model's step (two float32 scalars + global_step) over loopback TCP;
* the same round trip for large dense pushes (1 MB and 16 MB of gradients) -> applied bytes / s;
* concurrent workers against one ps task (Hogwild apply);
* toy-model training throughput: standalone session and ps session (steps / s).

    python bench/ps_transport_bench.py          # CPU only, writes gpurun_out/ps_transport.json
"""
import json
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from distributed_tensorflow_b200.cluster import ClusterSpec, Server
from distributed_tensorflow_b200.models.linear import LinearModel
from distributed_tensorflow_b200.optim import get_optimizer
from distributed_tensorflow_b200.session import LocalSession, PSClient, PSSession


def pct(xs, q):
    xs = sorted(xs)
    return xs[min(len(xs) - 1, int(q * len(xs)))]


def main():
    out = {}
    srv = Server(ClusterSpec({"ps": ["127.0.0.1:0"]}), "ps", 0)
    addr = "127.0.0.1:%d" % srv.port
    try:
        c = PSClient(addr)
        c.init_var("weight", np.zeros((), np.float32), overwrite=True)
        c.init_var("bias", np.zeros((), np.float32), overwrite=True)
        c.init_var("global_step", np.zeros((), np.int64), overwrite=True)
        g = [np.float32(0.01).reshape(()), np.float32(0.02).reshape(())]
        for _ in range(2000):
            c.push_pull(["weight", "bias"], g, 0, 0.01, "global_step")
        ts = []
        for _ in range(20000):
            t0 = time.perf_counter()
            c.push_pull(["weight", "bias"], g, 0, 0.01, "global_step")
            ts.append((time.perf_counter() - t0) * 1e6)
        out["toy_step_rtt_us"] = {"p50": pct(ts, 0.5), "p99": pct(ts, 0.99), "steps_per_s": 1e6 / (sum(ts) / len(ts))}
        print("toy step round trip", out["toy_step_rtt_us"], flush=True)

        for mb in (1, 16):
            n = mb * (1 << 20) // 4
            c.init_var("big%d" % mb, np.zeros(n, np.float32), overwrite=True)
            grad = np.full(n, 1e-3, np.float32)
            for opt_name, opt_id in (("sgd", 0), ("adagrad", 2), ("adam", 3)):
                for _ in range(3):
                    c.push_pull(["big%d" % mb], [grad], opt_id, 0.01, None)
                t0 = time.perf_counter()
                reps = 40 if mb == 1 else 8
                for _ in range(reps):
                    c.push_pull(["big%d" % mb], [grad], opt_id, 0.01, None)
                dt = (time.perf_counter() - t0) / reps
                out["push_pull_%dMB_%s" % (mb, opt_name)] = {"ms": dt * 1e3, "GBps_each_way": mb / 1024 / dt}
                print("push_pull %2d MB %-8s %.2f ms  (%.2f GB/s each way, apply included)" % (mb, opt_name, dt * 1e3, mb / 1024 / dt), flush=True)

        def worker(k, res):
            cl = PSClient(addr)
            t0 = time.perf_counter()
            for _ in range(5000):
                cl.push_pull(["weight", "bias"], g, 0, 0.01, "global_step")
            res[k] = 5000 / (time.perf_counter() - t0)
            cl.close()

        for nw in (2, 4, 8):
            res = {}
            th = [threading.Thread(target=worker, args=(k, res)) for k in range(nw)]
            for t in th:
                t.start()
            for t in th:
                t.join()
            out["concurrent_%d_workers_steps_per_s" % nw] = sum(res.values())
            print("%d workers: %.0f steps/s in total" % (nw, sum(res.values())), flush=True)
        c.close()

        x = np.linspace(-1, 1, 100)
        y = 2 * x + np.random.RandomState(0).randn(100) * 0.33 + 10
        for name, sess in (("standalone", LocalSession(LinearModel(), get_optimizer("sgd", 0.01))),
                           ("ps_session", PSSession(LinearModel(), get_optimizer("sgd", 0.01), ClusterSpec({"ps": [addr]})))):
            sess.run_init()
            t0 = time.perf_counter()
            for _ in range(10):
                for a, b in zip(x, y):
                    sess.train_step(np.array([[a]]), np.array([[b]]))
            dt = time.perf_counter() - t0
            out["toy_training_%s" % name] = {"steps_per_s": 1000 / dt, "w": float(sess.value("weight")), "b": float(sess.value("bias"))}
            print("toy training %-10s %.0f steps/s  w=%.3f b=%.3f" % (name, 1000 / dt, float(sess.value("weight")), float(sess.value("bias"))), flush=True)
            sess.close()
    finally:
        srv.stop()
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/ps_transport.json", "w"), indent=1)


if __name__ == "__main__":
    main()
