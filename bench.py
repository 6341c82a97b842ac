#!/usr/bin/env python
"""Headline benchmark: training examples/sec of the sharded wide-and-deep model (BASELINE.json
configs 3/4: 10M-row embedding sharded over the ranks, 26 sparse ids + 13 dense features,
MLP 3392-1024-512-256-1, per-rank batch 8192 => global batch 64K at 8 ranks, weak scaling).

    python bench.py --gpus N --steps K --warmup W [--mode async|sync] [--impl ours|reference]

N > 1 is launched by ``python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N``
(one rank per GPU).  Rank 0 prints ONE JSON line.  `value` is device-timed (CUDA events around the
K steps, max over ranks) with inputs resident on the device; `e2e` repeats the measurement through
the public API with a pinned-host -> device copy of every step's inputs and a device -> host read of
the step's loss inside the timed region (GraphedTrainLoop: the copies are kernels of the replayed graph
that read / write the pinned host buffers over PCIe).  Synthetic data, random-init weights (no network).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "torch"])
    ap.add_argument("--mode", default="async", choices=["async", "sync"])
    ap.add_argument("--optimizer", default="adagrad")
    ap.add_argument("--batch", type=int, default=8192, help="examples per rank per step")
    ap.add_argument("--emb_rows", type=int, default=10_000_000)
    ap.add_argument("--pool", type=int, default=8, help="distinct synthetic batches cycled through")
    ap.add_argument("--no_graph", action="store_true", help="do not replay the step as a CUDA graph")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 50 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(float(r[1]) for r in self.rows if len(r) > 2 and r[1].replace(".", "").isdigit())
        mx = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = set()
        for r in self.rows:
            for n, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


def reference_arm(args) -> int:
    """The unmodified reference is a TensorFlow-1.x script; it can only run where TF imports."""
    sys.path.insert(0, os.path.join(REPO, "baseline", "_ref"))
    why = None
    try:
        import tensorflow  # noqa: F401
    except Exception as e:  # noqa: BLE001
        why = ("reference needs TensorFlow 1.x (tf.app.flags, tf.train.Supervisor, contrib.session_bundle): "
               "import tensorflow -> %s; no TF wheel in /opt/wheelhouse and none for Python 3.12" % type(e).__name__)
    if why is None:
        why = "reference trains a scalar linear model only; it has no wide-and-deep config to time"
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


def main() -> int:
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo", rank=rank, world_size=world)   # host control plane only

    def host_barrier():
        if world > 1:
            dist.barrier()

    if args.impl == "torch":
        from baseline.wide_deep_torch import run_baseline
        return run_baseline(args, rank, world, dev, ClockSampler)

    from distributed_tensorflow_b200.models.wide_deep import WideDeepConfig, WideDeepTrainer
    from distributed_tensorflow_b200.ops import _lib

    verbose = bool(os.environ.get("DTF_BENCH_VERBOSE"))

    progress = [time.time()]

    def say(msg):
        progress[0] = time.time()
        if verbose:
            print("[rank %d] %s" % (rank, msg), file=sys.stderr, flush=True)

    tr_box = []
    if verbose:
        import faulthandler
        limit = int(os.environ.get("DTF_BENCH_WATCHDOG", "45"))

        def watchdog():
            # no progress message for `limit` seconds: say where the host is and, with DTF_STAMPS=1, the last
            # phase marks the device reached
            while time.time() - progress[0] < limit:
                time.sleep(1.0)
            faulthandler.dump_traceback(file=sys.stderr)
            if tr_box and tr_box[0].stamps_count() > 0:
                n = tr_box[0].stamps_count()
                print("[rank %d] stuck; last phase marks (pending = not reached):\n%s" % (rank, tr_box[0].stamps_read(max(0, n - 90), n)),
                      file=sys.stderr, flush=True)
            os._exit(1)

        threading.Thread(target=watchdog, daemon=True).start()
    cfg = WideDeepConfig(emb_rows=args.emb_rows, batch=args.batch, optimizer=args.optimizer, sync=(args.mode == "sync"),
                         lr=0.05)
    tr = WideDeepTrainer(cfg, rank=rank, world=world, device=dev)
    tr_box.append(tr)
    say("trainer created")
    if world > 1:
        tr.attach_peers(dist)
    say("peers attached")
    tr.init_params()
    say("params initialised")
    torch.cuda.synchronize()
    host_barrier()

    # synthetic data of the named shape: a pool of pinned host batches + device-resident copies
    g = torch.Generator().manual_seed(1000 + rank)
    pool_h = []
    for _ in range(args.pool):
        ids = torch.randint(0, cfg.emb_rows, (cfg.batch, cfg.slots), generator=g, dtype=torch.int32).pin_memory()
        dense = torch.randn(cfg.batch, cfg.dense_in, generator=g).pin_memory()
        # learnable target (same rule as models/wide_deep_trainer.synthetic_batch): depends on two dense
        # features and on the parity of the first id, so the reported loss moves when training works
        logit = dense[:, 0] - 0.5 * dense[:, 1] + (ids[:, 0] % 2).float() - 0.5
        labels = (torch.rand(cfg.batch, generator=g) < torch.sigmoid(logit)).float().pin_memory()
        pool_h.append((ids, dense, labels))
    pool_d = [tuple(t.to(dev) for t in b) for b in pool_h]
    s_ids, s_dense, s_labels = (torch.empty_like(t, device=dev) for t in pool_h[0])   # static graph inputs
    loss_host = torch.zeros(1).pin_memory()
    h2d_bytes = sum(t.numel() * t.element_size() for t in pool_h[0])
    # high priority: when the async ps-side apply shares the SMs with the next step's compute, the
    # compute kernels' CTAs are placed first and the apply fills the remaining room
    stream = torch.cuda.Stream(device=dev, priority=-1)

    def load_static(b):
        s_ids.copy_(b[0], non_blocking=True)
        s_dense.copy_(b[1], non_blocking=True)
        s_labels.copy_(b[2], non_blocking=True)

    first_loss = None
    with torch.cuda.stream(stream):
        for i in range(max(3, args.warmup)):
            load_static(pool_d[i % args.pool])
            tr.step(s_ids, s_dense, s_labels)
            if first_loss is None:
                first_loss = float(tr.last_loss_sum()) / cfg.batch      # loss of the very first step (random init)
        stream.synchronize()
        say("eager warm-up done")
        host_barrier()
        graph = None
        G = 1                                         # steps per graph replay
        prefetch = args.mode == "async" and os.environ.get("DTF_PREFETCH", "1") != "0"
        if not args.no_graph:
            if args.mode == "async":
                # Async mode runs the ps-side applies on a second stream that overlaps the next step's
                # compute.  Several consecutive steps (each with its own static input set) are captured
                # into one graph so that the overlap survives inside a replay; the capture ends with a
                # join, so replays are self-contained.
                G = next(g for g in (10, 8, 5, 4, 2, 1) if args.steps % g == 0)
            statics = [(s_ids, s_dense, s_labels)] + [tuple(torch.empty_like(t) for t in (s_ids, s_dense, s_labels))
                                                      for _ in range(G - 1)]
            tr.join()
            stream.synchronize()
            tr.overlap_reset()
            graph = torch.cuda.CUDAGraph()
            stamp_range = [tr.stamps_count(), 0]
            _lib.reset_launch_count()
            with torch.cuda.graph(graph, stream=stream):
                for gi in range(G):
                    # async: step gi announces step gi+1's inputs, whose rows are then gathered on a low-priority
                    # stream during step gi (the last step of a replay announces nothing: replays stay self-contained)
                    nxt = statics[gi + 1][:2] if (prefetch and gi + 1 < G) else None
                    tr.step(*statics[gi], next_batch=nxt)
                tr.join()
            graph_launches = _lib.launch_count()          # kernels of the G captured steps (prefetch kernels included)
            stamp_range[1] = tr.stamps_count()
            tr.overlap_reset()
            graph.replay()
            stream.synchronize()
            say("graph captured and replayed once")

        def load_group(pool, base):
            for gi in range(G):
                b = pool[(base + gi) % args.pool]
                for dst, src in zip(statics[gi] if graph is not None else (s_ids, s_dense, s_labels), b):
                    dst.copy_(src, non_blocking=True)

        def one_step():
            if graph is not None:
                graph.replay()
            else:
                tr.step(s_ids, s_dense, s_labels)

        # ---------------- device-timed region: inputs already in HBM ----------------
        _lib.reset_launch_count()
        tr.step(s_ids, s_dense, s_labels)          # count the launches of one step outside the graph
        launches_per_step = _lib.launch_count()
        stream.synchronize()
        launches_total = (graph_launches * (args.steps // G)) if graph is not None else launches_per_step * args.steps
        # Start the clock sampler BEFORE the barrier (spawning nvidia-smi takes tens of ms; doing it
        # after the barrier on rank 0 only would make every other rank's timed region wait for it),
        # then align the GPUs with one untimed step -- its device-side barriers bring all streams to
        # the same point within microseconds -- and only then record the start event.
        sampler = ClockSampler(local_rank) if rank == 0 else None
        torch.cuda.synchronize()
        host_barrier()
        one_step()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        if graph is not None:
            for i in range(0, args.steps, G):
                load_group(pool_d, i)              # device -> device, keeps every step on fresh ids
                graph.replay()                     # G steps
                progress[0] = time.time()
        else:
            for i in range(args.steps):
                load_static(pool_d[i % args.pool])
                one_step()
        tr.join()                                  # the timed region ends when the last apply has finished
        e1.record(stream)
        stream.synchronize()
        host_barrier()
        dev_ms = e0.elapsed_time(e1)
        if os.environ.get("DTF_BENCH_VERBOSE"):
            print("[rank %d] device-timed arm done: %.3f ms/step" % (rank, dev_ms / args.steps), file=sys.stderr, flush=True)
        if tr.stamps_count() > 0 and graph is not None:
            os.makedirs(os.path.join(REPO, "gpurun_out"), exist_ok=True)
            with open(os.path.join(REPO, "gpurun_out", "stamps_%s_n%d_rank%d.txt" % (args.mode, world, rank)), "w") as f:
                f.write(tr.stamps_read(stamp_range[0], stamp_range[1]))
        # ---------------- end-to-end region: pinned host -> device every step, loss read back -----
        # Through the public training-loop API (models/wide_deep_loop.py): Ge steps per CUDA-graph
        # replay, the graph holding the pinned-host -> device copies of the Ge input batches (copy kernels
        # on a second stream, running ahead of the compute) and the device -> host copies of the Ge losses.
        from distributed_tensorflow_b200.models.wide_deep_loop import GraphedTrainLoop
        Ge = next(g for g in (10, 8, 5, 4, 2, 1) if args.steps % g == 0)
        tr.join()
        stream.synchronize()
        loop = GraphedTrainLoop(tr, [pool_h[i % args.pool] for i in range(Ge)], stream=stream)
        host_barrier()
        loop.run()
        loop.losses()
        t0 = time.perf_counter()
        e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e2.record(stream)
        last_loss = 0.0
        for i in range(0, args.steps, Ge):
            if os.environ.get("DTF_BENCH_VERBOSE") and i % (100 * Ge) == 0:
                print("[rank %d] e2e replay %d" % (rank, i // Ge), file=sys.stderr, flush=True)
            progress[0] = time.time()
            loop.run()                             # Ge steps: H2D of their inputs, compute, apply, D2H of the losses
            last_loss = loop.losses()[-1]          # the caller looks at every step's loss (waits for the replay)
        tr.join()
        e3.record(stream)
        stream.synchronize()
        host_barrier()
        e2e_ms = e2.elapsed_time(e3)
        e2e_wall_ms = (time.perf_counter() - t0) * 1e3
        clocks = sampler.stop() if sampler else None
        progress[0] = float("inf")                 # the stall watchdog (verbose runs) covers the timed regions only

    t_dev = torch.tensor([dev_ms, max(e2e_ms, e2e_wall_ms)], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t_dev, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = float(t_dev[0]), float(t_dev[1])
    global_batch = cfg.batch * world
    out = None
    if rank == 0:
        out = {
            "metric": "train examples/sec (device-timed, max over ranks)",
            "value": global_batch * args.steps / (dev_ms / 1e3),
            "unit": "examples/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": dev_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "impl": "ours",
            "config": {"model": "wide_and_deep 10M x128 embedding + MLP 3392-1024-512-256-1",
                       "global_batch": global_batch, "per_gpu_batch": cfg.batch, "seq_len": None,
                       "parallelism": "ps%d (row-sharded embedding + range-sharded dense, every rank worker+ps)" % world,
                       "mode": args.mode, "optimizer": args.optimizer, "sparse_ids": cfg.slots, "dense_features": cfg.dense_in,
                       "emb_rows": cfg.emb_rows, "cuda_graph": graph is not None, "steps_per_graph": G,
                       "param_prefetch": bool(prefetch and graph is not None and G > 1),
                       "l2": "per-step working set (55 MB gathered operand + random rows of a 2.5 GB table + 54 MB "
                             "mailbox) exceeds the 126 MB L2; ids change every step"},
            "e2e": {"value": global_batch * args.steps / (e2e_ms / 1e3), "unit": "examples/s",
                    "ms_per_step": e2e_ms / args.steps, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": 4, "steps_per_graph": Ge,
                    "first_loss": first_loss, "last_loss": last_loss},
            "gpu_launches": int(launches_total),
            "gpu_launches_per_step": round(launches_total / args.steps, 2),
            "clocks": clocks,
        }
    tr.close()

    # ---------------- after the timed regions: correctness on the real ranks -----------------------------
    # A small sync-replica job on the same ranks (fd-attached slabs, device barriers, NVLS when available)
    # against the fp32 PyTorch oracle on rank 0 (models/wide_deep_verify.py).  It runs after the numbers are
    # taken, under a watchdog that prints the line anyway, so it can never cost the measurement.
    post = {"verify": None, "vs_torch_baseline": None}
    saved_stdout_box = []                            # set while stdout is redirected (baseline phase), for the watchdog

    def emit():
        if rank == 0:
            out.update(post)
            print(json.dumps(out), flush=True)

    limit = int(os.environ.get("DTF_BENCH_POST_TIMEOUT", "240"))

    def bail():
        try:
            os.dup2(saved_stdout_box[0], 1) if saved_stdout_box else None
        except OSError:
            pass
        for k in post:
            if post[k] is None:
                post[k] = {"ok": None, "error": "did not finish within %d s" % limit}
        emit()
        os._exit(0)
    dog = threading.Timer(limit, bail)
    dog.daemon = True
    dog.start()

    def fail_soft(key, e):
        post[key] = {"ok": None, "error": "%s: %s" % (type(e).__name__, str(e)[:300])}
        dog.cancel()
        emit()
        os._exit(0)                                  # peers may still sit in a collective: leave without teardown

    if os.environ.get("DTF_BENCH_VERIFY", "1") == "0":
        post["verify"] = {"ok": None, "skipped": "DTF_BENCH_VERIFY=0"}
    else:
        try:
            from distributed_tensorflow_b200.models.wide_deep_verify import verify_multi_rank
            post["verify"] = verify_multi_rank(rank, world, dev, dist if world > 1 else None)
        except Exception as e:  # noqa: BLE001  (reported, never fatal for the measurement)
            fail_soft("verify", e)

    # Same-box comparison arm: the PyTorch NCCL + cuBLAS baseline (baseline/wide_deep_torch.py, none of this
    # repository's kernels) timed on the same ranks for the same mode, after our own numbers were taken.
    # The contract's `vs_baseline` stays null (the reference publishes no number); this is a separate key.
    if os.environ.get("DTF_BENCH_BASELINE", "1") == "0":
        post["vs_torch_baseline"] = {"skipped": "DTF_BENCH_BASELINE=0"}
    else:
        saved_stdout = os.dup(1)
        saved_stdout_box.append(saved_stdout)
        try:
            from baseline.wide_deep_torch import time_baseline
            sys.stdout.flush()
            os.dup2(2, 1)                            # NCCL prints its version banner on stdout; keep stdout to the one JSON line
            grp = dist.new_group(backend="nccl") if world > 1 else None
            r = time_baseline(args, rank, world, dev, group=grp, mode=args.mode, steps=min(args.steps, 20))
            base_value = global_batch / (r["ms_per_step"] / 1e3)
            post["vs_torch_baseline"] = {
                "impl": "torch-nccl-cublas (fixed-capacity all_to_all, flat bf16 all-reduce, CUDA graph)",
                "value": base_value, "unit": "examples/s", "ms_per_step": r["ms_per_step"], "steps": r["steps"],
                "ratio": (out["value"] / base_value) if rank == 0 else None,
                "e2e_ratio": (out["e2e"]["value"] / (global_batch / (r["e2e_ms_per_step"] / 1e3))) if rank == 0 else None,
                "cuda_graph": r["cuda_graph"], "why_eager": r["why_eager"], "dropped_ids": r["dropped_ids"],
                "first_loss": r["first_loss"], "last_loss": r["last_loss"]}
        except Exception as e:  # noqa: BLE001
            os.dup2(saved_stdout, 1)
            fail_soft("vs_torch_baseline", e)
        os.dup2(saved_stdout, 1)
    dog.cancel()
    emit()
    verify = post["verify"] or {}
    # ok == False with an error several times beyond the tolerance is a wrong result, not noise: fail the run
    gross = verify.get("ok") is False and float(verify.get("max_rel_err") or 0.0) > float(os.environ.get("DTF_VERIFY_FAIL_AT", "5"))
    if gross and rank == 0:
        print("verify FAILED: %s" % json.dumps(verify), file=sys.stderr, flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 3 if gross else 0


if __name__ == "__main__":
    sys.exit(main())
