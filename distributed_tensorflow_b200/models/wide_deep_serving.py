"""Predictor for exported wide-and-deep models (the serving-batch forward).

On a GPU the exported shards are loaded into a single-rank plan and requests run the same fused
kernels as training (embedding gather + tcgen05 GEMM chain + logit kernel) with the request's batch
size; without a GPU a NumPy restatement serves (slow, for tests and tooling).  Serves the
``{keys, features}`` endpoint of the reference README (README.md:37-41): ``features`` holds, per
example, the F sparse ids followed by the dense features.
"""
from __future__ import annotations

import os
import threading
from typing import Dict

import numpy as np

# float32 represents every integer only below 2**24; ids at or above that must arrive as integers / float64
_F32_EXACT = 1 << 24


def assemble(variables: Dict[str, np.ndarray], name: str, world: int) -> np.ndarray:
    """Undo ``row r -> part_(r % world)[r // world]`` / range sharding."""
    parts = [variables["%s/part_%d" % (name, r)] for r in range(world)]
    if name == "dense_flat":
        return np.concatenate(parts)
    from ..placement import RowShard
    rows = sum(p.shape[0] for p in parts)
    shard = RowShard(rows=rows, world=world)
    out = np.empty((rows,) + parts[0].shape[1:], dtype=parts[0].dtype)
    for r, p in enumerate(parts):
        if p.shape[0] != shard.local_rows(r):
            raise ValueError("%s/part_%d has %d rows, the row sharding of %d rows over %d ranks gives it %d"
                             % (name, r, p.shape[0], rows, world, shard.local_rows(r)))
        out[shard.global_row(r, 0)::world] = p          # local row l of rank r is global row l * world + r
    return out


class WideDeepPredictor:
    def __init__(self, variables: Dict[str, np.ndarray], config: dict, device: str = "auto", max_batch: int = 4096):
        self.config = dict(config)
        world = int(config.get("world", 1))
        self.slots, self.dense_in = int(config["slots"]), int(config["dense_in"])
        emb = assemble(variables, "embedding", world)
        wide = assemble(variables, "wide_weights", world)
        flat = assemble(variables, "dense_flat", world)
        use_gpu = device in ("auto", "cuda")
        if use_gpu:
            try:
                import torch
                use_gpu = torch.cuda.is_available()
            except Exception:  # noqa: BLE001
                use_gpu = False
        if device == "cuda" and not use_gpu:
            raise RuntimeError("device=cuda requested but no GPU is visible")
        self.max_batch = max_batch
        self.emb_rows = int(config.get("emb_rows", emb.shape[0]))
        # One set of staging buffers and one plan serve every request: the HTTP layer is threaded
        # (serving/server.py), so the GPU path is serialised here.  Requests queue for ~0.1 ms each.
        self._gpu_lock = threading.Lock()
        if use_gpu:
            self._init_gpu(emb, wide, flat)
        else:
            self._init_cpu(emb, wide, flat)

    # ------------------------------------------------------------------ GPU
    def _init_gpu(self, emb, wide, flat):
        import torch
        from .wide_deep import WideDeepConfig, WideDeepTrainer
        c = self.config
        cfg = WideDeepConfig(emb_rows=int(c["emb_rows"]), emb_dim=int(c["emb_dim"]), slots=self.slots, dense_in=self.dense_in,
                             dense_pad=int(c["dense_pad"]), h1=int(c["h1"]), h2=int(c["h2"]), h3=int(c["h3"]),
                             batch=self.max_batch, optimizer="sgd", sync=False)
        self._torch = torch
        self._tr = WideDeepTrainer(cfg)
        n = self._tr.info.dense_hi - self._tr.info.dense_lo
        dense = np.zeros(n, np.float32)
        dense[: min(n, flat.size)] = flat[:n]
        self._tr.load_shards({"embedding/part_0": emb, "wide_weights/part_0": wide, "dense_flat/part_0": dense})
        dev = self._tr.device
        self._ids = torch.zeros(self.max_batch, self.slots, dtype=torch.int32, device=dev)
        self._dense = torch.zeros(self.max_batch, self.dense_in, dtype=torch.float32, device=dev)
        self._h_ids = torch.zeros(self.max_batch, self.slots, dtype=torch.int32).pin_memory()
        self._h_dense = torch.zeros(self.max_batch, self.dense_in, dtype=torch.float32).pin_memory()
        self._h_out = torch.zeros(self.max_batch, dtype=torch.float32).pin_memory()
        # One CUDA graph per batch bucket: host -> device copies of the bucket's rows, the forward kernel chain and the
        # device -> host copy of the probabilities replay as ONE launch (a request of n rows runs in the smallest
        # bucket >= n; the rows past n hold the previous request's values, which are valid inputs, and are ignored).
        # DTF_SERVE_GRAPHS=0 keeps the eager launches.
        self._stream = torch.cuda.Stream(device=dev)
        self._buckets = [b for b in (1, 4, 16, 64, 256, 1024, 4096) if b < self.max_batch] + [self.max_batch]
        self._graphs = {}
        self._use_graphs = os.environ.get("DTF_SERVE_GRAPHS", "1") != "0"
        self.backend = "cuda"

    def _enqueue(self, n: int) -> None:
        self._ids[:n].copy_(self._h_ids[:n], non_blocking=True)
        self._dense[:n].copy_(self._h_dense[:n], non_blocking=True)
        p = self._tr.predict(self._ids, self._dense, n=n)
        self._h_out[:n].copy_(p, non_blocking=True)

    def _graph_for(self, bucket: int):
        """The bucket's graph, captured on first use (after one eager run, so that nothing loads lazily under
        capture); None if capture is refused -- the caller then launches eagerly."""
        if bucket in self._graphs:
            return self._graphs[bucket]
        torch = self._torch
        g = None
        try:
            self._enqueue(bucket)
            self._stream.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=self._stream):
                self._enqueue(bucket)
        except Exception:  # noqa: BLE001
            g = None
            self._stream.synchronize()
        self._graphs[bucket] = g
        return g

    def _predict_gpu(self, ids: np.ndarray, dense: np.ndarray) -> np.ndarray:
        torch = self._torch
        out = np.empty(ids.shape[0], np.float32)
        with self._gpu_lock:
            self._predict_gpu_locked(torch, ids, dense, out)
        return out

    def _predict_gpu_locked(self, torch, ids, dense, out) -> None:
        for lo in range(0, ids.shape[0], self.max_batch):
            n = min(self.max_batch, ids.shape[0] - lo)
            self._h_ids[:n].copy_(torch.from_numpy(ids[lo:lo + n]))
            self._h_dense[:n].copy_(torch.from_numpy(dense[lo:lo + n]))
            with torch.cuda.stream(self._stream):
                g = None
                if self._use_graphs:
                    g = self._graph_for(next(b for b in self._buckets if b >= n))
                if g is not None:
                    g.replay()
                else:
                    self._enqueue(n)
                self._stream.synchronize()
            out[lo:lo + n] = self._h_out[:n].numpy()

    # ------------------------------------------------------------------ CPU
    def _init_cpu(self, emb, wide, flat):
        c = self.config
        K0 = self.slots * int(c["emb_dim"]) + int(c["dense_pad"])
        h1, h2, h3 = int(c["h1"]), int(c["h2"]), int(c["h3"])
        al = lambda v: (v + 255) // 256 * 256
        o = 0
        self.w1 = flat[o:o + h1 * K0].reshape(h1, K0); o = al(o + h1 * K0)
        self.w2 = flat[o:o + h2 * h1].reshape(h2, h1); o = al(o + h2 * h1)
        self.w3 = flat[o:o + h3 * h2].reshape(h3, h2); o = al(o + h3 * h2)
        self.b1 = flat[o:o + h1]; o = al(o + h1)
        self.b2 = flat[o:o + h2]; o = al(o + h2)
        self.b3 = flat[o:o + h3]; o = al(o + h3)
        self.w4 = flat[o:o + h3]; o = al(o + h3)
        self.b4 = flat[o]; self.wide_b = flat[o + 4]
        self.emb, self.wide = emb, wide
        self.backend = "numpy"

    def _predict_cpu(self, ids: np.ndarray, dense: np.ndarray) -> np.ndarray:
        c = self.config
        B = ids.shape[0]
        x = np.zeros((B, int(c["dense_pad"])), np.float32)
        x[:, : self.dense_in] = dense
        a = np.concatenate([self.emb[ids].reshape(B, -1), x], axis=1)
        h = np.maximum(a @ self.w1.T + self.b1, 0)
        h = np.maximum(h @ self.w2.T + self.b2, 0)
        h = np.maximum(h @ self.w3.T + self.b3, 0)
        logit = h @ self.w4 + self.b4 + self.wide[ids].sum(1) + self.wide_b
        return (1.0 / (1.0 + np.exp(-logit))).astype(np.float32)

    # ------------------------------------------------------------------ endpoint
    def predict(self, inputs: Dict[str, np.ndarray]) -> Dict[str, np.ndarray]:
        keys = np.asarray(inputs["keys"]).astype(np.int32)
        feats = np.asarray(inputs["features"])
        if feats.ndim != 2:
            raise ValueError("features must be a [batch, n] array")
        if "ids" in inputs:
            # integer ids travel on their own (exact for any table size); features then holds only the dense part
            ids = np.asarray(inputs["ids"])
            if ids.ndim != 2 or ids.shape[1] != self.slots or ids.shape[0] != feats.shape[0]:
                raise ValueError("ids must be [batch, %d]" % self.slots)
            if feats.shape[1] != self.dense_in:
                raise ValueError("with ids given, features must be [batch, %d] (dense values only)" % self.dense_in)
            dense = feats
        else:
            if feats.shape[1] != self.slots + self.dense_in:
                raise ValueError("features must be [batch, %d] (ids then dense values)" % (self.slots + self.dense_in))
            ids = feats[:, : self.slots]
            dense = feats[:, self.slots:]
        # one conversion, one comparison, one min / max: at batch 4096 the validation used to cost more than the GPU
        with np.errstate(invalid="ignore"):           # NaN / inf fail the equality test below
            ids64 = ids.astype(np.int64)
        if ids64.size:
            lo, hi = int(ids64.min()), int(ids64.max())
            if ids.dtype.kind == "f":
                if not bool((ids64 == ids).all()):
                    raise ValueError("sparse ids must be integers")
                if ids.dtype == np.float32 and max(-lo, hi) >= _F32_EXACT:
                    raise ValueError("sparse ids >= 2**24 are not exact in float32; send them as integers (input 'ids') or float64")
            if lo < 0 or hi >= self.emb_rows:
                ids64 = np.mod(ids64, self.emb_rows)
        ids = np.ascontiguousarray(ids64.astype(np.int32))
        dense = np.ascontiguousarray(dense, np.float32)
        p = self._predict_gpu(ids, dense) if self.backend == "cuda" else self._predict_cpu(ids, dense)
        return {"keys": keys, "prediction": p.reshape(-1, 1)}
