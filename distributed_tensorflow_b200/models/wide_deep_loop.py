"""Graph-replayed training loop for the wide-and-deep trainer.

The reference's worker loop feeds one example per ``sess.run`` through ``feed_dict``
(/root/reference/trainer/task.py:232-246): every step pays a host round trip.  Here ``G`` consecutive
steps are captured once into a CUDA graph *together with* the host->device copies of their input
batches (from pinned host buffers, on a second stream so that they run ahead of the compute) and the
device->host copies of the per-step losses; one ``run()`` then costs one graph launch for ``G`` steps.

    loop = GraphedTrainLoop(trainer, host_batches)      # G = len(host_batches) pinned (ids, dense, labels)
    for _ in range(n):
        # (refill host_batches[i] in place here: they are read again by the next run)
        loop.run()
        losses = loop.losses()                          # G mean losses, waits for the replay

The transfers inside the graph are small copy *kernels* that read / write the pinned host buffers
directly (UVA), not copy-engine memcpy nodes: with two or more ranks, graphs that mixed memcpy nodes
with the kernels that spin on peer signals stalled about once per thousand replays
(``bench/stress_overlap.py``); all-kernel graphs do not.

In async mode the ps-side applies of step t overlap the compute of step t+1 inside the graph
(:meth:`WideDeepTrainer.step` with ``overlap=True``), and step t announces step t+1's batch so that its
embedding rows are gathered while step t computes (``prefetch``, the default in async mode); the graph
ends with a join, so replays are self-contained.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch

from ..ops import _lib
from .wide_deep import WideDeepTrainer

HostBatch = Tuple[torch.Tensor, torch.Tensor, torch.Tensor]


class GraphedTrainLoop:
    def __init__(self, trainer: WideDeepTrainer, host_batches: Sequence[HostBatch],
                 stream: Optional[torch.cuda.Stream] = None, warmup: int = 3, prefetch: Optional[bool] = None):
        if not host_batches:
            raise ValueError("GraphedTrainLoop needs at least one host batch")
        cfg = trainer.cfg
        for ids, dense, labels in host_batches:
            if not (ids.is_pinned() and dense.is_pinned() and labels.is_pinned()):
                raise ValueError("host batches must live in pinned memory (tensor.pin_memory())")
            if tuple(ids.shape) != (cfg.batch, cfg.slots) or ids.dtype != torch.int32:
                raise ValueError("ids must be int32 [batch, slots]")
            if tuple(dense.shape) != (cfg.batch, cfg.dense_in) or dense.dtype != torch.float32:
                raise ValueError("dense must be float32 [batch, dense_in]")
            if tuple(labels.shape) != (cfg.batch,) or labels.dtype != torch.float32:
                raise ValueError("labels must be float32 [batch]")
        if prefetch is None:                    # async mode: gather step g+1's rows during step g (trainer.step docs)
            prefetch = not cfg.sync
        self.prefetch = bool(prefetch)
        self.trainer = trainer
        self.host_batches = list(host_batches)
        self.steps_per_run = len(self.host_batches)
        dev = trainer.device
        # high priority: the compute kernels get the SMs first, the overlapped applies fill the rest
        self.stream = stream or torch.cuda.Stream(device=dev, priority=-1)
        self._copy_stream = torch.cuda.Stream(device=dev)
        self._dev = [tuple(torch.empty_like(t, device=dev) for t in b) for b in self.host_batches]
        self._loss_host = torch.zeros(self.steps_per_run).pin_memory()
        self._done = torch.cuda.Event()
        self.h2d_bytes_per_step = sum(t.numel() * t.element_size() for t in self.host_batches[0])
        self.d2h_bytes_per_step = 4
        with torch.cuda.stream(self.stream):
            for i in range(max(1, warmup)):     # lazy initialisation (streams, kernels) before capture
                b = self.host_batches[i % self.steps_per_run]
                for dst, src in zip(self._dev[0], b):
                    dst.copy_(src, non_blocking=True)
                trainer.step(*self._dev[0])
            trainer.join()
            self.stream.synchronize()
            trainer.overlap_reset()
            L = _lib.lib()
            L.dtf_copy_async.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]

            def kcopy(dst: torch.Tensor, src: torch.Tensor, blocks: int, stream: torch.cuda.Stream) -> None:
                _lib.check(L.dtf_copy_async(C.c_void_p(dst.data_ptr()), C.c_void_p(src.data_ptr()),
                                            dst.numel() * dst.element_size(), blocks, C.c_void_p(stream.cuda_stream)),
                           "dtf_copy_async")

            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph, stream=self.stream):
                self._copy_stream.wait_stream(self.stream)
                ready: List[torch.cuda.Event] = []
                with torch.cuda.stream(self._copy_stream):
                    for gi, b in enumerate(self.host_batches):
                        for dst, src in zip(self._dev[gi], b):
                            kcopy(dst, src, 16, self._copy_stream)          # pinned host -> device over PCIe
                        ev = torch.cuda.Event()
                        ev.record(self._copy_stream)
                        ready.append(ev)
                for gi in range(self.steps_per_run):
                    self.stream.wait_event(ready[gi])
                    nxt = None
                    if prefetch and gi + 1 < self.steps_per_run:
                        self.stream.wait_event(ready[gi + 1])       # the prefetch reads the next batch's ids
                        nxt = self._dev[gi + 1][:2]
                    trainer.step(*self._dev[gi], next_batch=nxt)
                    kcopy(self._loss_host[gi:gi + 1], trainer.last_loss_sum(), 1, self.stream)   # device -> pinned host
                trainer.join()
                self.stream.wait_stream(self._copy_stream)
            trainer.overlap_reset()
        self.warmup_steps = max(1, warmup)

    def run(self) -> None:
        """Launch ``steps_per_run`` training steps (asynchronous; see :meth:`losses`)."""
        with torch.cuda.stream(self.stream):
            self.graph.replay()
            self._done.record(self.stream)

    def losses(self) -> List[float]:
        """Mean loss of every step of the last :meth:`run` (waits for it)."""
        self._done.synchronize()
        return (self._loss_host / self.trainer.cfg.batch).tolist()
