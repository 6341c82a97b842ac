"""Baseline: the same sharded wide-and-deep step the way a PyTorch user who cares about speed would
write it -- NCCL collectives and cuBLAS, none of this repository's kernels, engine or transport.

The north star asks for it because the reference (a TF-1.x script) publishes no numbers and cannot
run here: "always run a PyTorch NCCL(+cuBLAS) DDP/RPC build of the same model".  Same model, shapes,
batch, optimizer and synthetic data as ``bench.py``'s own arm; run with ``bench.py --impl torch``.

What "competent" means here (round 1's version measured Python, not PyTorch):

* no host synchronisation inside a step: the id exchange uses a FIXED capacity per (source, owner) pair
  (mean + 8 sigma of the multinomial split, padded with -1), so ``all_to_all_single`` runs with equal
  splits and nothing calls ``.tolist()`` / ``.item()``; an overflow counter on the device is read once
  after the run;
* every shape is static, so the whole step -- collectives included -- is captured in ONE CUDA graph
  (falls back to eager launches if capture is refused);
* dense parameters: one flat bf16 buffer the ``nn.functional.linear`` calls read views of, gradients
  accumulate in place into one flat bf16 buffer (no autocast casts, no flatten / copy round trip),
  which is all-reduced as it is; fp32 master + Adagrad accumulator updated by a handful of flat
  element-wise kernels;
* sparse update: sort by row + segment sum (no ``unique``), Adagrad on the touched rows;
* ``mode="async"``: the Hogwild counterpart that NCCL allows -- no dense all-reduce, every rank applies
  its own dense gradient to its replica; the sparse path (which needs the all-to-all) is unchanged.

The code is device-agnostic (``gloo`` on CPU tensors works), which is how tests/test_baseline_cpu.py
checks the routing and the update rules against a dense single-process restatement.
"""
from __future__ import annotations

import json
import math
import os
import time

import torch
import torch.distributed as dist
import torch.nn.functional as F


class TorchWideDeep:
    def __init__(self, emb_rows, batch, rank, world, dev, lr=0.05, slots=26, dense_in=13, dense_pad=64,
                 hidden=(1024, 512, 256), emb_dim=128, mode="sync", group=None, compute_dtype=torch.bfloat16, capacity=None):
        self.rank, self.world, self.dev, self.lr, self.mode, self.group = rank, world, dev, lr, mode, group
        self.B, self.F, self.D, self.dense_in, self.dense_pad = batch, slots, emb_dim, dense_in, dense_pad
        self.cdt = compute_dtype
        self.rows_local = (emb_rows - rank + world - 1) // world
        n = batch * slots
        mean = n / world
        self.cap = n if world == 1 else min(n, int(math.ceil(mean + 8.0 * math.sqrt(mean * (1 - 1.0 / world)) + 8)))
        if capacity is not None:
            self.cap = int(capacity)
        self.M = world * self.cap                        # entries this shard receives per step (padded)
        g = torch.Generator(device=dev).manual_seed(1234 + rank)
        # one extra "dummy" row at the end absorbs the padding entries
        self.emb = torch.zeros(self.rows_local + 1, emb_dim, device=dev)
        self.emb[: self.rows_local].uniform_(-0.05, 0.05, generator=g)
        self.emb_acc = torch.full_like(self.emb, 0.1)
        self.wide = torch.zeros(self.rows_local + 1, device=dev)
        self.wide_acc = torch.full_like(self.wide, 0.1)
        k0 = slots * emb_dim + dense_pad
        dims = (k0,) + tuple(hidden) + (1,)
        shapes = []
        for i in range(len(dims) - 1):
            shapes += [(dims[i + 1], dims[i]), (dims[i + 1],)]
        shapes.append((1,))                              # wide bias
        total = sum(math.prod(s) for s in shapes)
        gd = torch.Generator(device="cpu").manual_seed(7)  # identical dense init on every rank
        master = torch.zeros(total)
        o = 0
        for s in shapes:
            if len(s) == 2:
                master[o:o + math.prod(s)] = (torch.rand(math.prod(s), generator=gd) * 2 - 1) / math.sqrt(s[1])
            o += math.prod(s)
        self.master = master.to(dev)                     # fp32 master weights (flat)
        self.acc = torch.full_like(self.master, 0.1)     # Adagrad accumulator (flat)
        self.flat = self.master.to(self.cdt)             # working copy the forward reads (flat)
        self.flat_grad = torch.zeros_like(self.flat)
        self.views, o = [], 0
        for s in shapes:
            v = self.flat[o:o + math.prod(s)].view(s).requires_grad_(True)
            v.grad = self.flat_grad[o:o + math.prod(s)].view(s)   # autograd accumulates into the flat buffer in place
            self.views.append(v)
            o += math.prod(s)
        self.overflow = torch.zeros((), dtype=torch.long, device=dev)
        self._ar = torch.arange(n, device=dev)

    # ---------------------------------------------------------------- collectives (equal splits, static shapes)
    def _a2a(self, x):
        if self.world == 1:
            return x
        out = torch.empty_like(x)
        dist.all_to_all_single(out, x, group=self.group)
        return out

    def _route(self, flat):
        """ids -> (row index at the owner for every received entry [M], bookkeeping to bring rows back)."""
        R, cap = self.world, self.cap
        if R == 1:
            return flat, None                                  # one shard: ids are rows, nothing to route
        owner, local = flat % R, flat // R
        order = torch.argsort(owner, stable=True)
        so = owner[order]
        counts = torch.zeros(R, dtype=torch.long, device=self.dev).scatter_add_(0, owner, torch.ones_like(owner))   # no host sync
        starts = torch.cumsum(counts, 0) - counts
        pos = self._ar - starts[so]
        self.overflow += (pos >= cap).sum()
        slot = torch.where(pos < cap, so * cap + pos, torch.full_like(pos, R * cap))   # R*cap = discard slot
        send = torch.full((R * cap + 1,), -1, dtype=torch.long, device=self.dev)
        send.index_copy_(0, slot, local[order])
        recv = self._a2a(send[: R * cap].contiguous())
        rows = torch.where(recv >= 0, recv, torch.full_like(recv, self.rows_local))      # padding -> dummy row
        return rows, (order, slot)

    def _bring_back(self, x, route):
        """[M, ...] values computed at the owners -> [N, ...] in the order of this rank's ids."""
        if route is None:
            return x
        order, slot = route
        back = self._a2a(x.contiguous())
        back = torch.cat([back, torch.zeros_like(back[:1])], 0)      # index R*cap: entries that overflowed
        out = torch.empty((order.numel(),) + tuple(x.shape[1:]), dtype=x.dtype, device=self.dev)
        out.index_copy_(0, order, back.index_select(0, slot))
        return out

    def _send_to_owners(self, g, route):
        """[N, ...] gradients in id order -> [M, ...] aligned with the entries the owner received."""
        if route is None:
            return g
        order, slot = route
        send = torch.zeros((self.M + 1,) + tuple(g.shape[1:]), dtype=g.dtype, device=self.dev)
        send.index_copy_(0, slot, g.index_select(0, order))
        return self._a2a(send[: self.M].contiguous())

    # ---------------------------------------------------------------- one training step
    def _mlp(self, a):
        v = self.views
        h = a
        n_layers = (len(v) - 1) // 2
        for i in range(n_layers):
            h = F.linear(h, v[2 * i], v[2 * i + 1])
            if i < n_layers - 1:
                h = torch.relu(h)
        return h.float().squeeze(1)

    def _sparse_apply(self, rows, g_rows, g_wide):
        """Adagrad on the touched rows: sort by row, segment-sum duplicates, one update per distinct row.
        Static shapes throughout: unused segments point at the dummy row with a zero gradient."""
        M, dummy = rows.numel(), self.rows_local
        srt, perm = torch.sort(rows)
        first = torch.ones_like(srt, dtype=torch.bool)
        first[1:] = srt[1:] != srt[:-1]
        seg = torch.cumsum(first, 0) - 1
        ge = torch.zeros(M, self.D, device=self.dev).index_add_(0, seg, g_rows.index_select(0, perm).float())
        gw = torch.zeros(M, device=self.dev).index_add_(0, seg, g_wide.index_select(0, perm).float())
        urow = torch.full((M,), dummy, dtype=torch.long, device=self.dev).scatter_(0, seg, srt)
        real = (urow != dummy).float()
        ge, gw = ge * real[:, None], gw * real
        acc = self.emb_acc.index_select(0, urow) + ge * ge
        self.emb_acc.index_copy_(0, urow, acc)
        self.emb.index_add_(0, urow, -self.lr * ge / acc.sqrt())
        wacc = self.wide_acc.index_select(0, urow) + gw * gw
        self.wide_acc.index_copy_(0, urow, wacc)
        self.wide.index_add_(0, urow, -self.lr * gw / wacc.sqrt())

    def _dense_apply(self):
        g = self.flat_grad
        if self.world > 1 and self.mode == "sync":
            dist.all_reduce(g, group=self.group)         # the flat bf16 gradient buffer as it is
        g32 = g.float()
        self.acc.addcmul_(g32, g32)
        self.master.addcdiv_(g32, self.acc.sqrt(), value=-self.lr)
        self.flat.copy_(self.master)
        g.zero_()

    def step(self, ids, dense, labels):
        B, Fs, D = self.B, self.F, self.D
        rows, route = self._route(ids.reshape(-1).long())
        e = self._bring_back(self.emb.index_select(0, rows).to(self.cdt), route).requires_grad_(True)      # [B*F, D]
        w = self._bring_back(self.wide.index_select(0, rows), route).requires_grad_(True)                  # [B*F]
        x = torch.zeros(B, self.dense_pad, dtype=self.cdt, device=self.dev)
        x[:, : self.dense_in] = dense.to(self.cdt)
        deep = self._mlp(torch.cat([e.view(B, Fs * D), x], dim=1))
        logit = deep + w.view(B, Fs).sum(1) + self.views[-1].float()
        denom = B * (self.world if self.mode == "sync" else 1)
        loss = F.binary_cross_entropy_with_logits(logit, labels, reduction="sum")
        (loss / denom).backward()
        with torch.no_grad():
            g_rows = self._send_to_owners(e.grad, route)
            g_wide = self._send_to_owners(w.grad, route)
            self._sparse_apply(rows, g_rows, g_wide)
            self._dense_apply()
        return loss.detach()


class GraphedStep:
    """The whole step (collectives included) as one CUDA graph over static input buffers; eager fallback
    when capture is not possible (CPU, or a library refusing capture)."""

    def __init__(self, model: TorchWideDeep, example, warmup: int = 3, use_graph: bool = True):
        self.model = model
        self.s = tuple(torch.empty_like(t) for t in example)
        self.loss = None
        self.graph = None
        self.first_loss = None
        for _ in range(warmup):                          # allocator warm-up + NCCL connection set-up, outside capture
            self._copy(example)
            self.loss = model.step(*self.s)
            if self.first_loss is None:
                self.first_loss = float(self.loss)
        if use_graph and model.dev.type == "cuda":
            try:
                torch.cuda.synchronize()
                side = torch.cuda.Stream(device=model.dev)
                side.wait_stream(torch.cuda.current_stream(model.dev))
                with torch.cuda.stream(side):
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph, stream=side):
                        self.loss = model.step(*self.s)
                torch.cuda.current_stream(model.dev).wait_stream(side)
                torch.cuda.synchronize()
                self.graph = graph
            except Exception as e:  # noqa: BLE001
                self.graph = None
                self.why_eager = "%s: %s" % (type(e).__name__, str(e)[:200])
                torch.cuda.synchronize()

    def _copy(self, batch):
        for d, s in zip(self.s, batch):
            d.copy_(s, non_blocking=True)

    def __call__(self, batch):
        self._copy(batch)
        if self.graph is not None:
            self.graph.replay()
        else:
            self.loss = self.model.step(*self.s)
        return self.loss


def synthetic_pool(args, rank, pin: bool):
    """Same synthetic data rule as bench.py's own arm (learnable labels)."""
    g = torch.Generator().manual_seed(1000 + rank)
    pool = []
    for _ in range(args.pool):
        ids = torch.randint(0, args.emb_rows, (args.batch, 26), generator=g, dtype=torch.int32)
        dense = torch.randn(args.batch, 13, generator=g)
        logit = dense[:, 0] - 0.5 * dense[:, 1] + (ids[:, 0] % 2).float() - 0.5
        labels = (torch.rand(args.batch, generator=g) < torch.sigmoid(logit)).float()
        pool.append(tuple(t.pin_memory() if pin else t for t in (ids, dense, labels)))
    return pool


def time_baseline(args, rank, world, dev, group=None, mode="sync", steps=None, use_graph=True) -> dict:
    """Device-timed ms/step of the baseline on the calling ranks (collective).  ``group``: an NCCL process
    group (None = the default group).  Used by ``run_baseline`` and by bench.py's own arm for its
    same-box ratio."""
    steps = steps or args.steps
    model = TorchWideDeep(args.emb_rows, args.batch, rank, world, dev, mode=mode, group=group)
    pool_h = synthetic_pool(args, rank, pin=dev.type == "cuda")
    pool_d = [tuple(t.to(dev) for t in b) for b in pool_h]
    step = GraphedStep(model, pool_d[0], warmup=max(3, args.warmup), use_graph=use_graph)
    first_loss = step.first_loss / args.batch

    def barrier():
        if world > 1:
            dist.barrier(group=group)
    sync = torch.cuda.synchronize if dev.type == "cuda" else (lambda: None)
    sync()
    barrier()
    step(pool_d[0])                              # aligns the ranks (its collectives) before the start event
    if dev.type == "cuda":
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            step(pool_d[i % args.pool])
        e1.record()
        sync()
        dev_ms = e0.elapsed_time(e1)
    else:
        t0 = time.perf_counter()
        for i in range(steps):
            step(pool_d[i % args.pool])
        dev_ms = (time.perf_counter() - t0) * 1e3
    barrier()
    # end to end: pinned host -> device copy of every step's inputs, loss read back every step
    t0 = time.perf_counter()
    last = 0.0
    for i in range(steps):
        b = tuple(t.to(dev, non_blocking=True) for t in pool_h[i % args.pool])
        last = float(step(b)) / args.batch           # float(): device -> host read of the loss
    sync()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    ovf = model.overflow.clone()
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        dist.all_reduce(ovf, group=group)
    dev_ms, e2e_ms = t.tolist()
    return {"ms_per_step": dev_ms / steps, "e2e_ms_per_step": e2e_ms / steps, "steps": steps, "mode": mode,
            "cuda_graph": step.graph is not None, "why_eager": getattr(step, "why_eager", None),
            "first_loss": first_loss, "last_loss": last, "dropped_ids": int(ovf), "capacity_per_pair": model.cap,
            "h2d_bytes": sum(x.numel() * x.element_size() for x in pool_h[0])}


def run_baseline(args, rank, world, dev, ClockSampler) -> int:
    if world > 1 and dist.is_initialized():
        dist.destroy_process_group()                     # bench.py opened a gloo group; the baseline wants NCCL
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    sampler = ClockSampler(dev.index) if rank == 0 else None
    r = time_baseline(args, rank, world, dev, mode=args.mode)
    clocks = sampler.stop() if sampler else None
    gb = args.batch * world
    if rank == 0:
        print(json.dumps({
            "metric": "train examples/sec (device-timed, max over ranks)", "value": gb / (r["ms_per_step"] / 1e3),
            "unit": "examples/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "impl": "torch-nccl-cublas",
            "config": {"model": "wide_and_deep 10M x128 embedding + MLP 3392-1024-512-256-1", "global_batch": gb,
                       "per_gpu_batch": args.batch,
                       "parallelism": "row-sharded embedding via fixed-capacity NCCL all_to_all + flat bf16 all-reduce DP MLP",
                       "mode": args.mode, "optimizer": "adagrad", "emb_rows": args.emb_rows, "cuda_graph": r["cuda_graph"],
                       "why_eager": r["why_eager"], "capacity_per_pair": r["capacity_per_pair"], "dropped_ids": r["dropped_ids"]},
            "e2e": {"value": gb / (r["e2e_ms_per_step"] / 1e3), "unit": "examples/s", "ms_per_step": r["e2e_ms_per_step"],
                    "h2d_bytes_per_step": r["h2d_bytes"], "d2h_bytes_per_step": 4, "first_loss": r["first_loss"],
                    "last_loss": r["last_loss"]},
            "gpu_launches": 0, "clocks": clocks}))
    if world > 1:
        dist.destroy_process_group()
    return 0
