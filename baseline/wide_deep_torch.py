"""Baseline: the same sharded wide-and-deep step written the stock PyTorch way -- NCCL collectives
(`all_to_all_single` for the row-sharded embedding lookup and its gradient, `all_reduce` for the
dense gradients), cuBLAS GEMMs under bf16 autocast, torch index ops + `torch.optim.Adagrad` for the
updates.  None of this repository's kernels, engine or transport is on this path.

The north star asks for it because the reference (a TF-1.x script) publishes no numbers and cannot
run here: "always run a PyTorch NCCL(+cuBLAS) DDP/RPC build of the same model".  Same model, shapes,
batch, optimizer and synthetic data as `bench.py`'s own arm; run with `bench.py --impl torch`.
"""
from __future__ import annotations

import json
import os
import time

import torch
import torch.distributed as dist
import torch.nn as nn


class TorchWideDeep:
    def __init__(self, emb_rows, batch, rank, world, dev, lr=0.05, slots=26, dense_in=13, dense_pad=64,
                 hidden=(1024, 512, 256), emb_dim=128):
        self.rank, self.world, self.dev, self.lr = rank, world, dev, lr
        self.B, self.F, self.D, self.dense_in, self.dense_pad = batch, slots, emb_dim, dense_in, dense_pad
        rows_local = (emb_rows - rank + world - 1) // world
        g = torch.Generator(device=dev).manual_seed(1234 + rank)
        self.emb = torch.empty(rows_local, emb_dim, device=dev).uniform_(-0.05, 0.05, generator=g)
        self.emb_acc = torch.full_like(self.emb, 0.1)
        self.wide = torch.zeros(rows_local, device=dev)
        self.wide_acc = torch.full_like(self.wide, 0.1)
        k0 = slots * emb_dim + dense_pad
        dims = (k0,) + tuple(hidden)
        layers = []
        for i in range(len(hidden)):
            layers += [nn.Linear(dims[i], dims[i + 1]), nn.ReLU()]
        layers += [nn.Linear(hidden[-1], 1)]
        torch.manual_seed(7)                       # identical dense init on every rank
        self.mlp = nn.Sequential(*layers).to(dev)
        self.wide_b = torch.zeros(1, device=dev, requires_grad=True)
        self.params = list(self.mlp.parameters()) + [self.wide_b]
        self.opt = torch.optim.Adagrad(self.params, lr=lr, initial_accumulator_value=0.1)
        self.flat_grad = None

    # row-sharded lookup: ids -> owners (NCCL all-to-all) -> rows back (NCCL all-to-all)
    def _exchange_ids(self, flat):
        R = self.world
        owner = flat % R
        local = flat // R
        if R == 1:
            return local, None
        order = torch.argsort(owner, stable=True)
        send_counts = torch.bincount(owner, minlength=R)
        recv_counts = torch.empty_like(send_counts)
        dist.all_to_all_single(recv_counts, send_counts)
        sc, rc = send_counts.tolist(), recv_counts.tolist()
        recv_ids = torch.empty(sum(rc), dtype=local.dtype, device=self.dev)
        dist.all_to_all_single(recv_ids, local[order].contiguous(), rc, sc)
        return recv_ids, (order, sc, rc)

    def _return_rows(self, rows, route):
        if route is None:
            return rows
        order, sc, rc = route
        back = torch.empty(sum(sc), rows.shape[1], dtype=rows.dtype, device=self.dev)
        dist.all_to_all_single(back, rows.contiguous(), sc, rc)
        out = torch.empty_like(back)
        out[order] = back
        return out

    def _send_grads(self, grads, route):
        if route is None:
            return grads
        order, sc, rc = route
        recv = torch.empty(sum(rc), grads.shape[1], dtype=grads.dtype, device=self.dev)
        dist.all_to_all_single(recv, grads[order].contiguous(), rc, sc)
        return recv

    def step(self, ids, dense, labels):
        B, F, D = self.B, self.F, self.D
        flat = ids.reshape(-1).long()
        owner_rows, route = self._exchange_ids(flat)
        rows_bf16 = self.emb.index_select(0, owner_rows).to(torch.bfloat16)
        wide_rows = self.wide.index_select(0, owner_rows).unsqueeze(1)
        e = self._return_rows(rows_bf16, route).float().requires_grad_(True)        # [B*F, D]
        w = self._return_rows(wide_rows, route).float().requires_grad_(True)        # [B*F, 1]
        x = torch.zeros(B, self.dense_pad, device=self.dev)
        x[:, : self.dense_in] = dense
        with torch.autocast("cuda", dtype=torch.bfloat16):
            deep = self.mlp(torch.cat([e.view(B, F * D), x], dim=1)).float().squeeze(1)
        logit = deep + w.view(B, F).sum(1) + self.wide_b
        loss = nn.functional.binary_cross_entropy_with_logits(logit, labels, reduction="sum") / (B * self.world)
        self.opt.zero_grad(set_to_none=True)
        loss.backward()
        # dense: DDP-style gradient all-reduce, then the stock optimizer
        if self.world > 1:
            flat_g = torch.cat([p.grad.reshape(-1) for p in self.params])
            dist.all_reduce(flat_g)
            o = 0
            for p in self.params:
                p.grad.copy_(flat_g[o:o + p.numel()].view_as(p))
                o += p.numel()
        self.opt.step()
        # sparse: gradients travel to the owning shard, duplicates are summed, Adagrad on touched rows
        g_rows = self._send_grads(e.grad.to(torch.bfloat16), route).float()
        g_wide = self._send_grads(w.grad, route).squeeze(1)
        uniq, inv = torch.unique(owner_rows, return_inverse=True)
        ge = torch.zeros(uniq.numel(), D, device=self.dev).index_add_(0, inv, g_rows)
        gw = torch.zeros(uniq.numel(), device=self.dev).index_add_(0, inv, g_wide)
        acc = self.emb_acc.index_select(0, uniq) + ge * ge
        self.emb_acc.index_copy_(0, uniq, acc)
        self.emb.index_add_(0, uniq, -self.lr * ge / acc.sqrt())
        wacc = self.wide_acc.index_select(0, uniq) + gw * gw
        self.wide_acc.index_copy_(0, uniq, wacc)
        self.wide.index_add_(0, uniq, -self.lr * gw / wacc.sqrt())
        return loss.detach() * self.world


def run_baseline(args, rank, world, dev, ClockSampler) -> int:
    if world > 1 and dist.is_initialized():
        dist.destroy_process_group()                     # bench.py opened a gloo group; the baseline wants NCCL
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    model = TorchWideDeep(args.emb_rows, args.batch, rank, world, dev)
    g = torch.Generator().manual_seed(1000 + rank)
    pool_h = []
    for _ in range(args.pool):
        ids = torch.randint(0, args.emb_rows, (args.batch, 26), generator=g, dtype=torch.int32).pin_memory()
        dense = torch.randn(args.batch, 13, generator=g).pin_memory()
        labels = (torch.rand(args.batch, generator=g) < 0.25).float().pin_memory()
        pool_h.append((ids, dense, labels))
    pool_d = [tuple(t.to(dev) for t in b) for b in pool_h]
    h2d_bytes = sum(t.numel() * t.element_size() for t in pool_h[0])

    def barrier():
        if world > 1:
            dist.barrier()
    warm = max(3, args.warmup)
    for i in range(warm):
        model.step(*pool_d[i % args.pool])
    torch.cuda.synchronize()
    sampler = ClockSampler(dev.index) if rank == 0 else None
    barrier()
    model.step(*pool_d[0])                       # aligns the ranks (its collectives) before the start event
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        model.step(*pool_d[i % args.pool])
    e1.record()
    torch.cuda.synchronize()
    barrier()
    dev_ms = e0.elapsed_time(e1)
    t0 = time.perf_counter()
    last = 0.0
    for i in range(args.steps):
        b = tuple(t.to(dev, non_blocking=True) for t in pool_h[i % args.pool])
        last = float(model.step(*b))                      # .item(): device -> host read of the loss
    torch.cuda.synchronize()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = t.tolist()
    gb = args.batch * world
    if rank == 0:
        print(json.dumps({
            "metric": "train examples/sec (device-timed, max over ranks)", "value": gb * args.steps / (dev_ms / 1e3),
            "unit": "examples/s", "n_gpus": world, "steps": args.steps, "warmup": warm, "ms_per_step": dev_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "impl": "torch-nccl-cublas",
            "config": {"model": "wide_and_deep 10M x128 embedding + MLP 3392-1024-512-256-1", "global_batch": gb,
                       "per_gpu_batch": args.batch, "parallelism": "row-sharded embedding via NCCL all_to_all + all-reduce DP MLP",
                       "mode": "sync", "optimizer": "adagrad", "emb_rows": args.emb_rows},
            "e2e": {"value": gb * args.steps / (e2e_ms / 1e3), "unit": "examples/s", "ms_per_step": e2e_ms / args.steps,
                    "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": 4, "last_loss": last},
            "gpu_launches": 0, "clocks": clocks}))
    if world > 1:
        dist.destroy_process_group()
    return 0
