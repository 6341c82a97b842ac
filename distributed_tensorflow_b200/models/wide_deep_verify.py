"""Multi-rank correctness check on the REAL ranks of a running job.

The unit tests emulate ranks on one device with host-ordered phases; what only exists across real
processes -- slabs attached through passed file descriptors, device barriers on the signal pads, the
NVLS ``multimem`` path of the dense apply, the SM-split push next to wgrad-1 -- is exercised here:
a small sync-replica configuration (product layer sizes, 50 000 rows, 512 examples per rank) takes a
few steps on every rank, the global parameters are gathered before and after, and rank 0 replays the
same global batches through :class:`TorchReference` (plain fp32 PyTorch).  With a multicast mapping
attached the last step runs with it detached again, so both forms of the dense reduce-scatter /
all-gather are held against the same oracle (what ``bench/mc_probe.py`` printed, as an assertion).

``bench.py`` runs this at every N and puts the result into its JSON line (``"verify"``);
``tests/test_wide_deep_io_gpu.py`` runs it under torchrun when two GPUs are visible.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List

import numpy as np
import torch

from .wide_deep import TorchReference, WideDeepConfig, WideDeepTrainer

# same bounds as tests/test_wide_deep_gpu.py::test_two_emulated_ranks_sync_equals_one_rank_on_the_global_batch
_TOL = {"embedding": (2e-2, 2e-4), "wide": (2e-2, 2e-4), "dense": (2e-2, 3e-4)}


def verify_batch(cfg: WideDeepConfig, step: int, rank: int):
    """The micro-batch of ``rank`` at ``step``: a function of (step, rank) only, so every rank -- and the
    oracle on rank 0 -- can rebuild any rank's batch without communication."""
    g = torch.Generator(device="cpu").manual_seed(5000 + 131 * step + rank)
    ids = torch.randint(0, cfg.emb_rows, (cfg.batch, cfg.slots), generator=g, dtype=torch.int32)
    ids[:, 0] = ids[0, 0]                                   # one row shared by every example: duplicates across ranks
    dense = torch.randn(cfg.batch, cfg.dense_in, generator=g)
    labels = (torch.rand(cfg.batch, generator=g) < torch.sigmoid(dense[:, 0])).float()
    return ids, dense, labels


def _local_state(tr: WideDeepTrainer) -> Dict[str, np.ndarray]:
    c, n = tr.cfg, tr.info.rows_local
    torch.cuda.synchronize(tr.device)
    return {"emb": tr.buffer("emb_master", torch.float32)[: n * c.emb_dim].view(n, c.emb_dim).cpu().numpy(),
            "wide": tr.buffer("wide", torch.float32)[:n].cpu().numpy(),
            "dense": tr.buffer("dense_master", torch.float32).cpu().numpy(),
            "lo": tr.info.dense_lo, "hi": tr.info.dense_hi,
            "wsum": int(tr.buffer("dense_bf16", torch.int16).long().sum().item())}


def _assemble(states: List[dict], tr: WideDeepTrainer, dev):
    """Global fp32 parameters from the ranks' shards (row r on rank r % world, dense ranges by rank)."""
    c, inf, world = tr.cfg, tr.info, len(states)
    emb = torch.zeros(c.emb_rows, c.emb_dim, device=dev)
    wide = torch.zeros(c.emb_rows, device=dev)
    flat = torch.zeros(inf.P_pad, device=dev)
    for r, s in enumerate(states):
        emb[r::world] = torch.from_numpy(s["emb"]).to(dev)
        wide[r::world] = torch.from_numpy(s["wide"]).to(dev)
        flat[s["lo"]: s["hi"]] = torch.from_numpy(s["dense"]).to(dev)
    cut = lambda off, n: flat[off: off + n].clone()
    dense = {"w1": cut(inf.off_w1, c.h1 * inf.K0).view(c.h1, inf.K0), "w2": cut(inf.off_w2, c.h2 * c.h1).view(c.h2, c.h1),
             "w3": cut(inf.off_w3, c.h3 * c.h2).view(c.h3, c.h2), "b1": cut(inf.off_b1, c.h1), "b2": cut(inf.off_b2, c.h2),
             "b3": cut(inf.off_b3, c.h3), "w4": cut(inf.off_w4, c.h3), "b4": cut(inf.off_b4, 1),
             "wide_b": cut(inf.off_wide_b, 1)}
    return emb, wide, dense


def _ratio(a: torch.Tensor, b: torch.Tensor, rtol: float, atol: float) -> float:
    """max |a - b| as a fraction of the allowed error (<= 1 passes)."""
    err = (a.float() - b.float()).abs().max().item()
    return err / (atol + rtol * b.float().abs().max().item())


def verify_multi_rank(rank: int, world: int, dev: torch.device, dist=None, steps: int = 2, batch: int = 512,
                      emb_rows: int = 50000) -> dict:
    """Collective over all ranks (``dist``: torch.distributed or a group with its call shapes; None for one
    rank).  Returns the same dict on every rank: ``{"ok", "max_rel_err", "steps", "multicast", "detail"}``."""
    cfg = WideDeepConfig(emb_rows=emb_rows, batch=batch, optimizer="adagrad", lr=0.05, sync=True)
    tr = WideDeepTrainer(cfg, rank=rank, world=world, device=dev)

    def gather(obj):
        if world == 1:
            return [obj]
        out = [None] * world
        dist.all_gather_object(out, obj)
        return out

    try:
        if world > 1:
            tr.attach_peers(dist)
        tr.init_params()
        before = gather(_local_state(tr))
        total_steps = steps + (1 if tr.multicast else 0)
        for s in range(total_steps):
            if s == steps:                                  # last step: same trainer, NVLS path switched off
                tr._L.dtf_wd_attach_multicast.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong]
                tr._L.dtf_wd_attach_multicast(tr._h, None, 0)
            ids, dense, labels = (t.to(dev) for t in verify_batch(cfg, s, rank))
            tr.step(ids, dense, labels)
            torch.cuda.synchronize(dev)
            tr.check_error()
        after = gather(_local_state(tr))
        result = None
        if rank == 0:
            emb0, wide0, dense0 = _assemble(before, tr, dev)
            ref = TorchReference(cfg, emb0, wide0, dense0)
            for s in range(total_steps):
                ref.sync_step([tuple(t.to(dev) for t in verify_batch(cfg, s, r)) for r in range(world)])
            emb1, wide1, dense1 = _assemble(after, tr, dev)
            detail = {"embedding": _ratio(emb1, ref.emb, *_TOL["embedding"]), "wide": _ratio(wide1, ref.wide, *_TOL["wide"])}
            for k in dense1:
                detail["dense_" + k] = _ratio(dense1[k], ref.p[k], *_TOL["dense"])
            moved = float((emb1 - emb0).abs().max().item())
            same_copies = len({s["wsum"] for s in after}) == 1      # every rank holds the same dense working copy
            worst = max(detail.values())
            result = {"ok": bool(worst <= 1.0 and moved > 0 and same_copies), "max_rel_err": round(worst, 4),
                      "steps": total_steps, "multicast": bool(tr.multicast), "ranks": world,
                      "dense_copies_identical": same_copies,
                      "detail": {k: round(v, 4) for k, v in detail.items()}}
        if world > 1:
            box = [result]
            dist.broadcast_object_list(box, src=0)
            result = box[0]
        return result
    finally:
        tr.close()
