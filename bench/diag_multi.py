"""Per-phase device timing of the wide-and-deep step on N ranks (torchrun) or 1 rank:
compute | barrier+publish | apply | barrier, CUDA events, median over steps, max over ranks."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from distributed_tensorflow_b200.models.wide_deep import WideDeepConfig, WideDeepTrainer

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); lr = int(os.environ.get("LOCAL_RANK", "0"))
mode = sys.argv[1] if len(sys.argv) > 1 else "sync"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 12
back_to_back = len(sys.argv) > 3 and sys.argv[3] == "b2b"
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
if world > 1:
    dist.init_process_group("gloo", rank=rank, world_size=world)
cfg = WideDeepConfig(sync=(mode == "sync"))
tr = WideDeepTrainer(cfg, rank=rank, world=world, device=dev)
if world > 1:
    tr.attach_peers(dist)
tr.init_params()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
g = torch.Generator().manual_seed(5 + rank)
batches = [(torch.randint(0, cfg.emb_rows, (cfg.batch, cfg.slots), generator=g, dtype=torch.int32).to(dev),
            torch.randn(cfg.batch, cfg.dense_in, generator=g).to(dev), (torch.rand(cfg.batch, generator=g) < 0.25).float().to(dev))
           for _ in range(4)]
names = ["compute", "barrier1", "apply", "barrier2"]
rec = []
for s in range(steps):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    b = batches[s % 4]
    ev[0].record(); tr.step_compute(*b)
    ev[1].record(); tr.barrier(True)
    ev[2].record(); tr.step_apply()
    ev[3].record()
    if world > 1:
        tr.barrier(False)
    ev[4].record()
    if not back_to_back:
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
    rec.append(ev)
torch.cuda.synchronize()
ts = torch.tensor([[e[i].elapsed_time(e[i + 1]) for i in range(4)] for e in rec[3:]], dtype=torch.float64)
med = ts.median(0).values
tot = torch.tensor([rec[3][0].elapsed_time(rec[-1][4]) / (len(rec) - 3)], dtype=torch.float64)
if world > 1:
    dist.all_reduce(med, op=dist.ReduceOp.MAX); dist.all_reduce(tot, op=dist.ReduceOp.MAX)
if rank == 0:
    print("world", world, mode, "b2b" if back_to_back else "synced", " ".join("%s=%.3fms" % (n, v) for n, v in zip(names, med.tolist())),
          "sum=%.3fms" % med.sum().item(), "wall_per_step=%.3fms" % tot.item(), flush=True)
# per-kernel breakdown of a few more steps (CUDA events between launches)
tr.profile(True)
agg = {}
order = []
for s in range(4):
    b = batches[s % 4]
    tr.step_compute(*b); tr.barrier(True); tr.step_apply()
    if world > 1:
        tr.barrier(False)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    for k, v in tr.profile_read():
        if k in ("step", "begin"):
            continue
        if k not in agg:
            agg[k] = []
            order.append(k)
        agg[k].append(v)
tr.profile(False)
if rank == 0:
    print("world %d %s per-kernel us (median of 4): " % (world, mode) + "  ".join("%s=%.0f" % (k, sorted(agg[k])[len(agg[k]) // 2]) for k in order), flush=True)
tr.close()
if world > 1:
    dist.destroy_process_group()
