"""NVLS check (torchrun, >= 2 ranks): trains a few sync-replica steps and prints whether the multicast
path of the dense apply is active, the per-step losses and a checksum of this rank's dense shard.
Run once as is and once with DTF_MULTICAST=0: losses and checksums must agree to fp32 rounding."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from distributed_tensorflow_b200.models.wide_deep import WideDeepConfig, WideDeepTrainer

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); lr = int(os.environ.get("LOCAL_RANK", "0"))
mode = sys.argv[1] if len(sys.argv) > 1 else "sync"
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("gloo", rank=rank, world_size=world)
cfg = WideDeepConfig(emb_rows=1_000_000, sync=(mode == "sync"), optimizer="adam", lr=0.001)
tr = WideDeepTrainer(cfg, rank=rank, world=world, device=dev)
tr.attach_peers(dist)
tr.init_params()
torch.cuda.synchronize()
dist.barrier()
g = torch.Generator().manual_seed(17 + rank)
losses = []
for s in range(8):
    ids = torch.randint(0, cfg.emb_rows, (cfg.batch, cfg.slots), generator=g, dtype=torch.int32).to(dev)
    dense = torch.randn(cfg.batch, cfg.dense_in, generator=g).to(dev)
    labels = (torch.rand(cfg.batch, generator=g) < 0.3).float().to(dev)
    tr.step(ids, dense, labels)
    losses.append(float(tr.last_loss_sum()) / cfg.batch)
tr.join()
torch.cuda.synchronize()
dist.barrier()
shard = tr.named_shards()
chk = {k: float(v.double().abs().sum()) for k, v in shard.items() if v.dtype == torch.float32 and v.numel() < 10_000_000}
print("rank %d multicast=%s mode=%s losses=%s" % (rank, tr.multicast, mode, ["%.6f" % l for l in losses]), flush=True)
print("rank %d checksums %s" % (rank, {k: "%.6f" % v for k, v in sorted(chk.items())[:6]}), flush=True)
dist.barrier()
tr.close()
