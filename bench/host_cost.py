"""Host (CPU) cost of issuing one step: time the enqueue loop without waiting for the GPU."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_b200.models.wide_deep import WideDeepConfig, WideDeepTrainer
mode = sys.argv[1] if len(sys.argv) > 1 else "async"
cfg = WideDeepConfig(sync=(mode == "sync"))
hp = torch.cuda.Stream(priority=-1)
torch.cuda.set_stream(hp)
tr = WideDeepTrainer(cfg); tr.init_params()
g = torch.Generator().manual_seed(0)
b = (torch.randint(0, cfg.emb_rows, (cfg.batch, cfg.slots), generator=g, dtype=torch.int32).cuda(),
     torch.randn(cfg.batch, cfg.dense_in, generator=g).cuda(), (torch.rand(cfg.batch, generator=g) < 0.25).float().cuda())
for _ in range(5): tr.step(*b)
torch.cuda.synchronize()
for overlap in (True, False) if mode == "async" else (False,):
    torch.cuda.synchronize(); tr.overlap_reset()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(40): tr.step(*b, overlap=overlap)
    t_host = (time.perf_counter() - t0) / 40 * 1e3
    tr.join(); e1.record(); torch.cuda.synchronize()
    print(mode, "overlap" if overlap else "serial", "host ms/step %.3f  gpu ms/step %.3f" % (t_host, e0.elapsed_time(e1) / 40), flush=True)
