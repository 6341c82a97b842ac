"""Time line of overlapped async steps: absolute GPU time of every executor mark on both streams."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_b200.models.wide_deep import WideDeepConfig, WideDeepTrainer
torch.cuda.set_stream(torch.cuda.Stream(priority=-1))
cfg = WideDeepConfig(sync=False)
tr = WideDeepTrainer(cfg); tr.init_params()
g = torch.Generator().manual_seed(0)
b = (torch.randint(0, cfg.emb_rows, (cfg.batch, cfg.slots), generator=g, dtype=torch.int32).cuda(),
     torch.randn(cfg.batch, cfg.dense_in, generator=g).cuda(), (torch.rand(cfg.batch, generator=g) < 0.25).float().cuda())
for _ in range(6): tr.step(*b)
tr.join(); torch.cuda.synchronize()
tr.profile(True)
for _ in range(3): tr.step(*b)
tr.join(); torch.cuda.synchronize()
L = tr._L
L.dtf_wd_profile_read_abs.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
buf = C.create_string_buffer(1 << 16)
L.dtf_wd_profile_read_abs(tr._h, buf, len(buf))
rows = [(float(l.rpartition("=")[2]), l.rpartition("=")[0]) for l in buf.value.decode().splitlines()]
for t, n in sorted(rows):
    print("%8.1f  %s" % (t, n))
