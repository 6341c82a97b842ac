"""Fused pull and fused push with 7 of 8 rows in ANOTHER GPU's HBM, one process (so that ncu may wrap it and count
NVLink bytes: multi-rank commands are never run under ncu).  Mimics what one rank of an 8-rank job does per step:
54.5 MB of embedding rows pulled (47.7 MB over NVLink), 54.5 MB of gradient rows pushed.

    python bench/prof_pull_peer.py [remote_eighths=7]          # needs two visible GPUs
"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_tensorflow_b200.ops import _lib
from distributed_tensorflow_b200.ops import fused as F

remote8 = int(sys.argv[1]) if len(sys.argv) > 1 else 7
iters = int(os.environ.get("ITERS", "4"))
M, slots, dc, N, rows = 8192, 26, 64, 1024, 1_250_000
torch.cuda.set_device(0)
assert _lib.lib().dtf_symm_enable_peer(C.c_int(1)) == 0
torch.manual_seed(0)
t_local = torch.empty(rows, 128, device="cuda:0", dtype=torch.bfloat16).normal_(0, 0.05)
t_peer = torch.empty(rows, 128, device="cuda:1", dtype=torch.bfloat16).normal_(0, 0.05)
mb_local = torch.zeros(M * slots, 128, device="cuda:0", dtype=torch.bfloat16)        # mailboxes (push destinations)
mb_peer = torch.zeros(M * slots, 128, device="cuda:1", dtype=torch.bfloat16)
torch.cuda.synchronize(1)
dense = torch.randn(M, dc, device="cuda:0").bfloat16()
w = (torch.randn(N, slots * 128 + dc, device="cuda:0") * 0.02).bfloat16()
b = torch.zeros(N, device="cuda:0")
dz = (torch.randn(M, N, device="cuda:0") * 0.1).bfloat16()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
tp, tq = [], []
for i in range(iters):
    ids = torch.randint(0, rows, (M, slots), device="cuda:0")
    owner_remote = (torch.randint(0, 8, (M, slots), device="cuda:0") < remote8)
    src = torch.where(owner_remote, t_peer.data_ptr() + ids.to(torch.int64) * 256, t_local.data_ptr() + ids.to(torch.int64) * 256).contiguous()
    entry = torch.arange(M * slots, device="cuda:0", dtype=torch.int64).view(M, slots)
    dst = torch.where(owner_remote, mb_peer.data_ptr() + entry * 256, mb_local.data_ptr() + entry * 256).contiguous()
    flush.zero_()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record()
    H, a_buf = F.pull_gemm(src, dense, w, b, relu=True)
    e1.record()
    F.push_gemm(dz, w, dst)
    e2.record()
    torch.cuda.synchronize()
    tp.append(e0.elapsed_time(e1) * 1e3)
    tq.append(e1.elapsed_time(e2) * 1e3)
ref = torch.relu(torch.cat([torch.where(owner_remote[:64, :, None], t_peer.to("cuda:0")[ids[:64]], t_local[ids[:64]]).reshape(64, -1),
                            dense[:64]], 1).float() @ w.float().t() + b)
err = ((H[:64].float() - ref).abs().max() / ref.abs().max()).item()
print("remote %d/8: fused pull us %s | fused push us %s | pull rel err %.4f" % (remote8, ["%.0f" % t for t in tp], ["%.0f" % t for t in tq], err))
