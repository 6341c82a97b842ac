"""Remote-gather probe: table on cuda:1, fused pull kernel on cuda:0 (peer access), sequential vs
random row ids, and all-local for comparison.  python bench/p2p_probe.py"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_tensorflow_b200.ops import _lib
from distributed_tensorflow_b200.ops import fused as F

torch.cuda.set_device(0)
L = _lib.lib()
assert L.dtf_symm_enable_peer(C.c_int(1)) == 0, L.dtf_last_error()
M, slots, dc, N = 8192, 26, 64, 256
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
local = torch.empty(rows, 128, device="cuda:0", dtype=torch.bfloat16).normal_(0, 0.05)
remote = torch.empty(rows, 128, device="cuda:1", dtype=torch.bfloat16).normal_(0, 0.05)
torch.cuda.synchronize(1)
dense = torch.randn(M, dc, device="cuda:0").bfloat16()
w = (torch.randn(N, slots * 128 + dc, device="cuda:0") * 0.02).bfloat16()
b = torch.zeros(N, device="cuda:0")


def run(name, table, ids, frac_remote=None):
    addr = table.data_ptr() + ids.to(torch.int64) * 256
    if frac_remote is not None:
        la = local.data_ptr() + ids.to(torch.int64) * 256
        pick = torch.rand(ids.shape, device="cuda:0") < frac_remote
        addr = torch.where(pick, addr, la)
    addr = addr.contiguous()
    ts = []
    for _ in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); H, a = F.pull_gemm(addr, dense, w, b); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ref = table.to("cuda:0")[ids[:4, 0].long()] if frac_remote is None else None
    ok = True if ref is None else torch.equal(a[:4, :128], ref)
    print("%-34s us=%s  (%.0f GB/s gathered)  ok=%s" % (name, ["%.0f" % t for t in ts[1:]], M * slots * 256 / min(ts) / 1e3, ok), flush=True)


rand = torch.randint(0, rows, (M, slots), device="cuda:0")
seq = (torch.arange(M * slots, device="cuda:0") % rows).view(M, slots)
run("local  random rows", local, rand)
run("local  sequential rows", local, seq)
run("remote random rows", remote, rand)
run("remote sequential rows", remote, seq)
run("half remote random (2-rank mix)", remote, rand, frac_remote=0.5)
run("7/8 remote random (8-rank mix)", remote, rand, frac_remote=0.875)
small = torch.randint(0, 20000, (M, slots), device="cuda:0")
run("remote random rows in 5 MB window", remote, small)
