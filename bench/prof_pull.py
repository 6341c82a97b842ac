"""Time / profile the standalone fused pull (gather + GEMM-1):  python bench/prof_pull.py [N] [rows]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_tensorflow_b200.ops import fused as F
from distributed_tensorflow_b200.ops import gemm as G

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
rows = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
M, slots, dc = 8192, 26, 64
iters = int(os.environ.get("ITERS", "6"))
torch.manual_seed(0)
table = torch.empty(rows, 128, device="cuda", dtype=torch.bfloat16).normal_(0, 0.05)
dense = torch.randn(M, dc, device="cuda").bfloat16()
w = (torch.randn(N, slots * 128 + dc, device="cuda") * 0.02).bfloat16()
b = torch.zeros(N, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ts, tg = [], []
for i in range(iters):
    ids = torch.randint(0, rows, (M, slots), device="cuda")
    src = F.row_addresses(table, ids)
    flush.zero_()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record()
    H, a_buf = F.pull_gemm(src, dense, w, b, relu=True)
    e1.record()
    G.linear_fw(a_buf, w, b, relu=True)
    e2.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
    tg.append(e1.elapsed_time(e2) * 1e3)
print("pull_gemm N=%d rows=%d: fused us %s | plain gemm on gathered operand us %s" % (
    N, rows, ["%.0f" % t for t in ts], ["%.0f" % t for t in tg]))
