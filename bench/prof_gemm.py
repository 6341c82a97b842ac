"""Launch one GEMM flavour a few times (for ncu / quick timing).
    python bench/prof_gemm.py fw|dx|dw M N K [bn] [splits]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_tensorflow_b200.ops import gemm as G

op = sys.argv[1]
M, N, K = map(int, sys.argv[2:5])
bn = int(sys.argv[5]) if len(sys.argv) > 5 else 0
splits = int(sys.argv[6]) if len(sys.argv) > 6 else 0
iters = int(os.environ.get("ITERS", "5"))
dev = "cuda"
torch.manual_seed(0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
if op == "fw":
    x = torch.randn(M, K, device=dev).bfloat16(); w = torch.randn(N, K, device=dev).bfloat16(); b = torch.randn(N, device=dev)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    fn = lambda: G.linear_fw(x, w, b, relu=True, out=out, bn=bn)
elif op == "dx":
    dy = torch.randn(M, N, device=dev).bfloat16(); w = torch.randn(N, K, device=dev).bfloat16()
    h = G.pack_mask(torch.randn(M, K, device=dev) > 0); out = torch.empty(M, K, device=dev, dtype=torch.bfloat16)
    fn = lambda: G.linear_dx(dy, w, h, out=out, bn=bn)
elif op == "dxn":
    dy = torch.randn(M, N, device=dev).bfloat16(); w = torch.randn(N, K, device=dev).bfloat16()
    out = torch.empty(M, K, device=dev, dtype=torch.bfloat16)
    fn = lambda: G.linear_dx(dy, w, None, out=out, bn=bn)
else:
    dy = torch.randn(M, N, device=dev).bfloat16(); x = torch.randn(M, K, device=dev).bfloat16()
    out = torch.zeros(N, K, device=dev)
    fn = lambda: G.linear_dw(dy, x, out=out, accumulate=True, splits=splits, bn=bn)
ts = []
for i in range(iters):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
print(op, M, N, K, "bn", bn, "splits", splits, "us(L2-flushed):", ["%.1f" % t for t in ts],
      "TF/s best %.0f" % (2.0 * M * N * K / min(ts) / 1e6))
