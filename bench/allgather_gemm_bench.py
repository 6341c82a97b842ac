"""Fused dense pull (all-gather + GEMM, K1b) timed with half of the weight rows in another GPU's HBM.

One process, two GPUs (direct peer mapping): layer-1 shape x[8192, 3392] . W[1024, 3392]^T with W row-sharded over
the two GPUs.  Three ways to get act(x W^T + b) on GPU 0, each timed with CUDA events after warm-up, L2 flushed
between launches:

  fused   : dtf_allgather_gemm_bf16 -- every N-tile's TMA loads read their rows of W from the owning GPU
  gather  : copy the peer's shard into a local buffer (torch copy over NVLink), then the plain fused-epilogue GEMM
  local   : W entirely local (no communication) -- the floor

    python bench/allgather_gemm_bench.py            # needs 2 visible GPUs; writes gpurun_out/allgather_gemm.json
"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_tensorflow_b200.ops import _lib
from distributed_tensorflow_b200.ops import fused as F
from distributed_tensorflow_b200.ops import gemm as G


def timed(fn, flush, reps=30):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    if torch.cuda.device_count() < 2:
        print(json.dumps({"unavailable": "needs two GPUs"}))
        return 0
    torch.cuda.set_device(0)
    assert _lib.lib().dtf_symm_enable_peer(C.c_int(1)) == 0
    torch.manual_seed(0)
    out = []
    for M, N, K in ((8192, 1024, 3392), (8192, 512, 1024), (1024, 1024, 3392)):
        rows = N // 2
        x = (torch.randn(M, K, device="cuda:0") * 0.5).bfloat16()
        w0 = (torch.randn(rows, K, device="cuda:0") * 0.05).bfloat16()
        w1 = (torch.randn(rows, K, device="cuda:1") * 0.05).bfloat16()
        torch.cuda.synchronize(1)
        b = torch.randn(N, device="cuda:0") * 0.1
        w_local = torch.empty(N, K, dtype=torch.bfloat16, device="cuda:0")
        w_local[:rows].copy_(w0)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")

        def fused():
            return F.allgather_gemm(x, [w0, (w1.data_ptr(), rows)], b, relu=True)

        def gather_then_gemm():
            w_local[rows:].copy_(w1, non_blocking=True)
            return G.linear_fw(x, w_local, b, relu=True)

        def local():
            return G.linear_fw(x, w_local, b, relu=True)

        y = fused()
        w_local[rows:].copy_(w1)
        ref = torch.relu(x.float() @ w_local.float().t() + b)
        err = ((y.float() - ref).abs().max() / ref.abs().max()).item()
        rec = {"M": M, "N": N, "K": K, "remote_bytes": rows * K * 2, "rel_err": round(err, 5),
               "fused_us": round(timed(fused, flush), 1), "gather_then_gemm_us": round(timed(gather_then_gemm, flush), 1),
               "local_us": round(timed(local, flush), 1)}
        print(rec, flush=True)
        out.append(rec)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/allgather_gemm.json", "w"), indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
