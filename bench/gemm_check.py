"""Numerics + timing sweep of the tcgen05 GEMM against torch (cuBLAS) -- run on a B200:
    python bench/gemm_check.py [--quick]
Writes gpurun_out/gemm_check.json."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_tensorflow_b200.ops import gemm as G


def timeit(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us


def rel_err(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-9)).item()


def main():
    quick = "--quick" in sys.argv
    torch.manual_seed(0)
    dev = "cuda"
    res = []
    shapes_fw = [(128, 256, 64), (128, 256, 256), (256, 512, 128), (300, 200, 136), (8192, 1024, 3392),
                 (8192, 512, 1024), (8192, 256, 512), (8192, 8192, 8192)]
    if quick:
        shapes_fw = shapes_fw[:4]
    for (M, N, K) in shapes_fw:
        x = (torch.randn(M, K, device=dev) * 0.5).bfloat16()
        w = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
        b = torch.randn(N, device=dev)
        ref = torch.relu(x.float() @ w.float().t() + b)
        for bn in (0, 64, 128, 256):
            mo = torch.zeros(M, G.mask_words(N), dtype=torch.int32, device=dev)
            y = G.linear_fw(x, w, b, relu=True, bn=bn, mask_out=mo)
            torch.cuda.synchronize()
            e = rel_err(y, ref)
            if not torch.equal(G.unpack_mask(mo, N), y > 0):
                e = 1.0  # mask mismatch
            row = {"op": "fw", "M": M, "N": N, "K": K, "bn": bn, "rel_err": e}
            if bn == 0 or M >= 8192:
                t = timeit(lambda: G.linear_fw(x, w, b, relu=True, bn=bn))
                tc = timeit(lambda: torch.relu(torch.nn.functional.linear(x, w, b.bfloat16())))
                tm = timeit(lambda: x @ w.t())
                row.update(us=t, torch_us=tc, matmul_us=tm, tflops=2.0 * M * N * K / t / 1e6)
            res.append(row)
            print(row, flush=True)
    # dgrad: dx = (dy @ W) * mask
    for (M, N, K) in [(128, 256, 128), (300, 200, 136), (8192, 1024, 3392), (8192, 512, 1024), (8192, 256, 512)][: (2 if quick else 99)]:
        dy = (torch.randn(M, N, device=dev) * 0.5).bfloat16()
        w = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
        h = torch.relu(torch.randn(M, K, device=dev)).bfloat16()
        h = G.pack_mask(h > 0)
        ref = (dy.float() @ w.float()) * G.unpack_mask(h, K)
        for bn in (0, 128, 256):
            dx = G.linear_dx(dy, w, h, bn=bn)
            torch.cuda.synchronize()
            row = {"op": "dx", "M": M, "N": N, "K": K, "bn": bn, "rel_err": rel_err(dx, ref)}
            if bn == 0:
                t = timeit(lambda: G.linear_dx(dy, w, h))
                tm = timeit(lambda: dy @ w)
                row.update(us=t, matmul_us=tm, tflops=2.0 * M * N * K / t / 1e6)
            res.append(row)
            print(row, flush=True)
    # wgrad: dW = dy^T x
    for (M, N, K) in [(128, 128, 256), (256, 200, 136), (8192, 1024, 3392), (8192, 512, 1024), (8192, 256, 512)][: (2 if quick else 99)]:
        dy = (torch.randn(M, N, device=dev) * 0.5).bfloat16()
        x = (torch.randn(M, K, device=dev) * 0.5).bfloat16()
        ref = dy.float().t() @ x.float()
        for splits in (1, 0):
            dw = G.linear_dw(dy, x, splits=splits)
            torch.cuda.synchronize()
            row = {"op": "dw", "M": M, "N": N, "K": K, "splits": splits, "rel_err": rel_err(dw, ref)}
            out = torch.zeros(N, K, device=dev)
            t = timeit(lambda: G.linear_dw(dy, x, out=out, accumulate=True, splits=splits))
            tm = timeit(lambda: dy.t() @ x)
            row.update(us=t, matmul_us=tm, tflops=2.0 * M * N * K / t / 1e6)
            res.append(row)
            print(row, flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/gemm_check.json", "w") as f:
        json.dump(res, f, indent=1)
    bad = [r for r in res if not (r["rel_err"] < 2e-2)]
    print("BAD:", bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
