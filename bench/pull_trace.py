"""Device-side timeline of the fused pull kernel (gather warps + TMA producers), local table and,
with 2 GPUs, a table on cuda:1 read over NVLink.  Uses the dtf_debug_set_pull_trace hook: every
gather warp stamps %globaltimer at the start of an iteration (2 slices = 8 KB), after its loads are
consumed / stores issued, and after the release; every producer stamps the start / end of its wait
for a row-tile.   python bench/pull_trace.py [N]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_tensorflow_b200.ops import _lib
from distributed_tensorflow_b200.ops import fused as F

torch.cuda.set_device(0)
L = _lib.lib()
L.dtf_debug_set_pull_trace.restype = None
L.dtf_debug_set_pull_trace.argtypes = [C.c_void_p]
two = torch.cuda.device_count() > 1
if two:
    assert L.dtf_symm_enable_peer(C.c_int(1)) == 0, L.dtf_last_error()
M, slots, dc = 8192, 26, 64
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
rows = 5_000_000
dense = torch.randn(M, dc, device="cuda:0").bfloat16()
w = (torch.randn(N, slots * 128 + dc, device="cuda:0") * 0.02).bfloat16()
b = torch.zeros(N, device="cuda:0")
trace = torch.zeros(72000, dtype=torch.int64, device="cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")


def pct(t, q):
    return float(torch.quantile(t.double(), q))


def run(name, table):
    ids = torch.randint(0, rows, (M, slots), device="cuda:0")
    addr = (table.data_ptr() + ids.to(torch.int64) * 256).contiguous()
    for it in range(3):
        flush.fill_(it)
        trace.zero_()
        L.dtf_debug_set_pull_trace(C.c_void_p(trace.data_ptr() if it == 2 else 0))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); F.pull_gemm(addr, dense, w, b); e1.record(); torch.cuda.synchronize()
    L.dtf_debug_set_pull_trace(C.c_void_p(0))
    t = trace.cpu()
    g = t[:148 * 12 * 8 * 4].view(148 * 12, 8, 4)
    used = g[:, :, 0] > 0
    t00 = int(g[:, :, 0][used].min())
    start = (g[:, :, 0] - t00)[used] / 1e3
    load = (g[:, :, 1] - g[:, :, 0])[used] / 1e3
    rel = (g[:, :, 2] - g[:, :, 1])[used] / 1e3
    end = (g[:, :, 2] - t00)[used] / 1e3
    iters = used.sum(1)
    print("%s: kernel %.1f us (with trace stores)" % (name, e0.elapsed_time(e1) * 1e3))
    print("  gather slices per warp: min %d max %d; first retire p50 %.1f us, last end max %.1f us" %
          (int(iters.min()), int(iters.max()), pct(start, 0.5), float(end.max())))
    print("  wait for loads + issue stores                       us: p10 %.2f p50 %.2f p90 %.2f max %.2f" %
          (pct(load, 0.1), pct(load, 0.5), pct(load, 0.9), float(load.max())))
    print("  release phase (syncwarp + red.release.gpu)          us: p10 %.2f p50 %.2f p90 %.2f max %.2f" %
          (pct(rel, 0.1), pct(rel, 0.5), pct(rel, 0.9), float(rel.max())))
    for k in range(int(iters.max())):
        u = used[:, k]
        print("    iteration %d: start p50 %.1f  end p50 %.1f  end max %.1f" %
              (k, pct((g[:, k, 0] - t00)[u] / 1e3, 0.5), pct((g[:, k, 2] - t00)[u] / 1e3, 0.5), float(((g[:, k, 2] - t00)[u] / 1e3).max())))
    pr = t[65536:65536 + 148 * 8 * 4].view(148, 8, 4)
    pu = pr[:, :, 1] > 0
    wend = (pr[:, :, 1] - t00)[pu] / 1e3
    wstart = (pr[:, :, 0] - t00)[pu] / 1e3
    mt = pr[:, :, 3][pu]
    order = torch.argsort(mt * 1000000 + wend.long())
    print("  producer waits (row-tile: wait start -> end us): " +
          " ".join("%d:%.0f->%.0f" % (int(mt[i]), float(wstart[i]), float(wend[i])) for i in order[:: max(1, len(order) // 24)]))
    print("  last producer wait ends at %.1f us" % float(wend.max()), flush=True)
    mma = (pr[:, :, 2] - t00)[pu] / 1e3
    ep = (t[70272:70272 + 148 * 8].view(148, 8) - t00)[pu] / 1e3
    for k in range(int(pu.sum(1).max())):
        u = pu[:, k]
        print("    GEMM wave %d: MMAs done p10 %.1f p50 %.1f max %.1f | epilogue issued p50 %.1f max %.1f" % (
            k, pct(((pr[:, k, 2] - t00)[u]) / 1e3, 0.1), pct(((pr[:, k, 2] - t00)[u]) / 1e3, 0.5),
            float(((pr[:, k, 2] - t00)[u]).max()) / 1e3,
            pct((t[70272:70272 + 148 * 8].view(148, 8)[:, k] - t00)[u] / 1e3, 0.5),
            float((t[70272:70272 + 148 * 8].view(148, 8)[:, k] - t00)[u].max()) / 1e3), flush=True)


local = torch.empty(rows, 128, device="cuda:0", dtype=torch.bfloat16).normal_(0, 0.05)
run("local table", local)
if two:
    remote = torch.empty(rows, 128, device="cuda:1", dtype=torch.bfloat16).normal_(0, 0.05)
    torch.cuda.synchronize(1)
    run("remote table (NVLink)", remote)
