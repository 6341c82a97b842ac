"""IPC-mapped remote gather probe (torchrun, 2 ranks): each rank allocates a table with the symmetric
allocator, maps the peer's; then (a) only rank 1 gathers from rank 0's table, (b) both gather from
each other at the same time."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from distributed_tensorflow_b200.ops import _lib
from distributed_tensorflow_b200.ops import fused as F

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("gloo", rank=rank, world_size=world)
L = _lib.lib()
L.dtf_symm_handle_size.restype = C.c_int
rows = 5_000_000
nbytes = rows * 256
ptr = C.c_void_p()
assert L.dtf_symm_alloc(C.c_size_t(nbytes), C.byref(ptr)) == 0
hbuf = C.create_string_buffer(L.dtf_symm_handle_size())
assert L.dtf_symm_get_handle(ptr, hbuf) == 0
hs = [None] * world
dist.all_gather_object(hs, hbuf.raw)
peer = C.c_void_p()
assert L.dtf_symm_open_handle(hs[1 - rank], C.byref(peer)) == 0, L.dtf_last_error()
dist.barrier()
M, slots, dc, N = 8192, 26, 64, 256
dense = torch.randn(M, dc, device="cuda").bfloat16()
w = (torch.randn(N, slots * 128 + dc, device="cuda") * 0.02).bfloat16()
b = torch.zeros(N, device="cuda")
ids = torch.randint(0, rows, (M, slots), device="cuda")
remote_addr = (peer.value + ids.to(torch.int64) * 256).contiguous()
local_addr = (ptr.value + ids.to(torch.int64) * 256).contiguous()
mix_addr = torch.where(torch.rand(ids.shape, device="cuda") < 0.5, remote_addr, local_addr).contiguous()


def timed(addr):
    ts = []
    for _ in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); F.pull_gemm(addr, dense, w, b); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return ts[1:]


for name, addr in (("local", local_addr), ("remote (IPC)", remote_addr), ("half remote (IPC)", mix_addr)):
    dist.barrier()
    if rank == 1:
        print("only rank 1 active, %-18s us=%s" % (name, ["%.0f" % t for t in timed(addr)]), flush=True)
    dist.barrier()
for name, addr in (("remote (IPC)", remote_addr), ("half remote (IPC)", mix_addr)):
    dist.barrier()
    t = timed(addr)
    print("both ranks active, rank %d %-18s us=%s" % (rank, name, ["%.0f" % x for x in t]), flush=True)
    dist.barrier()
dist.destroy_process_group()
