"""Fused communication + compute ops (``csrc/kernels/gemm_sm100.cuh`` gather warps / push epilogue).

* :func:`pull_gemm` -- embedding rows are fetched through per-(row, slot) device addresses (local HBM
  or a peer's HBM mapped over NVLink) by gather warps of the same persistent kernel whose tcgen05
  warps run the first GEMM on row-tiles that already arrived.
* :func:`push_gemm` -- the data-gradient GEMM whose epilogue scatters every 128-wide output row to a
  per-(row, slot) destination address, e.g. a mailbox in the owning rank's HBM (row K3).

In the reference these are the ps->worker variable reads and worker->ps gradient sends around
``sess.run(train_op)`` (/root/reference/trainer/task.py:236).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from .gemm import mask_words


def row_addresses(table: torch.Tensor, ids: torch.Tensor) -> torch.Tensor:
    """int64 device addresses of ``table[ids]`` rows (table: [rows, 128] bf16, contiguous)."""
    assert table.dtype == torch.bfloat16 and table.shape[1] == 128 and table.is_contiguous()
    return (table.data_ptr() + ids.to(torch.int64) * 256).contiguous()


def pull_gemm_scratch(M: int, device) -> torch.Tensor:
    L = _lib.lib()
    L.dtf_pull_gemm_scratch_bytes.restype = C.c_size_t
    return torch.empty(int(L.dtf_pull_gemm_scratch_bytes(C.c_int(M))), dtype=torch.uint8, device=device)


def pull_gemm(row_src: torch.Tensor, dense_x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, relu: bool = True,
              mask_out: torch.Tensor = None, scratch: torch.Tensor = None):
    """Returns ``(H, a_buf)``; row_src int64 [M, slots], dense_x bf16 [M, dense_cols], w bf16 [N, K0]."""
    M, slots = row_src.shape
    dense_cols = dense_x.shape[1]
    N, K0 = w.shape
    assert K0 == slots * 128 + dense_cols and row_src.dtype == torch.int64 and row_src.is_contiguous()
    dev = w.device
    a_buf = torch.empty(M, K0, dtype=torch.bfloat16, device=dev)
    H = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    if scratch is None:
        scratch = pull_gemm_scratch(M, dev)
    L = _lib.lib()
    rc = L.dtf_pull_gemm_bf16(_lib.ptr(row_src), _lib.ptr(dense_x), C.c_int(slots), C.c_int(dense_cols), _lib.ptr(a_buf),
                              _lib.ptr(w), C.c_longlong(w.stride(0)), _lib.ptr(bias), _lib.ptr(H),
                              C.c_longlong(H.stride(0)), _lib.ptr(mask_out),
                              C.c_longlong(mask_out.stride(0) if mask_out is not None else mask_words(N)),
                              C.c_int(M), C.c_int(N), C.c_int(int(relu)), _lib.ptr(scratch), _lib.stream_ptr())
    _lib.check(rc, "dtf_pull_gemm_bf16")
    return H, a_buf


def push_gemm(dy: torch.Tensor, w: torch.Tensor, row_dst: torch.Tensor) -> None:
    """``(dy @ w[:, :slots*128])`` with row (m, slot) stored as bf16 at address ``row_dst[m, slot]``."""
    M, K = dy.shape
    slots = row_dst.shape[1]
    assert w.shape[0] == K and w.shape[1] >= slots * 128 and row_dst.dtype == torch.int64 and row_dst.is_contiguous()
    L = _lib.lib()
    rc = L.dtf_push_gemm_bf16(_lib.ptr(dy), C.c_longlong(dy.stride(0)), _lib.ptr(w), C.c_longlong(w.stride(0)),
                              _lib.ptr(row_dst), C.c_int(slots), C.c_int(M), C.c_int(K), _lib.stream_ptr())
    _lib.check(rc, "dtf_push_gemm_bf16")


def allgather_gemm(x: torch.Tensor, w_shards, bias: torch.Tensor, relu: bool = True, mask_out: torch.Tensor = None):
    """``act(x @ cat(w_shards).T + bias)`` without materialising the concatenation: every N-tile's TMA
    loads read their rows of W from the shard that owns them (shards may live in peers' HBM, mapped
    over NVLink).  ``w_shards``: list of bf16 [rows_per_shard, K] tensors or raw device addresses."""
    M, K = x.shape
    ptrs, rows = [], None
    for s in w_shards:
        if isinstance(s, torch.Tensor):
            assert s.dtype == torch.bfloat16 and s.shape[1] == K and s.is_contiguous()
            rows = s.shape[0] if rows is None else rows
            assert s.shape[0] == rows
            ptrs.append(s.data_ptr())
        else:
            ptrs.append(int(s[0]))
            rows = int(s[1]) if rows is None else rows
    N = rows * len(ptrs)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=x.device)
    scratch = torch.empty(128 * len(ptrs) + 64, dtype=torch.uint8, device=x.device)
    off = (-scratch.data_ptr()) % 64
    arr = (C.c_void_p * len(ptrs))(*ptrs)
    L = _lib.lib()
    rc = L.dtf_allgather_gemm_bf16(_lib.ptr(x), C.c_longlong(x.stride(0)), arr, C.c_int(len(ptrs)), C.c_int(rows),
                                   C.c_longlong(K), _lib.ptr(out), C.c_longlong(out.stride(0)), C.c_int(M), C.c_int(N),
                                   C.c_int(K), _lib.ptr(bias), C.c_int(int(relu)), _lib.ptr(mask_out),
                                   C.c_longlong(mask_out.stride(0) if mask_out is not None else mask_words(N)),
                                   C.c_void_p(scratch.data_ptr() + off), _lib.stream_ptr())
    _lib.check(rc, "dtf_allgather_gemm_bf16")
    return out
