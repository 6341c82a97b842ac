"""Dense bf16 GEMMs on the tcgen05 kernel (``csrc/kernels/gemm_sm100.cuh``).

Three layouts cover a linear layer end to end without any transposed copies:

=========  =====================================  ==========================================
function   math                                   operand layouts read by the kernel
=========  =====================================  ==========================================
linear_fw  ``y = act(x @ W.T + b)``               x [M,K] K-major, W [N,K] K-major
linear_dx  ``dx = (dy @ W) * relu'(h)``           dy [M,N] K-major, W [N,K] read MN-major
linear_dw  ``dW (+)= dy.T @ x``  (fp32)           dy [M,N] MN-major, x [M,K] MN-major
=========  =====================================  ==========================================

These stand in for the scalar ``tf.multiply``/``reduce_sum`` graph of the reference
(/root/reference/trainer/task.py:137-139) at wide-and-deep scale.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

EPI_BF16 = 0
EPI_BIAS_BF16 = 1
EPI_BIAS_RELU_BF16 = 2
EPI_DRELU_BF16 = 3
EPI_F32 = 4
EPI_F32_ATOMIC = 5


def _ld(t: torch.Tensor) -> int:
    assert t.dim() == 2 and t.stride(1) == 1, "row-major 2-D tensor expected"
    return t.stride(0)


def mask_words(n_cols: int) -> int:
    """uint32 words per row of a ReLU bit-mask covering ``n_cols`` columns (even, for 8-byte stores)."""
    w = (n_cols + 31) // 32
    return w + (w & 1)


def pack_mask(keep: torch.Tensor) -> torch.Tensor:
    """Pack a boolean [M, N] matrix into the int32 bit-mask layout of the GEMM epilogues (host helper
    for tests; the forward kernel writes this layout itself)."""
    M, N = keep.shape
    W = mask_words(N)
    pad = torch.zeros(M, W * 32, dtype=torch.int64, device=keep.device)
    pad[:, :N] = keep.to(torch.int64)
    weights = (1 << torch.arange(32, device=keep.device, dtype=torch.int64))
    words = (pad.view(M, W, 32) * weights).sum(-1)
    words = torch.where(words >= 2 ** 31, words - 2 ** 32, words)
    return words.to(torch.int32).contiguous()


def unpack_mask(mask: torch.Tensor, n_cols: int) -> torch.Tensor:
    M, W = mask.shape
    bits = (mask.to(torch.int64).unsqueeze(-1) >> torch.arange(32, device=mask.device)) & 1
    return bits.view(M, W * 32)[:, :n_cols].bool()


def gemm_raw(A, a_mn, B, b_mn, D, M, N, K, epilogue, bias=None, mask_in=None, mask_out=None, splits=1, bn=0,
             stream=None):
    L = _lib.lib()
    rc = L.dtf_gemm_bf16(
        _lib.ptr(A), ctypes.c_longlong(_ld(A)), ctypes.c_int(int(a_mn)),
        _lib.ptr(B), ctypes.c_longlong(_ld(B)), ctypes.c_int(int(b_mn)),
        _lib.ptr(D), ctypes.c_longlong(_ld(D)),
        ctypes.c_int(M), ctypes.c_int(N), ctypes.c_int(K), ctypes.c_int(epilogue),
        _lib.ptr(bias), _lib.ptr(mask_in), _lib.ptr(mask_out),
        ctypes.c_longlong(_ld(mask_in) if mask_in is not None else (_ld(mask_out) if mask_out is not None else 0)),
        ctypes.c_int(splits), ctypes.c_int(bn), _lib.stream_ptr(stream))
    _lib.check(rc, "dtf_gemm_bf16")
    return D


def linear_fw(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None, relu: bool = False,
              out: torch.Tensor | None = None, out_dtype=torch.bfloat16, bn: int = 0,
              mask_out: torch.Tensor | None = None) -> torch.Tensor:
    """``act(x @ w.T + bias)``; x [M,K] bf16, w [N,K] bf16, bias [N] fp32.  With ``relu`` and
    ``mask_out`` (int32 [M, mask_words(N)]) the kernel also records which outputs are > 0."""
    M, K = x.shape
    N = w.shape[0]
    assert w.shape[1] == K and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
    if out is None:
        out = torch.empty(M, N, device=x.device, dtype=out_dtype)
    if out.dtype == torch.float32:
        assert bias is None and not relu
        epi = EPI_F32
    elif bias is not None:
        epi = EPI_BIAS_RELU_BF16 if relu else EPI_BIAS_BF16
    else:
        assert not relu
        epi = EPI_BF16
    return gemm_raw(x, 0, w, 0, out, M, N, K, epi, bias=bias, mask_out=mask_out if relu else None, bn=bn)


def linear_dx(dy: torch.Tensor, w: torch.Tensor, mask: torch.Tensor | None = None,
              out: torch.Tensor | None = None, bn: int = 0) -> torch.Tensor:
    """``(dy @ w) * mask``; dy [M,N] bf16, w [N,K] bf16, mask = ReLU bit-mask [M, mask_words(K)] int32
    recorded by the forward of the layer that produced this layer's input (None = no mask)."""
    M, N = dy.shape
    K = w.shape[1]
    assert w.shape[0] == N
    if out is None:
        out = torch.empty(M, K, device=dy.device, dtype=torch.bfloat16)
    epi = EPI_DRELU_BF16 if mask is not None else EPI_BF16
    # GEMM view: D[M, K] = dy[M, N] * (w^T)[K, N]^T ; reduction over N; w is stored [N(red), K] = MN-major
    return gemm_raw(dy, 0, w, 1, out, M, K, N, epi, mask_in=mask, bn=bn)


def linear_dw(dy: torch.Tensor, x: torch.Tensor, out: torch.Tensor | None = None, accumulate: bool = False,
              splits: int = 0, bn: int = 0) -> torch.Tensor:
    """``dW[N,K] (+)= dy.T @ x`` in fp32; dy [M,N] bf16, x [M,K] bf16; reduction over the batch M."""
    M, N = dy.shape
    K = x.shape[1]
    assert x.shape[0] == M
    if out is None:
        out = torch.zeros(N, K, device=dy.device, dtype=torch.float32)
        accumulate = True
    if splits <= 0:
        tiles = ((N + 127) // 128) * ((K + 255) // 256)
        splits = max(1, min((M + 63) // 64, 148 // max(tiles, 1)))
    if splits > 1 and not accumulate:
        out.zero_()
        accumulate = True
    epi = EPI_F32_ATOMIC if accumulate else EPI_F32
    # GEMM view: D[N, K] = (dy^T)[N, M] * (x^T)[K, M]^T ; both stored with the reduction dim as rows
    return gemm_raw(dy, 1, x, 1, out, N, K, M, epi, splits=splits, bn=bn)


def set_pair_mode(on: bool) -> None:
    """CTA-pair tiles (``tcgen05.mma.cta_group::2``: the two SMs of a TPC share one 256 x BN tile) for
    every GEMM launch that has a pair instantiation; off = single-CTA 128 x BN tiles everywhere.
    Process-wide; the initial value comes from ``DTF_PAIR`` (default on)."""
    L = _lib.lib()
    L.dtf_gemm_set_pair_mode.restype = None
    L.dtf_gemm_set_pair_mode(ctypes.c_int(int(bool(on))))


def pair_mode() -> bool:
    return bool(_lib.lib().dtf_gemm_get_pair_mode())
