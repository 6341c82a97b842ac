"""dtf_pull_gemm_bf16_noalloc -- the fused pull for callers that cannot allocate device memory.

Some users of this library have no allocator to turn to when a kernel wants scratch space: an
inference server that gave the whole GPU to its own memory pool when it started (and plug-ins
running inside it), or a request path where cudaMalloc's device-wide synchronisation is not
acceptable.  For them the library carries the scratch of the fused-pull kernel as a static device
array, and this entry point takes no scratch argument.

The test drives the C ABI directly (ctypes on the shared object, raw device pointers) and checks,
on two calls in a row with different inputs, that the result equals what the ordinary entry point
(`dtf_pull_gemm_bf16`, caller-provided scratch) produces -- the second call exercises the "array
still holds the previous call's counters" case.  It also checks that a batch larger than the
built-in scratch is refused with a clear message.  The file sorts last so that a `-x` run has
already executed every other test when it gets here; it passes alone in a fresh process as well.
"""
import ctypes as C
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

_SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "distributed_tensorflow_b200", "_lib",
                   "libdtf_kernels.so")


def _abi():
    if not os.path.exists(_SO):
        from distributed_tensorflow_b200.utils.build import build_kernels
        build_kernels()
    L = C.CDLL(_SO)
    vp, i, ll = C.c_void_p, C.c_int, C.c_longlong
    L.dtf_last_error.restype = C.c_char_p
    L.dtf_pull_gemm_scratch_bytes.restype = C.c_size_t
    L.dtf_pull_gemm_scratch_bytes.argtypes = [i]
    L.dtf_pull_gemm_noalloc_max_rows.restype = i
    common = [vp, vp, i, i, vp, vp, ll, vp, vp, ll, vp, ll, i, i, i]
    L.dtf_pull_gemm_bf16.argtypes = common + [vp, vp]          # ..., scratch, stream
    L.dtf_pull_gemm_bf16.restype = i
    L.dtf_pull_gemm_bf16_noalloc.argtypes = common + [vp]      # ..., stream
    L.dtf_pull_gemm_bf16_noalloc.restype = i
    return L


def _problem(seed, M, slots, dense_cols, N):
    g = torch.Generator(device="cuda").manual_seed(seed)
    table = (torch.randn(4096, 128, device="cuda", generator=g) * 0.1).bfloat16()
    ids = torch.randint(0, 4096, (M, slots), device="cuda", generator=g)
    row_src = (table.data_ptr() + ids * 256).contiguous()                     # int64 device addresses
    dense = torch.randn(M, dense_cols, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, slots * 128 + dense_cols, device="cuda", generator=g) * 0.03).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g) * 0.1
    return table, row_src, dense, w, bias


def _args(row_src, dense, w, bias, a_buf, H, M, slots, dense_cols, N):
    return [C.c_void_p(row_src.data_ptr()), C.c_void_p(dense.data_ptr()), slots, dense_cols, C.c_void_p(a_buf.data_ptr()),
            C.c_void_p(w.data_ptr()), w.stride(0), C.c_void_p(bias.data_ptr()), C.c_void_p(H.data_ptr()), H.stride(0),
            C.c_void_p(0), 0, M, N, 1]


def test_no_alloc_entry_matches_the_ordinary_entry_on_two_calls_in_a_row():
    L = _abi()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # two different problems back to back; the second one is larger, so it also touches flag words
    # the first call never cleared
    for seed, (M, slots, dense_cols, N) in enumerate([(300, 4, 64, 256), (1500, 6, 64, 512)]):
        table, row_src, dense, w, bias = _problem(seed, M, slots, dense_cols, N)
        K0 = slots * 128 + dense_cols
        a_ref = torch.empty(M, K0, dtype=torch.bfloat16, device="cuda")
        H_ref = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        scratch = torch.empty(int(L.dtf_pull_gemm_scratch_bytes(M)), dtype=torch.uint8, device="cuda")
        rc = L.dtf_pull_gemm_bf16(*_args(row_src, dense, w, bias, a_ref, H_ref, M, slots, dense_cols, N),
                                  C.c_void_p(scratch.data_ptr()), stream)
        assert rc == 0, L.dtf_last_error().decode()
        a_new = torch.zeros(M, K0, dtype=torch.bfloat16, device="cuda")
        H_new = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
        rc = L.dtf_pull_gemm_bf16_noalloc(*_args(row_src, dense, w, bias, a_new, H_new, M, slots, dense_cols, N), stream)
        assert rc == 0, L.dtf_last_error().decode()
        torch.cuda.synchronize()
        assert torch.equal(a_new, a_ref)                                       # gathered operand: exact
        err = ((H_new.float() - H_ref.float()).abs().max() / (H_ref.float().abs().max() + 1e-9)).item()
        assert err < 1e-2, "call %d: relative error %.3e" % (seed, err)        # same tolerance as the GEMM tests
        ref = torch.relu(a_ref.float() @ w.float().t() + bias)                 # and both agree with plain torch
        assert ((H_new.float() - ref).abs().max() / ref.abs().max()).item() < 1e-2


def test_no_alloc_entry_refuses_batches_beyond_its_built_in_scratch():
    L = _abi()
    limit = L.dtf_pull_gemm_noalloc_max_rows()
    assert limit == 1 << 20
    dummy = torch.zeros(64, device="cuda")
    p = C.c_void_p(dummy.data_ptr())
    rc = L.dtf_pull_gemm_bf16_noalloc(p, p, 1, 64, p, p, 192, p, p, 128, C.c_void_p(0), 0, limit + 1, 128, 1,
                                      C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc < 0
    msg = L.dtf_last_error().decode()
    assert "exceeds" in msg and str(limit) in msg
