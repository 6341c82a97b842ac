"""tcgen05 GEMM family vs plain PyTorch fp32: forward / data-gradient / weight-gradient layouts,
fused epilogues (bias, ReLU + bit-mask, masked dgrad, split-K TMA reduce-add) and the fused
pull (gather + GEMM) and push (GEMM + row scatter) kernels."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-9)).item()


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (300, 200, 136), (1000, 512, 1024), (77, 64, 3392)])
@pytest.mark.parametrize("bn", [0, 64, 128, 256])
def test_forward_bias_relu_and_mask(M, N, K, bn):
    from distributed_tensorflow_b200.ops import gemm as G
    torch.manual_seed(0)
    x = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    b = torch.randn(N, device="cuda")
    mask = torch.zeros(M, G.mask_words(N), dtype=torch.int32, device="cuda")
    y = G.linear_fw(x, w, b, relu=True, bn=bn, mask_out=mask)
    ref = torch.relu(x.float() @ w.float().t() + b)
    assert _rel(y, ref) < 1e-2
    assert torch.equal(G.unpack_mask(mask, N), y > 0)
    y2 = G.linear_fw(x, w, b, relu=False, bn=bn)
    assert _rel(y2, x.float() @ w.float().t() + b) < 1e-2
    y3 = G.linear_fw(x, w, out_dtype=torch.float32, bn=bn)
    assert _rel(y3, x.float() @ w.float().t()) < 1e-5


@pytest.mark.parametrize("M,N,K", [(128, 256, 128), (300, 200, 136), (1000, 512, 1024)])
def test_data_gradient_with_relu_mask(M, N, K):
    from distributed_tensorflow_b200.ops import gemm as G
    torch.manual_seed(1)
    dy = (torch.randn(M, N, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    keep = torch.rand(M, K, device="cuda") > 0.5
    for bn in (0, 128, 256):
        dx = G.linear_dx(dy, w, G.pack_mask(keep), bn=bn)
        assert _rel(dx, (dy.float() @ w.float()) * keep) < 1e-2
        assert _rel(G.linear_dx(dy, w, None, bn=bn), dy.float() @ w.float()) < 1e-2


@pytest.mark.parametrize("M,N,K", [(128, 128, 256), (256, 200, 136), (2048, 512, 1024)])
@pytest.mark.parametrize("splits", [1, 0, 7])
def test_weight_gradient_split_k(M, N, K, splits):
    from distributed_tensorflow_b200.ops import gemm as G
    torch.manual_seed(2)
    dy = (torch.randn(M, N, device="cuda") * 0.5).bfloat16()
    x = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    ref = dy.float().t() @ x.float()
    assert _rel(G.linear_dw(dy, x, splits=splits), ref) < 1e-4
    acc = torch.ones(N, K, device="cuda")
    G.linear_dw(dy, x, out=acc, accumulate=True, splits=splits)          # accumulates into existing values
    assert _rel(acc, ref + 1) < 1e-4


def test_argument_errors_are_reported():
    from distributed_tensorflow_b200.ops import _lib
    from distributed_tensorflow_b200.ops import gemm as G
    x = torch.zeros(128, 72, device="cuda", dtype=torch.bfloat16)[:, :68]      # row stride not 16-byte aligned... ok
    w = torch.zeros(64, 68, device="cuda", dtype=torch.bfloat16)
    with pytest.raises(_lib.KernelError, match="16-byte"):
        G.linear_fw(x.contiguous()[:, :67], w[:, :67])
    with pytest.raises(_lib.KernelError, match="splits"):
        G.gemm_raw(w, 1, w, 1, torch.zeros(68, 68, device="cuda"), 68, 68, 64, G.EPI_F32, splits=4)


@pytest.mark.parametrize("M,slots,dense_cols,N", [(300, 4, 64, 128), (1024, 26, 64, 1024), (128, 2, 0, 256)])
def test_fused_pull_gemm(M, slots, dense_cols, N):
    from distributed_tensorflow_b200.ops import fused as F
    from distributed_tensorflow_b200.ops import gemm as G
    torch.manual_seed(3)
    table = (torch.randn(5000, 128, device="cuda") * 0.1).bfloat16()
    ids = torch.randint(0, 5000, (M, slots), device="cuda")
    dense = torch.randn(M, dense_cols, device="cuda").bfloat16()
    K0 = slots * 128 + dense_cols
    w = (torch.randn(N, K0, device="cuda") * 0.03).bfloat16()
    b = torch.randn(N, device="cuda") * 0.1
    mask = torch.zeros(M, G.mask_words(N), dtype=torch.int32, device="cuda")
    H, a_buf = F.pull_gemm(F.row_addresses(table, ids), dense, w, b, relu=True, mask_out=mask)
    a_ref = torch.cat([table[ids].reshape(M, slots * 128), dense], dim=1)
    assert torch.equal(a_buf, a_ref)
    ref = torch.relu(a_ref.float() @ w.float().t() + b)
    assert _rel(H, ref) < 1e-2
    assert torch.equal(G.unpack_mask(mask, N), H > 0)


@pytest.mark.parametrize("M,slots,K", [(300, 4, 128), (1024, 26, 1024)])
def test_fused_push_gemm(M, slots, K):
    from distributed_tensorflow_b200.ops import fused as F
    torch.manual_seed(4)
    dy = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(K, slots * 128 + 64, device="cuda") * 0.05).bfloat16()
    mailbox = torch.zeros(M * slots + 3, 128, device="cuda", dtype=torch.bfloat16)
    perm = torch.randperm(M * slots, device="cuda").view(M, slots)
    dst = (mailbox.data_ptr() + perm * 256).contiguous()
    dst[5, 1] = 0                                                       # address 0 = this row is not pushed
    F.push_gemm(dy, w, dst)
    torch.cuda.synchronize()
    ref = (dy.float() @ w.float()[:, : slots * 128]).view(M, slots, 128)
    got = mailbox[perm.view(-1)].view(M, slots, 128).float()
    ref[5, 1] = 0
    assert _rel(got, ref) < 1e-2
    assert mailbox[M * slots:].abs().max() == 0


@pytest.mark.parametrize("M,rows,shards,K", [(300, 64, 3, 136), (2048, 256, 4, 1024), (1024, 128, 8, 512)])
def test_allgather_gemm_reads_row_shards_in_place(M, rows, shards, K):
    from distributed_tensorflow_b200.ops import fused as F
    torch.manual_seed(5)
    x = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    ws = [(torch.randn(rows, K, device="cuda") * 0.05).bfloat16() for _ in range(shards)]
    b = torch.randn(rows * shards, device="cuda") * 0.1
    y = F.allgather_gemm(x, ws, b, relu=True)
    ref = torch.relu(x.float() @ torch.cat(ws).float().t() + b)
    assert _rel(y, ref) < 1e-2


def test_allgather_gemm_over_peer_memory():
    """Half of the weight rows live in another GPU's HBM (direct peer mapping)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import ctypes as C
    from distributed_tensorflow_b200.ops import _lib
    from distributed_tensorflow_b200.ops import fused as F
    torch.cuda.set_device(0)
    assert _lib.lib().dtf_symm_enable_peer(C.c_int(1)) == 0
    torch.manual_seed(6)
    M, rows, K = 1024, 256, 512
    x = (torch.randn(M, K, device="cuda:0") * 0.5).bfloat16()
    w0 = (torch.randn(rows, K, device="cuda:0") * 0.05).bfloat16()
    w1 = (torch.randn(rows, K, device="cuda:1") * 0.05).bfloat16()
    torch.cuda.synchronize(1)
    b = torch.zeros(2 * rows, device="cuda:0")
    y = F.allgather_gemm(x, [w0, (w1.data_ptr(), rows)], b, relu=False)
    ref = x.float() @ torch.cat([w0, w1.to("cuda:0")]).float().t()
    assert _rel(y, ref) < 1e-2


def test_cta_pair_tiles_match_single_cta_tiles():
    """Every layout / epilogue with a cta_group::2 instantiation: same numbers as the single-CTA
    kernels, including an odd number of 128-row tiles (last pair half empty) and ragged N / K."""
    from distributed_tensorflow_b200.ops import fused as F
    from distributed_tensorflow_b200.ops import gemm as G
    torch.manual_seed(5)
    before = G.pair_mode()
    try:
        for M, N, K in [(256, 256, 128), (640, 512, 200), (1100, 1024, 1024), (2048, 296, 520)]:
            x = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
            w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
            b = torch.randn(N, device="cuda")
            dy = (torch.randn(M, N, device="cuda") * 0.5).bfloat16()
            keep = torch.rand(M, K, device="cuda") > 0.5
            got = {}
            for mode in (False, True):
                G.set_pair_mode(mode)
                assert G.pair_mode() == mode
                m = torch.zeros(M, G.mask_words(N), dtype=torch.int32, device="cuda")
                got[mode] = (G.linear_fw(x, w, b, relu=True, bn=256, mask_out=m), m,
                             G.linear_fw(x, w, None, out_dtype=torch.float32, bn=256),
                             G.linear_dx(dy, w, G.pack_mask(keep), bn=256),
                             G.linear_dw(dy, x, splits=0), G.linear_dw(dy, x, splits=3))
            for a, c in zip(got[False], got[True]):
                if a.dtype == torch.int32:
                    assert torch.equal(a, c)
                else:
                    assert _rel(c, a) < 2e-3
            assert _rel(got[True][0], torch.relu(x.float() @ w.float().t() + b)) < 1e-2
            assert _rel(got[True][4], dy.float().t() @ x.float()) < 1e-4
        # fused pull and push on pairs
        M, slots, N = 1000, 6, 512
        table = (torch.randn(5000, 128, device="cuda") * 0.1).bfloat16()
        ids = torch.randint(0, 5000, (M, slots), device="cuda")
        dense = torch.randn(M, 64, device="cuda").bfloat16()
        w = (torch.randn(N, slots * 128 + 64, device="cuda") * 0.05).bfloat16()
        b = torch.randn(N, device="cuda")
        outs = {}
        for mode in (False, True):
            G.set_pair_mode(mode)
            H, a = F.pull_gemm(F.row_addresses(table, ids), dense, w, b, relu=True)
            outs[mode] = (H.clone(), a.clone())
        assert torch.equal(outs[True][1], outs[False][1])
        assert _rel(outs[True][0], outs[False][0]) < 2e-3
        ref_a = torch.cat([table[ids.view(-1)].view(M, slots * 128), dense], dim=1)
        assert torch.equal(outs[True][1], ref_a)
    finally:
        G.set_pair_mode(before)
