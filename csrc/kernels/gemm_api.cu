// C-ABI entry points for the dense GEMM family.
#include "../../include/dtf_kernels.h"
#include "gemm_launch.cuh"

using namespace dtf;

extern "C" int dtf_gemm_bf16(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn,
                             void* D, long long ldd, int M, int N, int K, int epilogue, const float* bias,
                             const void* mask_in, void* mask_out, long long ldmask, int splits, int bn,
                             void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) {
    set_error("dtf_gemm_bf16: empty problem %dx%dx%d", M, N, K);
    return -2;
  }
  if (splits > 1 && epilogue != EPI_F32_ATOMIC) {
    set_error("dtf_gemm_bf16: splits=%d needs DTF_EPI_F32_ATOMIC", splits);
    return -2;
  }
  if ((epilogue == EPI_BIAS_BF16 || epilogue == EPI_BIAS_RELU_BF16) && bias == nullptr) {
    set_error("dtf_gemm_bf16: bias epilogue without bias pointer");
    return -2;
  }
  if (epilogue == EPI_DRELU_BF16 && mask_in == nullptr) {
    set_error("dtf_gemm_bf16: DRELU epilogue without mask pointer");
    return -2;
  }
  if ((mask_in || mask_out) && ((ldmask & 1) || ldmask * 32 < N)) {
    set_error("dtf_gemm_bf16: ldmask=%lld must be even and cover N=%d columns", ldmask, N);
    return -2;
  }
  if (epilogue < 0 || epilogue > EPI_F32_ATOMIC) {
    set_error("dtf_gemm_bf16: unknown epilogue %d", epilogue);
    return -2;
  }
  GemmArgs a{};
  a.A = A; a.lda = lda; a.a_mn = a_mn != 0;
  a.B = B; a.ldb = ldb; a.b_mn = b_mn != 0;
  a.D = D; a.ldd = ldd;
  a.M = M; a.N = N; a.K = K;
  a.epi = epilogue; a.bias = bias;
  a.mask_in = reinterpret_cast<const uint32_t*>(mask_in);
  a.mask_out = reinterpret_cast<uint32_t*>(mask_out);
  a.ldmask = ldmask;
  a.splits = splits; a.bn = bn;
  a.stream = reinterpret_cast<cudaStream_t>(stream);
  if (!a.a_mn && !a.b_mn) return launch_gemm_kk(a);
  if (!a.a_mn && a.b_mn) return launch_gemm_kmn(a);
  if (a.a_mn && a.b_mn) return launch_gemm_mnmn(a);
  set_error("dtf_gemm_bf16: layout (a_mn=1, b_mn=0) is not provided");
  return -2;
}

// Fused pull: H = act(gather(rows) | dense) W^T + bias.  `row_src` holds one device address per
// (row, slot) of a 128-wide bf16 embedding row (local or peer-mapped HBM); the gathered operand is
// also left in `a_buf` [M, slots*128 + dense_cols] for the weight-gradient GEMM.  `scratch` is
// caller-provided device memory of dtf_pull_gemm_scratch_bytes(M) bytes; it is cleared here.
extern "C" size_t dtf_pull_gemm_scratch_bytes(int M) {
  return (static_cast<size_t>((M + BM - 1) / BM) + 64) * sizeof(unsigned int);
}

extern "C" int dtf_pull_gemm_bf16(const unsigned long long* row_src, const void* dense_x, int slots, int dense_cols,
                                  void* a_buf, const void* W, long long ldw, const float* bias, void* H,
                                  long long ldh, void* mask_out, long long ldmask, int M, int N, int relu,
                                  void* scratch, void* stream) {
  if (M <= 0 || N <= 0 || slots <= 0 || dense_cols < 0 || (dense_cols % 64) != 0) {
    set_error("dtf_pull_gemm_bf16: bad shape M=%d N=%d slots=%d dense_cols=%d (dense_cols multiple of 64)", M, N,
              slots, dense_cols);
    return -2;
  }
  if (bias == nullptr || scratch == nullptr || row_src == nullptr || (dense_cols > 0 && dense_x == nullptr)) {
    set_error("dtf_pull_gemm_bf16: bias, scratch, row_src (and dense_x when dense_cols > 0) are required");
    return -2;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  DTF_CUDA_OK(cudaMemsetAsync(scratch, 0, dtf_pull_gemm_scratch_bytes(M), st));
  GemmArgs a{};
  a.stream = st;
  a.splits = 1;
  const int K0 = slots * 128 + dense_cols;
  a.A = a_buf; a.lda = K0; a.B = W; a.ldb = ldw; a.D = H; a.ldd = ldh;
  a.M = M; a.N = N; a.K = K0;
  a.epi = relu ? EPI_BIAS_RELU_BF16 : EPI_BIAS_BF16;
  a.bias = bias;
  a.mask_out = relu ? reinterpret_cast<uint32_t*>(mask_out) : nullptr;
  a.ldmask = ldmask;
  a.gather = true;
  a.row_src = row_src; a.dense_x = dense_x; a.slots = slots; a.dense_cols = dense_cols;
  a.chunk_counter = reinterpret_cast<unsigned int*>(scratch);
  a.tile_flags = reinterpret_cast<unsigned int*>(scratch) + 64;
  a.bn = (N % 256 == 0 || N > 512) ? 256 : 128;
  return launch_gemm_gather(a);
}

// Fused push: dA = dY W[:, :slots*128]; each 128-wide output row (m, slot) is written as bf16 to the
// address row_dst[m*slots + slot] (0 = skip) -- a mailbox entry in the owning rank's HBM.
extern "C" int dtf_push_gemm_bf16(const void* dY, long long lddy, const void* W, long long ldw,
                                  const unsigned long long* row_dst, int slots, int M, int K, void* stream) {
  if (M <= 0 || K <= 0 || slots <= 0 || row_dst == nullptr) {
    set_error("dtf_push_gemm_bf16: bad arguments");
    return -2;
  }
  GemmArgs a{};
  a.stream = reinterpret_cast<cudaStream_t>(stream);
  a.splits = 1;
  a.A = dY; a.lda = lddy; a.B = W; a.ldb = ldw; a.b_mn = true;
  a.M = M; a.N = slots * 128; a.K = K; a.epi = EPI_PUSH_ROWS; a.row_dst = row_dst; a.slots = slots;
  a.bn = 256;
  return launch_gemm_kmn(a);
}
