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
