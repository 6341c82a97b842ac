// Forward-layout GEMMs: A [M,K] K-major, B [N,K] K-major.
#include "gemm_launch.cuh"

namespace dtf {

#define DTF_CASE(BN_, EPI_) \
  if (bn == BN_ && a.epi == EPI_) return launch_gemm_t<BN_, false, false, EPI_, false>(a);

#define DTF_PAIR_CASE(BN_, EPI_) \
  if (bn == BN_ && a.epi == EPI_) return launch_gemm_t<BN_, false, false, EPI_, false, 2>(a);

int launch_gemm_kk(const GemmArgs& a) {
  const int bn = a.bn ? a.bn : choose_bn(a.M, a.N, a.splits);
  if (a.pair) {
    DTF_PAIR_CASE(256, EPI_BF16) DTF_PAIR_CASE(256, EPI_BIAS_BF16) DTF_PAIR_CASE(256, EPI_BIAS_RELU_BF16)
    DTF_PAIR_CASE(256, EPI_F32)
    DTF_PAIR_CASE(128, EPI_BF16) DTF_PAIR_CASE(128, EPI_BIAS_BF16) DTF_PAIR_CASE(128, EPI_BIAS_RELU_BF16)
    set_error("gemm (K-major x K-major, CTA pairs): unsupported bn=%d epilogue=%d", bn, a.epi);
    return -2;
  }
  DTF_CASE(256, EPI_BF16) DTF_CASE(128, EPI_BF16) DTF_CASE(64, EPI_BF16)
  DTF_CASE(256, EPI_BIAS_BF16) DTF_CASE(128, EPI_BIAS_BF16) DTF_CASE(64, EPI_BIAS_BF16)
  DTF_CASE(256, EPI_BIAS_RELU_BF16) DTF_CASE(128, EPI_BIAS_RELU_BF16) DTF_CASE(64, EPI_BIAS_RELU_BF16)
  DTF_CASE(256, EPI_F32) DTF_CASE(128, EPI_F32) DTF_CASE(64, EPI_F32)
  set_error("gemm (K-major x K-major): unsupported bn=%d epilogue=%d", bn, a.epi);
  return -2;
}

}  // namespace dtf
