// Weight-gradient GEMMs: dW[N_out, K_in] = dY^T X with both operands read in their forward
// layouts (MN-major), reduction over the batch; split-K partials are added with red.global.
#include "gemm_launch.cuh"

namespace dtf {

#define DTF_CASE(BN_, EPI_) \
  if (bn == BN_ && a.epi == EPI_) return launch_gemm_t<BN_, true, true, EPI_, false>(a);

int launch_gemm_mnmn(const GemmArgs& a) {
  const int bn = a.bn ? a.bn : choose_bn(a.M, a.N, a.splits);
  if (a.pair) {
#define DTF_PAIR_CASE(BN_, EPI_) \
  if (bn == BN_ && a.epi == EPI_) return launch_gemm_t<BN_, true, true, EPI_, false, 2>(a);
    DTF_PAIR_CASE(256, EPI_F32) DTF_PAIR_CASE(256, EPI_F32_ATOMIC) DTF_PAIR_CASE(128, EPI_F32_ATOMIC)
    set_error("gemm (MN-major x MN-major, CTA pairs): unsupported bn=%d epilogue=%d", bn, a.epi);
    return -2;
  }
  DTF_CASE(256, EPI_F32) DTF_CASE(128, EPI_F32) DTF_CASE(64, EPI_F32)
  DTF_CASE(256, EPI_F32_ATOMIC) DTF_CASE(128, EPI_F32_ATOMIC) DTF_CASE(64, EPI_F32_ATOMIC)
  DTF_CASE(256, EPI_BF16) DTF_CASE(128, EPI_BF16)
  set_error("gemm (MN-major x MN-major): unsupported bn=%d epilogue=%d", bn, a.epi);
  return -2;
}

}  // namespace dtf
