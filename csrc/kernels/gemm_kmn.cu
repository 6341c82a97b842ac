// Data-gradient GEMMs: A [M,K] K-major (dY), B stored [K,N] i.e. MN-major (the forward weight
// read without a transpose).  Includes the fused push epilogue.
#include "gemm_launch.cuh"

namespace dtf {

#define DTF_CASE(BN_, EPI_) \
  if (bn == BN_ && a.epi == EPI_) return launch_gemm_t<BN_, false, true, EPI_, false>(a);

int launch_gemm_kmn(const GemmArgs& a) {
  int bn = a.bn ? a.bn : choose_bn(a.M, a.N, a.splits);
  if (a.epi == EPI_PUSH_ROWS && bn == 64) bn = 128;
  if (a.pair) {
#define DTF_PAIR_CASE(BN_, EPI_) \
  if (bn == BN_ && a.epi == EPI_) return launch_gemm_t<BN_, false, true, EPI_, false, 2>(a);
    DTF_PAIR_CASE(256, EPI_BF16) DTF_PAIR_CASE(256, EPI_DRELU_BF16) DTF_PAIR_CASE(256, EPI_PUSH_ROWS)
    DTF_PAIR_CASE(128, EPI_BF16) DTF_PAIR_CASE(128, EPI_DRELU_BF16)
    set_error("gemm (K-major x MN-major, CTA pairs): unsupported bn=%d epilogue=%d", bn, a.epi);
    return -2;
  }
  DTF_CASE(256, EPI_BF16) DTF_CASE(128, EPI_BF16) DTF_CASE(64, EPI_BF16)
  DTF_CASE(256, EPI_DRELU_BF16) DTF_CASE(128, EPI_DRELU_BF16) DTF_CASE(64, EPI_DRELU_BF16)
  DTF_CASE(256, EPI_PUSH_ROWS) DTF_CASE(128, EPI_PUSH_ROWS)
  DTF_CASE(256, EPI_F32) DTF_CASE(128, EPI_F32)
  set_error("gemm (K-major x MN-major): unsupported bn=%d epilogue=%d", bn, a.epi);
  return -2;
}

}  // namespace dtf
