// Fused pull: embedding rows are gathered from their owning ranks' HBM (peer-mapped pointers) by
// dedicated warps of the same persistent kernel whose MMA warps run GEMM-1 on row-tiles gathered
// earlier.
#include "gemm_launch.cuh"

namespace dtf {

#define DTF_CASE(BN_, EPI_) \
  if (bn == BN_ && a.epi == EPI_) return launch_gemm_t<BN_, false, false, EPI_, true>(a);

int launch_gemm_gather(const GemmArgs& a) {
  const int bn = a.bn ? a.bn : 256;
  if (a.pair) {
#define DTF_PAIR_CASE(BN_, EPI_) \
  if (bn == BN_ && a.epi == EPI_) return launch_gemm_t<BN_, false, false, EPI_, true, 2>(a);
    DTF_PAIR_CASE(256, EPI_BIAS_RELU_BF16) DTF_PAIR_CASE(256, EPI_BIAS_BF16)
    set_error("fused pull gemm (CTA pairs): unsupported bn=%d epilogue=%d", bn, a.epi);
    return -2;
  }
  DTF_CASE(256, EPI_BIAS_RELU_BF16) DTF_CASE(128, EPI_BIAS_RELU_BF16)
  DTF_CASE(256, EPI_BIAS_BF16) DTF_CASE(128, EPI_BIAS_BF16)
  set_error("fused pull gemm: unsupported bn=%d epilogue=%d", bn, a.epi);
  return -2;
}

}  // namespace dtf
