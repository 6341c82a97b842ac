// Host-side launch plumbing for gemm_sm100_kernel: tensor maps, tiling decisions, dispatch.
#pragma once

#include "common.cuh"
#include "gemm_sm100.cuh"

namespace dtf {

struct GemmArgs {
  const void* A; long long lda; bool a_mn;
  const void* B; long long ldb; bool b_mn;
  void* D; long long ldd;
  int M, N, K;
  int epi;
  const float* bias;
  const uint32_t* mask_in; uint32_t* mask_out; long long ldmask;
  int splits;
  int bn;  // 0 = choose
  // fused pull
  bool gather;
  const unsigned long long* row_src;
  const void* dense_x;
  int slots, dense_cols;
  unsigned long long* tile_flags;
  unsigned int* chunk_counter;
  unsigned long long* trace = nullptr;
  unsigned long long* dbg = nullptr;   // DTF_GEMM_DBG=1 (see GemmParams::dbg)
  int dbg_id = 0;
  const unsigned long long* wide_src = nullptr; float* wide_out = nullptr;   // fused pull: side gather of fp32 scalars
  // fused dense pull (B row-sharded over peers)
  const CUtensorMap* b_maps; int b_rows_per_shard;
  // fused push
  const unsigned long long* row_dst;
  bool pdl;   // launch with programmatic stream serialization (step executor)
  int max_ctas = 0;    // cap on the persistent grid (0 = all SMs): lets two GEMMs share the GPU side by side
  bool pair = false;   // CTA pairs (cta_group::2, 256 x BN tiles); BN = 256 or 128 only
  cudaStream_t stream;
};

template <int BN, bool A_MN, bool B_MN, int EPI, bool GATHER, int CTAS = 1>
int launch_gemm_t(const GemmArgs& a) {
  using S = GemmSmem<BN, GATHER, CTAS>;
  auto kern = gemm_sm100_kernel<BN, A_MN, B_MN, EPI, GATHER, CTAS>;
  static bool attr_set[64] = {false};
  int dev = 0;
  DTF_CUDA_OK(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    DTF_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
    attr_set[dev & 63] = true;
  }
  CUtensorMap tmA, tmB;
  if (!A_MN) {
    if (!make_tmap_bf16_2d(&tmA, a.A, (uint64_t)a.K, (uint64_t)a.M, (uint64_t)a.lda * 2, BK, BM)) return -1;
  } else {
    if (!make_tmap_bf16_2d(&tmA, a.A, (uint64_t)a.M, (uint64_t)a.K, (uint64_t)a.lda * 2, 64, BK)) return -1;
  }
  if (!B_MN) {
    if (!make_tmap_bf16_2d(&tmB, a.B, (uint64_t)a.K, (uint64_t)a.N, (uint64_t)a.ldb * 2, BK, BN / CTAS)) return -1;
  } else {
    if (!make_tmap_bf16_2d(&tmB, a.B, (uint64_t)a.N, (uint64_t)a.K, (uint64_t)a.ldb * 2, 64, BK)) return -1;
  }
  CUtensorMap tmD;
  constexpr bool kOutF32 = (EPI == EPI_F32 || EPI == EPI_F32_ATOMIC);
  if (EPI != EPI_PUSH_ROWS) {
    if (!make_tmap_2d(&tmD, a.D, kOutF32 ? 4 : 2, (uint64_t)a.N, (uint64_t)a.M,
                      (uint64_t)a.ldd * (kOutF32 ? 4 : 2), kOutF32 ? 32 : 64, 32))
      return -1;
  } else {
    tmD = tmA;  // unused by the push epilogue
  }
  if (GATHER && a.slots + 1 > kFlagWords * 16) {
    set_error("fused pull: at most %d slots per example (got %d)", kFlagWords * 16 - 1, a.slots);
    return -2;
  }
  GemmParams p{};
  p.M = a.M; p.N = a.N; p.K = a.K;
  p.m_tiles = (a.M + BM - 1) / BM;
  p.n_tiles = (a.N + BN - 1) / BN;
  const int total_kb = (a.K + BK - 1) / BK;
  int splits = a.splits < 1 ? 1 : a.splits;
  if (splits > total_kb) splits = total_kb;
  p.k_blocks_per_split = (total_kb + splits - 1) / splits;
  p.splits = (total_kb + p.k_blocks_per_split - 1) / p.k_blocks_per_split;
  p.D = a.D; p.ldd = a.ldd;
  p.bias = a.bias;
  p.mask_in = a.mask_in; p.mask_out = a.mask_out; p.ldmask = a.ldmask;
  p.row_src = a.row_src;
  p.dense_x = reinterpret_cast<const __nv_bfloat16*>(a.dense_x);
  p.a_buf = reinterpret_cast<__nv_bfloat16*>(const_cast<void*>(a.A));
  p.lda_buf = a.lda;
  p.slots = a.slots; p.dense_cols = a.dense_cols;
  p.tile_flags = a.tile_flags; p.chunk_counter = a.chunk_counter; p.trace = a.trace; p.dbg = a.dbg; p.dbg_id = a.dbg_id; p.wide_src = a.wide_src; p.wide_out = a.wide_out;
  p.row_dst = a.row_dst;
  p.b_maps = a.b_maps; p.b_rows_per_shard = a.b_rows_per_shard;
  const int tiles = ((p.m_tiles + CTAS - 1) / CTAS) * p.n_tiles * p.splits * CTAS;   // in CTAs
  int sms = num_sms();
  if (a.max_ctas > 0 && a.max_ctas < sms && !GATHER) sms = a.max_ctas;
  sms = sms / CTAS * CTAS;
  if (sms < CTAS) sms = CTAS;
  int grid = tiles < sms ? tiles : sms;
  if (GATHER && grid < sms) grid = sms;  // gather warps of CTAs without tiles still pull rows
  const int threads = GATHER ? kNumMainThreads + kNumGatherThreads : kNumMainThreads;
  DTF_CUDA_OK(launch_kernel_cluster(a.pdl, CTAS, kern, dim3(grid), dim3(threads), S::kTotal, a.stream, tmA, tmB, tmD, p));
  count_launch();
  return 0;
}

// picks the N tile that wastes the fewest SM-waves for this shape
inline int choose_bn(int M, int N, int splits) {
  const int sms = num_sms();
  const int cand[3] = {256, 128, 64};
  int best = 256;
  double best_cost = 1e30;
  for (int i = 0; i < 3; ++i) {
    const int bn = cand[i];
    const long long tiles = (long long)((M + BM - 1) / BM) * ((N + bn - 1) / bn) * (splits < 1 ? 1 : splits);
    const long long waves = (tiles + sms - 1) / sms;
    // per-tile cost ~ bn (MMA time) plus a fixed overhead that penalises tiny tiles
    const double cost = (double)waves * (bn + 24.0);
    if (cost < best_cost) { best_cost = cost; best = bn; }
  }
  return best;
}

bool pair_mode();                                     // DTF_PAIR / dtf_gemm_set_pair_mode
bool pair_eligible(const GemmArgs& a, int bn);        // is there a CTA-pair instantiation for this launch
int launch_gemm_kk(const GemmArgs& a);    // A K-major,  B K-major   (forward)
int launch_gemm_kmn(const GemmArgs& a);   // A K-major,  B MN-major  (dgrad)
int launch_gemm_mnmn(const GemmArgs& a);  // A MN-major, B MN-major  (wgrad)
int launch_gemm_gather(const GemmArgs& a);  // fused pull: gather warps + forward GEMM

}  // namespace dtf
