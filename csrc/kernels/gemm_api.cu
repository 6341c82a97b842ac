// C-ABI entry points for the dense GEMM family.
#include "../../include/dtf_kernels.h"
#include "gemm_launch.cuh"

using namespace dtf;

// CTA-pair mode (tcgen05.mma.cta_group::2, 256 x BN tiles on the two SMs of a TPC): process-wide
// switch, initialised from DTF_PAIR (default on); shapes / epilogues without a pair instantiation
// keep using single-CTA tiles.
static int g_pair_mode = -1;
namespace dtf {
bool pair_mode() {
  if (g_pair_mode < 0) {
    const char* e = getenv("DTF_PAIR");
    g_pair_mode = (e == nullptr || e[0] != '0') ? 1 : 0;
  }
  return g_pair_mode == 1;
}
bool pair_eligible(const GemmArgs& a, int bn) {
  if (!pair_mode() || a.M < 2 * BM || a.b_maps != nullptr) return false;
  const int e = a.epi;
  if (a.gather) return bn == 256 && (e == EPI_BIAS_RELU_BF16 || e == EPI_BIAS_BF16);
  if (!a.a_mn && !a.b_mn)
    return ((bn == 256 || bn == 128) && (e == EPI_BF16 || e == EPI_BIAS_BF16 || e == EPI_BIAS_RELU_BF16)) ||
           (bn == 256 && e == EPI_F32);
  if (!a.a_mn && a.b_mn)
    return (bn == 256 && (e == EPI_BF16 || e == EPI_DRELU_BF16 || e == EPI_PUSH_ROWS)) ||
           (bn == 128 && (e == EPI_BF16 || e == EPI_DRELU_BF16));
  if (a.a_mn && a.b_mn) return (bn == 256 && (e == EPI_F32 || e == EPI_F32_ATOMIC)) || (bn == 128 && e == EPI_F32_ATOMIC);
  return false;
}
}  // namespace dtf
extern "C" void dtf_gemm_set_pair_mode(int on) { g_pair_mode = on ? 1 : 0; }
extern "C" int dtf_gemm_get_pair_mode(void) { return dtf::pair_mode() ? 1 : 0; }

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
  a.splits = splits; a.bn = bn ? bn : choose_bn(M, N, splits);
  a.pair = pair_eligible(a, a.bn);
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
  return static_cast<size_t>((M + BM - 1) / BM) * kFlagWords * sizeof(unsigned long long) + 256;
}

// Debugging hook for bench/pull_trace.py: device buffer (>= 71 000 x 8 bytes) that receives globaltimer
// stamps of the gather warps and TMA producers of the next fused-pull launches; nullptr switches it off.
static unsigned long long* g_pull_trace = nullptr;
extern "C" void dtf_debug_set_pull_trace(void* buf) { g_pull_trace = reinterpret_cast<unsigned long long*>(buf); }

extern "C" int dtf_pull_gemm_bf16(const unsigned long long* row_src, const void* dense_x, int slots, int dense_cols,
                                  void* a_buf, const void* W, long long ldw, const float* bias, void* H,
                                  long long ldh, void* mask_out, long long ldmask, int M, int N, int relu,
                                  void* scratch, void* stream) {
  // the gather warps move the dense block as ONE 64-column chunk per row (gemm_sm100.cuh): wider blocks would
  // leave rows 8..15 of every slice unwritten, so only "no dense block" and "64 columns" are accepted
  if (M <= 0 || N <= 0 || slots <= 0 || (dense_cols != 0 && dense_cols != 64)) {
    set_error("dtf_pull_gemm_bf16: bad shape M=%d N=%d slots=%d dense_cols=%d (dense_cols must be 0 or 64)", M, N,
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
  a.tile_flags = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(scratch) + 256);
  a.trace = g_pull_trace;
  a.bn = (N % 256 == 0 || N > 512) ? 256 : 128;
  a.pair = pair_eligible(a, a.bn);
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
  a.pair = pair_eligible(a, a.bn);
  return launch_gemm_kmn(a);
}

// Fused dense-parameter pull: D = act(A W^T + bias) where W [N,K] is row-sharded over `shards`
// ranks (shard s holds rows [s*rows_per_shard, (s+1)*rows_per_shard)) and every shard pointer may be
// a peer-mapped address.  The TMA producer of each N-tile loads its B rows directly from the owning
// rank through that rank's tensor map, i.e. the all-gather happens inside the GEMM and no gathered
// copy of W is ever materialised.  maps_scratch: shards * 128 bytes of device memory (64-byte aligned)
// that receives the encoded tensor maps.  SURVEY.md 2E row K1b.
extern "C" int dtf_allgather_gemm_bf16(const void* A, long long lda, const void* const* W_shards, int shards,
                                       int rows_per_shard, long long ldw, void* D, long long ldd, int M, int N, int K,
                                       const float* bias, int relu, void* mask_out, long long ldmask,
                                       void* maps_scratch, void* stream) {
  if (shards < 1 || shards > 16 || rows_per_shard <= 0 || static_cast<long long>(shards) * rows_per_shard < N) {
    set_error("dtf_allgather_gemm_bf16: %d shards x %d rows do not cover N=%d (at most 16 shards)", shards,
              rows_per_shard, N);
    return -2;
  }
  int bn = 256;
  while (bn > 64 && (rows_per_shard % bn) != 0) bn >>= 1;
  if (rows_per_shard % bn != 0) {
    set_error("dtf_allgather_gemm_bf16: rows_per_shard=%d must be a multiple of 64", rows_per_shard);
    return -2;
  }
  if (bias == nullptr || maps_scratch == nullptr || (reinterpret_cast<uintptr_t>(maps_scratch) & 63)) {
    set_error("dtf_allgather_gemm_bf16: bias and a 64-byte aligned maps_scratch are required");
    return -2;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  CUtensorMap maps[16];
  for (int s = 0; s < shards; ++s) {
    const int rows = (s + 1) * rows_per_shard <= N ? rows_per_shard : (N - s * rows_per_shard > 0 ? N - s * rows_per_shard : 0);
    if (rows == 0) { maps[s] = maps[0]; continue; }
    if (!make_tmap_bf16_2d(&maps[s], W_shards[s], (uint64_t)K, (uint64_t)rows, (uint64_t)ldw * 2, BK, bn)) return -1;
  }
  DTF_CUDA_OK(cudaMemcpyAsync(maps_scratch, maps, sizeof(CUtensorMap) * shards, cudaMemcpyHostToDevice, st));
  GemmArgs a{};
  a.stream = st;
  a.splits = 1;
  a.A = A; a.lda = lda; a.B = W_shards[0]; a.ldb = ldw; a.D = D; a.ldd = ldd;
  a.M = M; a.N = N; a.K = K;
  a.epi = relu ? EPI_BIAS_RELU_BF16 : EPI_BIAS_BF16;
  a.bias = bias;
  a.mask_out = relu ? reinterpret_cast<uint32_t*>(mask_out) : nullptr;
  a.ldmask = ldmask;
  a.bn = bn;
  a.b_maps = reinterpret_cast<const CUtensorMap*>(maps_scratch);
  a.b_rows_per_shard = rows_per_shard;
  return launch_gemm_kk(a);
}
