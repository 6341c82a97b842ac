// dtf_pull_gemm_bf16_noalloc: the fused pull (embedding gather + GEMM-1) for callers that must
// not allocate device memory.
//
// The ordinary entry point, dtf_pull_gemm_bf16, asks its caller for a scratch buffer (32 bytes of
// arrival counters per 128-row tile plus a 256-byte header).  Some callers have nowhere to get one: a
// serving process that handed the whole device to its own memory pool when it started, a plug-in
// living inside such a process, or a latency-critical request path where cudaMalloc's device-wide
// synchronisation and unbounded duration are not acceptable.  For them the scratch lives here, in a
// file-scope __device__ array that is part of the library image: it costs kScratchWords * 4 bytes
// (about 256 KiB) on every GPU that loads the library, from load until the process exits, and it
// caps the problem at kMaxRows rows per call.
//
// Rules that follow from sharing one array:
//   * Every call on a device uses the same array, and the array still holds whatever the previous
//     call left in it; the part a call needs is cleared (stream-ordered) before its kernel starts.
//   * Two calls whose GPU work overlaps in time (different streams, or different host threads
//     feeding different streams) race on the counters and produce garbage or hang.  Keep calls on
//     one device stream-ordered with respect to each other.
//   * The first call on each device looks the array's device address up with
//     cudaGetSymbolAddress (in host code the array's name is only a host-side handle, not a device
//     pointer, so it cannot be handed to a kernel or to cudaMemsetAsync directly).  Make that
//     first call during start-up, outside any stream capture.
//   * The call performs no cudaMalloc / cudaFree and uses no allocating library routine; inputs
//     and outputs remain the caller's buffers.
#include <mutex>

#include "../../include/dtf_kernels.h"
#include "gemm_launch.cuh"

using namespace dtf;

namespace {

constexpr int kMaxRows = 1 << 20;                          // 1,048,576 rows per call
constexpr int kMaxTiles = kMaxRows / BM;                   // 8192 row tiles
constexpr int kScratchWords = kMaxTiles * 8 + 64;          // word 0: reserved counter, words 64..: 32 bytes of chunk counters per tile
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned int g_pull_scratch[kScratchWords];

std::mutex g_mu;
unsigned int* g_scratch_addr[kMaxDevices] = {nullptr};     // device address of g_pull_scratch, per device

// Device address of the static scratch on the current device (looked up once per device).
int scratch_address(unsigned int** out) {
  int dev = -1;
  DTF_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("dtf_pull_gemm_bf16_noalloc: device ordinal %d outside [0, %d)", dev, kMaxDevices);
    return -2;
  }
  std::lock_guard<std::mutex> lock(g_mu);
  if (g_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_pull_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error("dtf_pull_gemm_bf16_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                cudaGetErrorString(e));
      (void)cudaGetLastError();
      return -1;
    }
    g_scratch_addr[dev] = static_cast<unsigned int*>(p);
  }
  *out = g_scratch_addr[dev];
  return 0;
}

}  // namespace

extern "C" int dtf_pull_gemm_noalloc_max_rows(void) { return kMaxRows; }

extern "C" int dtf_pull_gemm_bf16_noalloc(const unsigned long long* row_src, const void* dense_x, int slots,
                                          int dense_cols, void* a_buf, const void* W, long long ldw,
                                          const float* bias, void* H, long long ldh, void* mask_out,
                                          long long ldmask, int M, int N, int relu, void* stream) {
  if (M <= 0 || N <= 0 || slots <= 0 || (dense_cols != 0 && dense_cols != 64)) {
    set_error("dtf_pull_gemm_bf16_noalloc: bad shape M=%d N=%d slots=%d dense_cols=%d (dense_cols must be 0 or 64)",
              M, N, slots, dense_cols);
    return -2;
  }
  if (M > kMaxRows) {
    set_error("dtf_pull_gemm_bf16_noalloc: M=%d exceeds the built-in scratch (at most %d rows per call); "
              "split the batch or use dtf_pull_gemm_bf16 with a caller-provided scratch buffer",
              M, kMaxRows);
    return -2;
  }
  if (bias == nullptr || row_src == nullptr || a_buf == nullptr || W == nullptr || H == nullptr ||
      (dense_cols > 0 && dense_x == nullptr)) {
    set_error("dtf_pull_gemm_bf16_noalloc: row_src, a_buf, W, bias, H (and dense_x when dense_cols > 0) are required");
    return -2;
  }
  unsigned int* scratch = nullptr;
  const int rc = scratch_address(&scratch);
  if (rc != 0) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // the array keeps the previous call's counters: clear exactly what this call will use
  const size_t used_words = static_cast<size_t>((M + BM - 1) / BM) * 8 + 64;
  DTF_CUDA_OK(cudaMemsetAsync(scratch, 0, used_words * sizeof(unsigned int), st));
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
  a.chunk_counter = scratch;
  a.tile_flags = reinterpret_cast<unsigned long long*>(scratch + 64);
  a.bn = (N % 256 == 0 || N > 512) ? 256 : 128;
  a.pair = pair_eligible(a, a.bn);   // same tile shape choice as dtf_pull_gemm_bf16 (CTA pairs where the shape allows)
  return launch_gemm_gather(a);
}
