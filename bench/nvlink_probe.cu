// SM-issued peer-memory access patterns over NVLink, GPU 0 <-> GPU 1 (needs 2 GPUs).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/nvlink_probe bench/nvlink_probe.cu && /tmp/nvlink_probe
// Every case moves the same number of bytes (ROWS_TOTAL x 256 B); rows are `seg` bytes long when
// random.  Reads land in a local buffer (like the gather), writes take their data from a local buffer
// (like the push).  Timed with CUDA events, best of 5.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define OK(x)                                                                  \
  do {                                                                         \
    cudaError_t e_ = (x);                                                      \
    if (e_ != cudaSuccess) {                                                   \
      printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));        \
      exit(1);                                                                 \
    }                                                                          \
  } while (0)

// One "row" = seg bytes handled by seg/16 consecutive lanes; a warp covers 512 B per instruction and
// keeps UNROLL independent instructions in flight.  idx[i] = source row of output row i.
template <int UNROLL, bool WRITE>
__global__ void __launch_bounds__(256) rows_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst,
                                                   const unsigned int* __restrict__ idx, long long nrows, int lanes_per_row) {
  const int lane = threadIdx.x & 31;
  const long long warp = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  const int rows_per_inst = 32 / lanes_per_row;
  const int sub = lane / lanes_per_row, off = lane % lanes_per_row;
  const long long ninst = nrows / rows_per_inst;
  for (long long i = warp * UNROLL; i < ninst; i += nwarps * UNROLL) {
    uint4 v[UNROLL];
    long long r[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const long long o = (i + u) * rows_per_inst + sub;
      r[u] = o < nrows ? static_cast<long long>(idx[o]) : -1;
    }
    if (!WRITE) {
#pragma unroll
      for (int u = 0; u < UNROLL; ++u)
        if (r[u] >= 0) v[u] = __ldg(src + r[u] * lanes_per_row + off);
#pragma unroll
      for (int u = 0; u < UNROLL; ++u)
        if (r[u] >= 0) dst[((i + u) * rows_per_inst + sub) * lanes_per_row + off] = v[u];
    } else {
#pragma unroll
      for (int u = 0; u < UNROLL; ++u)
        if (r[u] >= 0) v[u] = __ldg(src + ((i + u) * rows_per_inst + sub) * lanes_per_row + off);
#pragma unroll
      for (int u = 0; u < UNROLL; ++u)
        if (r[u] >= 0) dst[r[u] * lanes_per_row + off] = v[u];
    }
  }
}

struct Case {
  const char* name;
  bool write;
  bool remote;
  bool random;
  int seg;
  int unroll;
  int blocks_per_sm;
};

int main() {
  int n = 0;
  OK(cudaGetDeviceCount(&n));
  if (n < 2) {
    printf("needs 2 GPUs\n");
    return 0;
  }
  OK(cudaSetDevice(0));
  int can = 0;
  OK(cudaDeviceCanAccessPeer(&can, 0, 1));
  if (!can) {
    printf("no peer access\n");
    return 0;
  }
  OK(cudaDeviceEnablePeerAccess(1, 0));
  int sms = 0;
  OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const long long table_bytes = 1280ll << 20;      // 1.25 GB table, as one shard of the 10M x 128 bf16 table at 2 ranks
  const long long move_bytes = 8192ll * 26 * 256;  // one step's gather
  uint4 *local_tab, *remote_tab, *local_buf;
  OK(cudaMalloc(&local_tab, table_bytes));
  OK(cudaMalloc(&local_buf, move_bytes));
  OK(cudaSetDevice(1));
  OK(cudaMalloc(&remote_tab, table_bytes));
  OK(cudaMemset(remote_tab, 1, table_bytes));
  OK(cudaDeviceSynchronize());
  OK(cudaSetDevice(0));
  OK(cudaMemset(local_tab, 1, table_bytes));
  OK(cudaMemset(local_buf, 2, move_bytes));
  unsigned int* idx;
  OK(cudaMalloc(&idx, (move_bytes / 64) * sizeof(unsigned int)));
  char* flush;
  OK(cudaMalloc(&flush, 256 << 20));
  cudaEvent_t e0, e1;
  OK(cudaEventCreate(&e0));
  OK(cudaEventCreate(&e1));

  const Case cases[] = {
      {"read  local  seq        ", false, false, false, 256, 4, 8}, {"read  remote seq        ", false, true, false, 256, 4, 8},
      {"read  remote seq  u8    ", false, true, false, 256, 8, 8},  {"read  local  rand 256B  ", false, false, true, 256, 4, 8},
      {"read  remote rand 64B   ", false, true, true, 64, 4, 8},    {"read  remote rand 128B  ", false, true, true, 128, 4, 8},
      {"read  remote rand 256B  ", false, true, true, 256, 4, 8},   {"read  remote rand 256B u8", false, true, true, 256, 8, 8},
      {"read  remote rand 256B u2", false, true, true, 256, 2, 8},  {"read  remote rand 256B b2", false, true, true, 256, 4, 2},
      {"read  remote rand 256B b1", false, true, true, 256, 4, 1},  {"read  remote rand 512B  ", false, true, true, 512, 4, 8},
      {"write local  rand 256B  ", true, false, true, 256, 4, 8},   {"write remote seq        ", true, true, false, 256, 4, 8},
      {"write remote rand 128B  ", true, true, true, 128, 4, 8},    {"write remote rand 256B  ", true, true, true, 256, 4, 8},
      {"write remote rand 256B b2", true, true, true, 256, 4, 2},   {"write remote rand 512B  ", true, true, true, 512, 4, 8},
  };
  for (const Case& c : cases) {
    const long long nrows = move_bytes / c.seg;
    const long long table_rows = table_bytes / c.seg;
    std::vector<unsigned int> h(nrows);
    unsigned long long s = 88172645463325252ull;
    for (long long i = 0; i < nrows; ++i) {
      s ^= s << 13;
      s ^= s >> 7;
      s ^= s << 17;
      h[i] = c.random ? static_cast<unsigned int>(s % table_rows) : static_cast<unsigned int>(i);
    }
    OK(cudaMemcpy(idx, h.data(), nrows * sizeof(unsigned int), cudaMemcpyHostToDevice));
    uint4* tab = c.remote ? remote_tab : local_tab;
    const uint4* src = c.write ? local_buf : tab;
    uint4* dst = c.write ? tab : local_buf;
    const int lanes = c.seg / 16;
    float best = 1e9f;
    for (int it = 0; it < 6; ++it) {
      OK(cudaMemsetAsync(flush, it, 256 << 20));
      OK(cudaEventRecord(e0));
      const int grid = sms * c.blocks_per_sm;
#define LAUNCH(U)                                                                      \
  if (c.write)                                                                         \
    rows_kernel<U, true><<<grid, 256>>>(src, dst, idx, nrows, lanes);                  \
  else                                                                                 \
    rows_kernel<U, false><<<grid, 256>>>(src, dst, idx, nrows, lanes);
      if (c.unroll == 2) {
        LAUNCH(2)
      } else if (c.unroll == 8) {
        LAUNCH(8)
      } else {
        LAUNCH(4)
      }
      OK(cudaEventRecord(e1));
      OK(cudaEventSynchronize(e1));
      float ms;
      OK(cudaEventElapsedTime(&ms, e0, e1));
      if (it > 0 && ms < best) best = ms;
    }
    printf("%s  %7.1f us  %6.1f GB/s\n", c.name, best * 1e3f, move_bytes / (best * 1e-3f) / 1e9f);
    fflush(stdout);
  }
  return 0;
}
