#include "common.cuh"

#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>

#include "../../include/dtf_kernels.h"

namespace dtf {

static thread_local char g_err[1024] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("DTF_PDL");
    v = (e != nullptr && e[0] == '1') ? 1 : 0;   // off unless DTF_PDL=1, see common.cuh
  }
  return v == 1;
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    // The driver is reached through the runtime so that the library carries no link-time
    // dependency on libcuda (it must load on the GPU-less build box).
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

bool make_tmap_2d_ex(CUtensorMap* out, const void* base, int elem_bytes, uint64_t inner, uint64_t outer,
                     uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer, bool swizzle128) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
    return false;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15) != 0 || (row_stride_bytes & 15) != 0) {
    set_error("TMA operand must be 16-byte aligned (base %p, row stride %llu B)", base,
              (unsigned long long)row_stride_bytes);
    return false;
  }
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box,
                   estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (inner=%llu outer=%llu stride=%llu box=%ux%u)",
              (int)r, (unsigned long long)inner, (unsigned long long)outer,
              (unsigned long long)row_stride_bytes, box_inner, box_outer);
    return false;
  }
  return true;
}

}  // namespace dtf

// Copy by kernel.  dst / src may be device memory or pinned host memory (directly addressable under UVA),
// so this is a host->device or device->host transfer driven by a few SMs instead of a copy-engine
// operation: inside CUDA graphs that also contain kernels spinning on peer signals, copy-engine memcpy
// nodes stalled intermittently (bench/stress_overlap.py), kernel nodes do not.
__global__ void dtf_copy_kernel(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, size_t bytes) {
  const size_t n16 = ((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15) == 0 ? bytes / 16 : 0;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t t = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  // four independent 16-byte loads in flight per thread: PCIe round trips are long
  for (size_t i = t; i < n16; i += 4 * stride) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < n16) v[u] = reinterpret_cast<const uint4*>(src)[i + u * stride];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < n16) reinterpret_cast<uint4*>(dst)[i + u * stride] = v[u];
  }
  for (size_t i = n16 * 16 + t; i < bytes; i += stride) dst[i] = src[i];
}

extern "C" {

int dtf_copy_async(void* dst, const void* src, size_t bytes, int blocks, void* stream) {
  if (bytes == 0) return 0;
  if (dst == nullptr || src == nullptr) {
    dtf::set_error("dtf_copy_async: null pointer");
    return -2;
  }
  if (blocks < 1) blocks = 1;
  dtf_copy_kernel<<<blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(static_cast<uint8_t*>(dst),
                                                                              static_cast<const uint8_t*>(src), bytes);
  DTF_CUDA_OK(cudaGetLastError());
  dtf::count_launch();
  return 0;
}

const char* dtf_last_error(void) { return dtf::g_err; }

long long dtf_launch_count(void) { return dtf::g_launches.load(std::memory_order_relaxed); }

void dtf_reset_launch_count(void) { dtf::g_launches.store(0, std::memory_order_relaxed); }

int dtf_kernels_abi_version(void) { return DTF_KERNELS_ABI_VERSION; }

// sha256 over csrc/kernels + include at build time (utils/build.py passes it in): lets build() and the
// loader tell a library that was built from other sources than the ones in the tree
#ifndef DTF_SOURCE_HASH
#define DTF_SOURCE_HASH "unstamped"
#endif
const char* dtf_kernels_source_hash(void) { return DTF_SOURCE_HASH; }

}  // extern "C"
