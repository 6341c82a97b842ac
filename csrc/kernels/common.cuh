// Host-side helpers shared by every .cu in libdtf_kernels: error reporting through the C ABI,
// launch accounting (bench.py reports "gpu_launches" from it) and TMA tensor-map construction.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace dtf {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);
int num_sms();

// 2-D bf16 tensor map with 128-byte swizzle.  `inner` is the contiguous dimension.
// Returns false (and sets the error string) on failure.
bool make_tmap_2d_ex(CUtensorMap* out, const void* base, int elem_bytes /*2=bf16, 4=f32*/, uint64_t inner,
                     uint64_t outer, uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer, bool swizzle128);
inline bool make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t inner, uint64_t outer,
                         uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer) {
  return make_tmap_2d_ex(out, base, elem_bytes, inner, outer, row_stride_bytes, box_inner, box_outer, true);
}
inline bool make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer,
                              uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer) {
  return make_tmap_2d(out, base, 2, inner, outer, row_stride_bytes, box_inner, box_outer);
}

// Launch with the programmatic-stream-serialization attribute (PDL): the kernel may be scheduled while
// its predecessor in the stream drains; it must execute griddepcontrol.wait before touching memory
// the predecessor writes.  Measured on the wide-and-deep step (1 x B200): serial eager 0.470 -> 0.445 ms,
// no change under CUDA-graph replay, and the async mode that overlaps the ps-side apply on a second
// stream gets slower (0.388 -> 0.438 ms: early-resident dependents take the room the apply blocks use),
// so the attribute is opt-in: DTF_PDL=1.
bool pdl_enabled();
template <typename... KArgs, typename... Args>
cudaError_t launch_kernel(bool pdl, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                          Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = (pdl && pdl_enabled()) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// same, with the grid cut into clusters of `cluster` CTAs along x (1 = no cluster attribute)
template <typename... KArgs, typename... Args>
cudaError_t launch_kernel_cluster(bool pdl, int cluster, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                  cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (pdl && pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (cluster > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#define DTF_CUDA_OK(expr)                                                               \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      ::dtf::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                       __LINE__);                                                       \
      return -1;                                                                        \
    }                                                                                   \
  } while (0)

}  // namespace dtf
