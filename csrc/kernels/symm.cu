// Symmetric heap: one cudaMalloc'd slab per rank whose CUDA-IPC handle is exchanged over the host
// control plane and mapped by every peer, so that device code can address any rank's parameter
// shard, mailbox or signal pad with plain loads/stores over NVLink 5 / NVSwitch.
//
// This is the data plane that replaces the gRPC RecvTensor traffic behind every sess.run of the
// reference (reference trainer/task.py:236).  The host side only exchanges 64-byte
// handles once at start-up.
#include "../../include/dtf_kernels.h"
#include "common.cuh"

using namespace dtf;

extern "C" {

int dtf_symm_alloc(size_t bytes, void** out_ptr) {
  void* p = nullptr;
  DTF_CUDA_OK(cudaMalloc(&p, bytes));
  DTF_CUDA_OK(cudaMemset(p, 0, bytes));
  *out_ptr = p;
  return 0;
}

int dtf_symm_free(void* ptr) {
  DTF_CUDA_OK(cudaFree(ptr));
  return 0;
}

int dtf_symm_handle_size(void) { return static_cast<int>(sizeof(cudaIpcMemHandle_t)); }

int dtf_symm_get_handle(void* ptr, void* handle_out) {
  cudaIpcMemHandle_t h;
  DTF_CUDA_OK(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle_out, &h, sizeof(h));
  return 0;
}

int dtf_symm_open_handle(const void* handle, void** out_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  DTF_CUDA_OK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *out_ptr = p;
  return 0;
}

int dtf_symm_close_handle(void* ptr) {
  DTF_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

// Same-process peers (several "ranks" driven from one process, e.g. tests on one box with several
// GPUs): enable direct peer access instead of IPC.
int dtf_symm_enable_peer(int peer_device) {
  int dev = 0;
  DTF_CUDA_OK(cudaGetDevice(&dev));
  if (dev == peer_device) return 0;
  int can = 0;
  DTF_CUDA_OK(cudaDeviceCanAccessPeer(&can, dev, peer_device));
  if (!can) {
    set_error("device %d cannot access peer %d", dev, peer_device);
    return -3;
  }
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
    set_error("cudaDeviceEnablePeerAccess(%d) failed: %s", peer_device, cudaGetErrorString(e));
    return -1;
  }
  (void)cudaGetLastError();
  return 0;
}

}  // extern "C"
