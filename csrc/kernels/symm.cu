// Symmetric heap: one slab of device memory per rank that every peer maps, so that device code can
// address any rank's parameter shard, mailbox or signal pad with plain loads/stores over NVLink 5 /
// NVSwitch.  This is the data plane that replaces the gRPC RecvTensor traffic behind every sess.run
// of the reference (reference trainer/task.py:236); the host side only passes one
// file descriptor per rank at start-up.
//
// Slabs are created with the virtual-memory-management API (cuMemCreate, 2 MB granularity) and
// shared as POSIX file descriptors.  Measured on 2 x B200: a peer mapping made by
// cudaIpcOpenMemHandle serves the fused pull's 256-byte row reads 3x slower than a VMM / direct
// peer mapping (486 us vs 165 us for 54 MB), so the legacy IPC handle path is kept only as a
// fallback for drivers that cannot export file descriptors.
#include <unistd.h>

#include <map>
#include <mutex>

#include "../../include/dtf_kernels.h"
#include "common.cuh"

using namespace dtf;

namespace {

struct Drv {
  CUresult (*GetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*Create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*Release)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*AddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*AddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*Map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*Unmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*SetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*Export)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*Import)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
  // multicast objects (NVLS): optional, mc_ok says whether the entry points exist
  CUresult (*McCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*McAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                        unsigned long long) = nullptr;
  CUresult (*McGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
  CUresult (*DeviceGet)(CUdevice*, int) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  bool ok = false;
  bool mc_ok = false;
};

template <typename T>
bool load(const char* name, T* fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
    (void)cudaGetLastError();
    return false;
  }
  *fn = reinterpret_cast<T>(p);
  return true;
}

Drv& drv() {
  static Drv d;
  static std::once_flag once;
  std::call_once(once, [] {
    d.ok = load("cuMemGetAllocationGranularity", &d.GetGranularity) && load("cuMemCreate", &d.Create) &&
           load("cuMemRelease", &d.Release) && load("cuMemAddressReserve", &d.AddressReserve) &&
           load("cuMemAddressFree", &d.AddressFree) && load("cuMemMap", &d.Map) && load("cuMemUnmap", &d.Unmap) &&
           load("cuMemSetAccess", &d.SetAccess) && load("cuMemExportToShareableHandle", &d.Export) &&
           load("cuMemImportFromShareableHandle", &d.Import);
    d.mc_ok = d.ok && load("cuMulticastCreate", &d.McCreate) && load("cuMulticastAddDevice", &d.McAddDevice) &&
              load("cuMulticastBindMem", &d.McBindMem) && load("cuMulticastGetGranularity", &d.McGetGranularity) &&
              load("cuDeviceGet", &d.DeviceGet) && load("cuDeviceGetAttribute", &d.DeviceGetAttribute);
  });
  return d;
}

struct Mapping {
  CUmemGenericAllocationHandle handle = 0;
  size_t size = 0;   // rounded to the granularity
  bool vmm = false;  // false: legacy cudaMalloc allocation
  bool imported = false;
  bool owns_handle = true;   // false: the handle belongs to a multicast object (released by dtf_mc_destroy)
};
std::mutex g_mu;
std::map<void*, Mapping> g_maps;

int map_handle(CUmemGenericAllocationHandle h, size_t size, int dev, void** out) {
  Drv& d = drv();
  CUdeviceptr va = 0;
  CUresult r = d.AddressReserve(&va, size, 0, 0, 0);
  if (r != CUDA_SUCCESS) {
    set_error("symm: cuMemAddressReserve(%zu) failed (%d)", size, (int)r);
    return -1;
  }
  r = d.Map(va, size, 0, h, 0);
  if (r != CUDA_SUCCESS) {
    d.AddressFree(va, size);
    set_error("symm: cuMemMap failed (%d)", (int)r);
    return -1;
  }
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d.SetAccess(va, size, &acc, 1);
  if (r != CUDA_SUCCESS) {
    d.Unmap(va, size);
    d.AddressFree(va, size);
    set_error("symm: cuMemSetAccess for device %d failed (%d) -- no peer access between the GPUs?", dev, (int)r);
    return -1;
  }
  *out = reinterpret_cast<void*>(va);
  return 0;
}

}  // namespace

extern "C" {

int dtf_symm_alloc(size_t bytes, void** out_ptr) {
  int dev = 0;
  DTF_CUDA_OK(cudaGetDevice(&dev));
  DTF_CUDA_OK(cudaFree(nullptr));  // make sure the primary context exists before driver calls
  Drv& d = drv();
  if (d.ok) {
    CUmemAllocationProp prop{};
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = dev;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t gran = 0;
    if (d.GetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && gran > 0) {
      const size_t size = (bytes + gran - 1) / gran * gran;
      CUmemGenericAllocationHandle h = 0;
      const CUresult r = d.Create(&h, size, &prop, 0);
      if (r == CUDA_SUCCESS) {
        void* p = nullptr;
        if (map_handle(h, size, dev, &p) != 0) {
          d.Release(h);
          return -1;
        }
        DTF_CUDA_OK(cudaMemset(p, 0, size));
        std::lock_guard<std::mutex> l(g_mu);
        g_maps[p] = Mapping{h, size, true, false};
        *out_ptr = p;
        return 0;
      }
    }
  }
  // fallback: plain cudaMalloc (shareable through the legacy IPC handle only)
  void* p = nullptr;
  DTF_CUDA_OK(cudaMalloc(&p, bytes));
  DTF_CUDA_OK(cudaMemset(p, 0, bytes));
  std::lock_guard<std::mutex> l(g_mu);
  g_maps[p] = Mapping{0, bytes, false, false};
  *out_ptr = p;
  return 0;
}

int dtf_symm_free(void* ptr) {
  Mapping m;
  {
    std::lock_guard<std::mutex> l(g_mu);
    auto it = g_maps.find(ptr);
    if (it == g_maps.end()) {
      set_error("symm: %p was not allocated by dtf_symm_alloc / imported", ptr);
      return -2;
    }
    m = it->second;
    g_maps.erase(it);
  }
  if (!m.vmm) {
    DTF_CUDA_OK(cudaFree(ptr));
    return 0;
  }
  Drv& d = drv();
  d.Unmap(reinterpret_cast<CUdeviceptr>(ptr), m.size);
  d.AddressFree(reinterpret_cast<CUdeviceptr>(ptr), m.size);
  if (m.owns_handle) d.Release(m.handle);
  return 0;
}

// 1 when slabs are VMM allocations that can be shared as file descriptors
int dtf_symm_is_vmm(void* ptr) {
  std::lock_guard<std::mutex> l(g_mu);
  auto it = g_maps.find(ptr);
  return it != g_maps.end() && it->second.vmm ? 1 : 0;
}

long long dtf_symm_mapped_size(void* ptr) {
  std::lock_guard<std::mutex> l(g_mu);
  auto it = g_maps.find(ptr);
  return it == g_maps.end() ? -1 : static_cast<long long>(it->second.size);
}

// POSIX file descriptor of a VMM slab (caller closes it after it has been sent to the peers)
int dtf_symm_export_fd(void* ptr, int* fd_out) {
  Mapping m;
  {
    std::lock_guard<std::mutex> l(g_mu);
    auto it = g_maps.find(ptr);
    if (it == g_maps.end() || !it->second.vmm) {
      set_error("symm: %p is not a VMM allocation", ptr);
      return -2;
    }
    m = it->second;
  }
  int fd = -1;
  const CUresult r = drv().Export(&fd, m.handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    set_error("symm: cuMemExportToShareableHandle failed (%d)", (int)r);
    return -1;
  }
  *fd_out = fd;
  return 0;
}

// map a peer's slab received as a file descriptor; `size` is dtf_symm_mapped_size() on the owner
int dtf_symm_import_fd(int fd, size_t size, void** out_ptr) {
  int dev = 0;
  DTF_CUDA_OK(cudaGetDevice(&dev));
  DTF_CUDA_OK(cudaFree(nullptr));
  Drv& d = drv();
  if (!d.ok) {
    set_error("symm: VMM driver entry points unavailable");
    return -1;
  }
  CUmemGenericAllocationHandle h = 0;
  const CUresult r = d.Import(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) {
    set_error("symm: cuMemImportFromShareableHandle failed (%d)", (int)r);
    return -1;
  }
  void* p = nullptr;
  if (map_handle(h, size, dev, &p) != 0) {
    d.Release(h);
    return -1;
  }
  std::lock_guard<std::mutex> l(g_mu);
  g_maps[p] = Mapping{h, size, true, true};
  *out_ptr = p;
  return 0;
}

// ---- multicast objects (NVLS: in-switch reduction and broadcast) ---------------------------------
// One multicast object spans the same byte range of every rank's slab.  A multimem.ld_reduce on its
// mapping returns the sum of all ranks' values (reduced inside NVSwitch), a multimem.st writes all
// ranks' copies with one store.  Life cycle (all ranks, with host barriers between the phases):
// rank 0 creates the object and passes its file descriptor around; every rank adds its device; every
// rank binds the leading `bytes` of its slab and maps the object.
namespace {
struct McObject {
  CUmemGenericAllocationHandle handle = 0;
  size_t bytes = 0;
  void* mapping = nullptr;   // this process's mapping of the object (dtf_mc_bind_map)
};
std::map<long long, McObject> g_mc;
long long g_mc_next = 1;

CUmulticastObjectProp mc_prop(int num_devices, size_t bytes) {
  CUmulticastObjectProp prop{};
  prop.numDevices = static_cast<unsigned int>(num_devices);
  prop.size = bytes;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  prop.flags = 0;
  return prop;
}
}  // namespace

// 1 when the current device and driver support multicast objects
int dtf_mc_supported(void) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  (void)cudaFree(nullptr);
  Drv& d = drv();
  if (!d.mc_ok) return 0;
  CUdevice cd;
  int v = 0;
  if (d.DeviceGet(&cd, dev) != CUDA_SUCCESS) return 0;
  if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cd) != CUDA_SUCCESS) return 0;
  return v ? 1 : 0;
}

// largest multiple of the multicast granularity that fits into `bytes` (0: unsupported)
long long dtf_mc_usable_bytes(int num_devices, size_t bytes) {
  Drv& d = drv();
  if (!d.mc_ok) return 0;
  CUmulticastObjectProp prop = mc_prop(num_devices, bytes);
  size_t gran = 0;
  if (d.McGetGranularity(&gran, &prop, CU_MULTICAST_GRANULARITY_MINIMUM) != CUDA_SUCCESS || gran == 0) return 0;
  return static_cast<long long>(bytes / gran * gran);
}

// rank 0: create the object (bytes from dtf_mc_usable_bytes) and export it as a file descriptor
int dtf_mc_create(int num_devices, size_t bytes, long long* id_out, int* fd_out) {
  Drv& d = drv();
  if (!d.mc_ok) {
    set_error("multicast: driver entry points unavailable");
    return -1;
  }
  CUmulticastObjectProp prop = mc_prop(num_devices, bytes);
  CUmemGenericAllocationHandle h = 0;
  CUresult r = d.McCreate(&h, &prop);
  if (r != CUDA_SUCCESS) {
    set_error("multicast: cuMulticastCreate(%d devices, %zu bytes) failed (%d)", num_devices, bytes, (int)r);
    return -1;
  }
  int fd = -1;
  r = d.Export(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    d.Release(h);
    set_error("multicast: export of the object failed (%d)", (int)r);
    return -1;
  }
  std::lock_guard<std::mutex> l(g_mu);
  const long long id = g_mc_next++;
  g_mc[id] = McObject{h, bytes};
  *id_out = id;
  *fd_out = fd;
  return 0;
}

int dtf_mc_import(int fd, size_t bytes, long long* id_out) {
  Drv& d = drv();
  if (!d.mc_ok) {
    set_error("multicast: driver entry points unavailable");
    return -1;
  }
  (void)cudaFree(nullptr);
  CUmemGenericAllocationHandle h = 0;
  const CUresult r = d.Import(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) {
    set_error("multicast: import of the object failed (%d)", (int)r);
    return -1;
  }
  std::lock_guard<std::mutex> l(g_mu);
  const long long id = g_mc_next++;
  g_mc[id] = McObject{h, bytes};
  *id_out = id;
  return 0;
}

int dtf_mc_add_device(long long id) {
  Drv& d = drv();
  McObject o;
  {
    std::lock_guard<std::mutex> l(g_mu);
    auto it = g_mc.find(id);
    if (it == g_mc.end()) {
      set_error("multicast: unknown object %lld", id);
      return -2;
    }
    o = it->second;
  }
  int dev = 0;
  DTF_CUDA_OK(cudaGetDevice(&dev));
  CUdevice cd;
  if (d.DeviceGet(&cd, dev) != CUDA_SUCCESS) {
    set_error("multicast: cuDeviceGet(%d) failed", dev);
    return -1;
  }
  const CUresult r = d.McAddDevice(o.handle, cd);
  if (r != CUDA_SUCCESS) {
    set_error("multicast: cuMulticastAddDevice(device %d) failed (%d)", dev, (int)r);
    return -1;
  }
  return 0;
}

// after every rank has added its device: bind the leading bytes of `slab` (a dtf_symm_alloc VMM
// allocation) and map the object; *mc_ptr_out addresses all ranks' copies at once
int dtf_mc_bind_map(long long id, void* slab, void** mc_ptr_out) {
  Drv& d = drv();
  McObject o;
  Mapping m;
  {
    std::lock_guard<std::mutex> l(g_mu);
    auto it = g_mc.find(id);
    auto im = g_maps.find(slab);
    if (it == g_mc.end() || im == g_maps.end() || !im->second.vmm || im->second.imported) {
      set_error("multicast: unknown object or %p is not a local VMM slab", slab);
      return -2;
    }
    o = it->second;
    m = im->second;
  }
  if (o.bytes > m.size) {
    set_error("multicast: object (%zu bytes) larger than the slab mapping (%zu)", o.bytes, m.size);
    return -2;
  }
  int dev = 0;
  DTF_CUDA_OK(cudaGetDevice(&dev));
  CUresult r = d.McBindMem(o.handle, 0, m.handle, 0, o.bytes, 0);
  if (r != CUDA_SUCCESS) {
    set_error("multicast: cuMulticastBindMem(%zu bytes) failed (%d)", o.bytes, (int)r);
    return -1;
  }
  void* p = nullptr;
  if (map_handle(o.handle, o.bytes, dev, &p) != 0) return -1;
  {
    std::lock_guard<std::mutex> l(g_mu);
    g_maps[p] = Mapping{o.handle, o.bytes, true, true, /*owns_handle=*/false};
    g_mc[id].mapping = p;
  }
  *mc_ptr_out = p;
  return 0;
}

// Tear one multicast object down exactly once: unmap this process's view of it (if it was mapped) and
// release the object handle.  The slabs bound to it stay valid; the driver drops the bindings with the
// last reference to the object.
int dtf_mc_destroy(long long id) {
  McObject o;
  {
    std::lock_guard<std::mutex> l(g_mu);
    auto it = g_mc.find(id);
    if (it == g_mc.end()) {
      set_error("multicast: unknown object %lld", id);
      return -2;
    }
    o = it->second;
    g_mc.erase(it);
  }
  if (o.mapping != nullptr) dtf_symm_free(o.mapping);   // unmaps only: the mapping does not own the handle
  drv().Release(o.handle);
  return 0;
}

// ---- legacy CUDA IPC handles (fallback) -------------------------------------------------------
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
  {
    std::lock_guard<std::mutex> l(g_mu);
    if (g_maps.count(ptr)) {
      // imported VMM mapping
    } else {
      DTF_CUDA_OK(cudaIpcCloseMemHandle(ptr));
      return 0;
    }
  }
  return dtf_symm_free(ptr);
}

// Same-process peers (several ranks driven from one process): enable direct peer access.
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
