// Host-side runtime of libkubedl_b200: error reporting and the symmetric-memory plumbing used
// by the fused compute+collective kernels.  Two ways to get a peer-mapped
// heap: (a) cudaMalloc + legacy CUDA IPC handles, (b) cuMemCreate (VMM) + POSIX file
// descriptors, which additionally allows binding the heap to an NVLS multicast object so that
// kernels can use multimem.ld_reduce / multimem.st.  Driver entry points are resolved at run
// time (cudaGetDriverEntryPoint) so that the library loads on machines without libcuda.so.1.
// The reference operator has no communication backend (reference
// pkg/job_controller/service.go:237-295 only creates the headless Services peers resolve).
#include "common.cuh"
#include "../include/kubedl_b200.h"
#include <cuda.h>
#include <stdarg.h>
#include <string.h>
#include <unistd.h>

static thread_local char g_err[1024] = "";

void kdl_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* kdl_last_error(void) { return g_err; }
extern "C" int kdl_version(void) { return 100; }

// ---------------------------------------------------------------- legacy IPC path
extern "C" int kdl_symm_malloc(size_t bytes, void** ptr) {
  KDL_CUDA(cudaMalloc(ptr, bytes));
  KDL_CUDA(cudaMemset(*ptr, 0, bytes));
  return KDL_OK;
}
extern "C" int kdl_symm_free(void* ptr) {
  KDL_CUDA(cudaFree(ptr));
  return KDL_OK;
}
extern "C" int kdl_ipc_get_handle(void* ptr, void* handle64_out) {
  cudaIpcMemHandle_t h;
  KDL_CUDA(cudaIpcGetMemHandle(&h, ptr));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64_out, &h, 64);
  return KDL_OK;
}
extern "C" int kdl_ipc_open_handle(const void* handle64, void** ptr_out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  KDL_CUDA(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
  return KDL_OK;
}
extern "C" int kdl_ipc_close_handle(void* ptr) {
  KDL_CUDA(cudaIpcCloseMemHandle(ptr));
  return KDL_OK;
}
extern "C" int kdl_enable_peer_access(int peer_device) {
  int cur = 0;
  KDL_CUDA(cudaGetDevice(&cur));
  if (cur == peer_device) return KDL_OK;
  int can = 0;
  KDL_CUDA(cudaDeviceCanAccessPeer(&can, cur, peer_device));
  if (!can) {
    kdl_set_error("device %d cannot access peer %d", cur, peer_device);
    return KDL_ERR_UNSUPPORTED;
  }
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return KDL_OK;
  }
  KDL_CUDA(e);
  return KDL_OK;
}

// ---------------------------------------------------------------- VMM + multicast path
namespace {

template <typename F>
bool resolve(const char* name, F* fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || p == nullptr) {
    cudaGetLastError();
    kdl_set_error("driver entry point %s unavailable", name);
    return false;
  }
  *fn = reinterpret_cast<F>(p);
  return true;
}

struct Drv {
  CUresult (*GetErrorString)(CUresult, const char**) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*,
                                          CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*,
                        unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr,
                                unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle,
                     unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle,
                                         CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*,
                                           CUmemAllocationHandleType) = nullptr;
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle,
                               size_t, size_t, unsigned long long) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*,
                                      CUmulticastGranularity_flags) = nullptr;
  bool ok = false;
  bool mc_ok = false;
};

Drv* drv() {
  static Drv d;
  static bool tried = false;
  if (!tried) {
    tried = true;
    cudaFree(0);  // make sure a context exists
    d.ok = resolve("cuGetErrorString", &d.GetErrorString) &&
           resolve("cuDeviceGetAttribute", &d.DeviceGetAttribute) &&
           resolve("cuMemGetAllocationGranularity", &d.MemGetAllocationGranularity) &&
           resolve("cuMemCreate", &d.MemCreate) && resolve("cuMemRelease", &d.MemRelease) &&
           resolve("cuMemAddressReserve", &d.MemAddressReserve) &&
           resolve("cuMemAddressFree", &d.MemAddressFree) && resolve("cuMemMap", &d.MemMap) &&
           resolve("cuMemUnmap", &d.MemUnmap) && resolve("cuMemSetAccess", &d.MemSetAccess) &&
           resolve("cuMemExportToShareableHandle", &d.MemExportToShareableHandle) &&
           resolve("cuMemImportFromShareableHandle", &d.MemImportFromShareableHandle);
    if (d.ok)
      d.mc_ok = resolve("cuMulticastCreate", &d.MulticastCreate) &&
                resolve("cuMulticastAddDevice", &d.MulticastAddDevice) &&
                resolve("cuMulticastBindMem", &d.MulticastBindMem) &&
                resolve("cuMulticastGetGranularity", &d.MulticastGetGranularity);
  }
  return &d;
}

#define KDL_DRV(d, call)                                                   \
  do {                                                                     \
    CUresult r__ = (call);                                                 \
    if (r__ != CUDA_SUCCESS) {                                             \
      const char* s__ = "?";                                               \
      if ((d)->GetErrorString) (d)->GetErrorString(r__, &s__);             \
      kdl_set_error("%s:%d %s: %s", __FILE__, __LINE__, #call, s__);       \
      return KDL_ERR_CUDA;                                                 \
    }                                                                      \
  } while (0)

CUmemAllocationProp make_prop(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

int map_rw(Drv* d, int device, CUmemGenericAllocationHandle h, size_t bytes, void** ptr_out) {
  CUdeviceptr va = 0;
  size_t gran = kdl_vmm_granularity(device, 0);
  KDL_DRV(d, d->MemAddressReserve(&va, bytes, gran, 0, 0));
  KDL_DRV(d, d->MemMap(va, bytes, 0, h, 0));
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  KDL_DRV(d, d->MemSetAccess(va, bytes, &acc, 1));
  *ptr_out = (void*)va;
  return KDL_OK;
}

}  // namespace

extern "C" int kdl_vmm_supported(int device, int* posix_fd_ok, int* multicast_ok) {
  Drv* d = drv();
  *posix_fd_ok = 0;
  *multicast_ok = 0;
  if (!d->ok) return KDL_OK;
  int v = 0;
  if (d->DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED,
                            device) == CUDA_SUCCESS)
    *posix_fd_ok = v;
  v = 0;
  if (d->mc_ok && d->DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device) ==
                      CUDA_SUCCESS)
    *multicast_ok = v;
  return KDL_OK;
}

extern "C" size_t kdl_vmm_granularity(int device, int multicast_world) {
  Drv* d = drv();
  if (!d->ok) return 0;
  size_t gran = 0;
  CUmemAllocationProp prop = make_prop(device);
  if (d->MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) !=
      CUDA_SUCCESS)
    return 0;
  if (multicast_world > 0 && d->mc_ok) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = multicast_world;
    mp.size = gran;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (d->MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) ==
            CUDA_SUCCESS &&
        mg > gran)
      gran = mg;
  }
  return gran;
}

extern "C" int kdl_vmm_alloc(int device, size_t bytes, unsigned long long* handle_out,
                             void** ptr_out, int* fd_out) {
  Drv* d = drv();
  if (!d->ok) return KDL_ERR_UNSUPPORTED;
  CUmemAllocationProp prop = make_prop(device);
  CUmemGenericAllocationHandle h;
  KDL_DRV(d, d->MemCreate(&h, bytes, &prop, 0));
  int rc = map_rw(d, device, h, bytes, ptr_out);
  if (rc) return rc;
  KDL_CUDA(cudaMemset(*ptr_out, 0, bytes));
  int fd = -1;
  KDL_DRV(d, d->MemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle_out = (unsigned long long)h;
  *fd_out = fd;
  return KDL_OK;
}

extern "C" int kdl_vmm_import(int device, int fd, size_t bytes, unsigned long long* handle_out,
                              void** ptr_out) {
  Drv* d = drv();
  if (!d->ok) return KDL_ERR_UNSUPPORTED;
  CUmemGenericAllocationHandle h;
  KDL_DRV(d, d->MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd,
                                             CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  int rc = map_rw(d, device, h, bytes, ptr_out);
  if (rc) return rc;
  *handle_out = (unsigned long long)h;
  return KDL_OK;
}

extern "C" int kdl_vmm_release(unsigned long long handle, void* ptr, size_t bytes) {
  Drv* d = drv();
  if (!d->ok) return KDL_ERR_UNSUPPORTED;
  if (ptr) {
    KDL_DRV(d, d->MemUnmap((CUdeviceptr)ptr, bytes));
    KDL_DRV(d, d->MemAddressFree((CUdeviceptr)ptr, bytes));
  }
  if (handle) KDL_DRV(d, d->MemRelease((CUmemGenericAllocationHandle)handle));
  return KDL_OK;
}

extern "C" int kdl_mc_create(int world, size_t bytes, unsigned long long* mc_out, int* fd_out) {
  Drv* d = drv();
  if (!d->ok || !d->mc_ok) return KDL_ERR_UNSUPPORTED;
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = world;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle mc;
  KDL_DRV(d, d->MulticastCreate(&mc, &mp));
  int fd = -1;
  KDL_DRV(d, d->MemExportToShareableHandle(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *mc_out = (unsigned long long)mc;
  *fd_out = fd;
  return KDL_OK;
}

extern "C" int kdl_mc_import(int fd, unsigned long long* mc_out) {
  Drv* d = drv();
  if (!d->ok || !d->mc_ok) return KDL_ERR_UNSUPPORTED;
  CUmemGenericAllocationHandle mc;
  KDL_DRV(d, d->MemImportFromShareableHandle(&mc, (void*)(uintptr_t)fd,
                                             CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *mc_out = (unsigned long long)mc;
  return KDL_OK;
}

extern "C" int kdl_mc_add_device(unsigned long long mc, int device) {
  Drv* d = drv();
  if (!d->ok || !d->mc_ok) return KDL_ERR_UNSUPPORTED;
  KDL_DRV(d, d->MulticastAddDevice((CUmemGenericAllocationHandle)mc, (CUdevice)device));
  return KDL_OK;
}

extern "C" int kdl_mc_bind_and_map(unsigned long long mc, unsigned long long mem_handle,
                                   int device, size_t bytes, void** mc_ptr_out) {
  Drv* d = drv();
  if (!d->ok || !d->mc_ok) return KDL_ERR_UNSUPPORTED;
  KDL_DRV(d, d->MulticastBindMem((CUmemGenericAllocationHandle)mc, 0,
                                 (CUmemGenericAllocationHandle)mem_handle, 0, bytes, 0));
  return map_rw(d, device, (CUmemGenericAllocationHandle)mc, bytes, mc_ptr_out);
}

extern "C" int kdl_mc_release(unsigned long long mc, void* mc_ptr, size_t bytes) {
  return kdl_vmm_release(mc, mc_ptr, bytes);
}
