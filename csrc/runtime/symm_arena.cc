// Symmetric-memory arena for the peer-memory collectives (NVLink 5 / NVSwitch).
//
// Each rank creates its arena with the CUDA virtual-memory-management API, exports it as a POSIX
// file descriptor, and maps every peer's arena into its own address space; on NVSwitch systems a
// multicast object is bound over the same physical memory so that `multimem.*` instructions can
// reduce in the switch (NVLS).  The descriptor exchange itself (SCM_RIGHTS over AF_UNIX sockets)
// is done by the Python layer (cloud_b200/parallel/symm.py); this file owns every driver call.
//
// libcuda is loaded with dlopen at first use so the runtime library also loads on CPU-only boxes.
//
// The reference has no communication layer at all: this replaces what
// TensorFlow's NcclAllReduce / CollectiveAllReduce cross-device ops did underneath
// MirroredStrategy()/MultiWorkerMirroredStrategy() (core/preprocess.py:139-143).
#include <cuda.h>
#include <dlfcn.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <unistd.h>

#include <mutex>

#include "rt_common.h"
#include "cloud_b200_runtime.h"   // the public declarations must match the definitions below

namespace {

struct Driver {
  void* lib = nullptr;
  bool ok = false;
#define CB2_DRV(name) decltype(&name) p_##name = nullptr;
  CB2_DRV(cuInit)
  CB2_DRV(cuDeviceGet)
  CB2_DRV(cuDeviceGetAttribute)
  CB2_DRV(cuGetErrorString)
  CB2_DRV(cuMemGetAllocationGranularity)
  CB2_DRV(cuMemCreate)
  CB2_DRV(cuMemRelease)
  CB2_DRV(cuMemAddressReserve)
  CB2_DRV(cuMemAddressFree)
  CB2_DRV(cuMemMap)
  CB2_DRV(cuMemUnmap)
  CB2_DRV(cuMemSetAccess)
  CB2_DRV(cuMemExportToShareableHandle)
  CB2_DRV(cuMemImportFromShareableHandle)
  CB2_DRV(cuMulticastCreate)
  CB2_DRV(cuMulticastAddDevice)
  CB2_DRV(cuMulticastBindMem)
  CB2_DRV(cuMulticastGetGranularity)
  CB2_DRV(cuDevicePrimaryCtxRetain)
  CB2_DRV(cuCtxSetCurrent)
  CB2_DRV(cuCtxGetCurrent)
#undef CB2_DRV
};

Driver g_drv;
std::once_flag g_drv_once;

template <typename T>
bool load_sym(void* lib, const char* name, T* out, bool required) {
  void* s = dlsym(lib, name);
  *out = reinterpret_cast<T>(s);
  return s != nullptr || !required;
}

void load_driver() {
  g_drv.lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
  if (!g_drv.lib) g_drv.lib = dlopen("libcuda.so", RTLD_NOW | RTLD_GLOBAL);
  if (!g_drv.lib) return;
  bool ok = true;
#define CB2_LOAD(name, sym, req) ok &= load_sym(g_drv.lib, sym, &g_drv.p_##name, req);
  CB2_LOAD(cuInit, "cuInit", true)
  CB2_LOAD(cuDeviceGet, "cuDeviceGet", true)
  CB2_LOAD(cuDeviceGetAttribute, "cuDeviceGetAttribute", true)
  CB2_LOAD(cuGetErrorString, "cuGetErrorString", true)
  CB2_LOAD(cuMemGetAllocationGranularity, "cuMemGetAllocationGranularity", true)
  CB2_LOAD(cuMemCreate, "cuMemCreate", true)
  CB2_LOAD(cuMemRelease, "cuMemRelease", true)
  CB2_LOAD(cuMemAddressReserve, "cuMemAddressReserve", true)
  CB2_LOAD(cuMemAddressFree, "cuMemAddressFree", true)
  CB2_LOAD(cuMemMap, "cuMemMap", true)
  CB2_LOAD(cuMemUnmap, "cuMemUnmap", true)
  CB2_LOAD(cuMemSetAccess, "cuMemSetAccess", true)
  CB2_LOAD(cuMemExportToShareableHandle, "cuMemExportToShareableHandle", true)
  CB2_LOAD(cuMemImportFromShareableHandle, "cuMemImportFromShareableHandle", true)
  CB2_LOAD(cuMulticastCreate, "cuMulticastCreate", false)
  CB2_LOAD(cuMulticastAddDevice, "cuMulticastAddDevice", false)
  CB2_LOAD(cuMulticastBindMem, "cuMulticastBindMem", false)
  CB2_LOAD(cuMulticastGetGranularity, "cuMulticastGetGranularity", false)
  CB2_LOAD(cuDevicePrimaryCtxRetain, "cuDevicePrimaryCtxRetain", true)
  CB2_LOAD(cuCtxSetCurrent, "cuCtxSetCurrent", true)
  CB2_LOAD(cuCtxGetCurrent, "cuCtxGetCurrent", true)
#undef CB2_LOAD
  if (ok && g_drv.p_cuInit(0) == CUDA_SUCCESS) g_drv.ok = true;
}

bool driver() {
  std::call_once(g_drv_once, load_driver);
  if (!g_drv.ok) cb2rt::set_error("CUDA driver (libcuda.so.1) not available");
  return g_drv.ok;
}

int fail(CUresult r, const char* what) {
  const char* msg = nullptr;
  if (g_drv.p_cuGetErrorString) g_drv.p_cuGetErrorString(r, &msg);
  cb2rt::set_error("%s failed: CUDA driver error %d (%s)", what, static_cast<int>(r),
                   msg ? msg : "?");
  return 2;
}

#define DRV_OK(call, what)                       \
  do {                                           \
    CUresult _r = (call);                        \
    if (_r != CUDA_SUCCESS) return fail(_r, what); \
  } while (0)

int ensure_ctx(int dev) {
  CUcontext cur = nullptr;
  g_drv.p_cuCtxGetCurrent(&cur);
  if (cur) return 0;
  CUdevice d;
  DRV_OK(g_drv.p_cuDeviceGet(&d, dev), "cuDeviceGet");
  CUcontext ctx;
  DRV_OK(g_drv.p_cuDevicePrimaryCtxRetain(&ctx, d), "cuDevicePrimaryCtxRetain");
  DRV_OK(g_drv.p_cuCtxSetCurrent(ctx), "cuCtxSetCurrent");
  return 0;
}

CUmemAllocationProp alloc_prop(int dev) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = dev;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

}  // namespace

extern "C" {

// vmm: device supports VMM + POSIX fd handles; multicast: device supports NVLS multicast objects.
int cb2rt_symm_capabilities(int dev, int* vmm, int* multicast) {
  *vmm = 0;
  *multicast = 0;
  if (!driver()) return 1;
  CUdevice d;
  DRV_OK(g_drv.p_cuDeviceGet(&d, dev), "cuDeviceGet");
  int v = 0, h = 0, m = 0;
  g_drv.p_cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, d);
  g_drv.p_cuDeviceGetAttribute(&h, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, d);
  if (g_drv.p_cuMulticastCreate)
    g_drv.p_cuDeviceGetAttribute(&m, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d);
  *vmm = (v && h) ? 1 : 0;
  *multicast = m ? 1 : 0;
  return 0;
}

// Granularity every arena size must be rounded to (max of allocation and multicast minimum).
int cb2rt_symm_granularity(int dev, int world, size_t* out) {
  if (!driver()) return 1;
  CUmemAllocationProp prop = alloc_prop(dev);
  size_t g = 0;
  DRV_OK(g_drv.p_cuMemGetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED),
         "cuMemGetAllocationGranularity");
  if (g_drv.p_cuMulticastGetGranularity && world > 1) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = world;
    mp.size = g;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (g_drv.p_cuMulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_MINIMUM) ==
            CUDA_SUCCESS && mg > g)
      g = mg;
  }
  *out = g;
  return 0;
}

// Create `bytes` (multiple of the granularity) of device memory and export it as an fd.
int cb2rt_symm_create(int dev, size_t bytes, uint64_t* handle_out, int* fd_out) {
  if (!driver() || ensure_ctx(dev)) return 1;
  CUmemAllocationProp prop = alloc_prop(dev);
  CUmemGenericAllocationHandle h;
  DRV_OK(g_drv.p_cuMemCreate(&h, bytes, &prop, 0), "cuMemCreate");
  int fd = -1;
  CUresult r = g_drv.p_cuMemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    g_drv.p_cuMemRelease(h);
    return fail(r, "cuMemExportToShareableHandle");
  }
  *handle_out = static_cast<uint64_t>(h);
  *fd_out = fd;
  return 0;
}

// Import a peer's (or the multicast object's) fd.  The fd stays owned by the caller.
int cb2rt_symm_import(int fd, uint64_t* handle_out) {
  if (!driver()) return 1;
  CUmemGenericAllocationHandle h;
  DRV_OK(g_drv.p_cuMemImportFromShareableHandle(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                                CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
         "cuMemImportFromShareableHandle");
  *handle_out = static_cast<uint64_t>(h);
  return 0;
}

// Reserve address space, map the handle and enable read/write from `dev`.
int cb2rt_symm_map(uint64_t handle, int dev, size_t bytes, size_t align, uint64_t* ptr_out) {
  if (!driver() || ensure_ctx(dev)) return 1;
  CUdeviceptr ptr = 0;
  DRV_OK(g_drv.p_cuMemAddressReserve(&ptr, bytes, align, 0, 0), "cuMemAddressReserve");
  CUresult r = g_drv.p_cuMemMap(ptr, bytes, 0, static_cast<CUmemGenericAllocationHandle>(handle), 0);
  if (r != CUDA_SUCCESS) {
    g_drv.p_cuMemAddressFree(ptr, bytes);
    return fail(r, "cuMemMap");
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = g_drv.p_cuMemSetAccess(ptr, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) {
    g_drv.p_cuMemUnmap(ptr, bytes);
    g_drv.p_cuMemAddressFree(ptr, bytes);
    return fail(r, "cuMemSetAccess");
  }
  *ptr_out = static_cast<uint64_t>(ptr);
  return 0;
}

int cb2rt_symm_unmap(uint64_t ptr, size_t bytes) {
  if (!driver()) return 1;
  DRV_OK(g_drv.p_cuMemUnmap(static_cast<CUdeviceptr>(ptr), bytes), "cuMemUnmap");
  DRV_OK(g_drv.p_cuMemAddressFree(static_cast<CUdeviceptr>(ptr), bytes), "cuMemAddressFree");
  return 0;
}

int cb2rt_symm_release(uint64_t handle) {
  if (!driver()) return 1;
  DRV_OK(g_drv.p_cuMemRelease(static_cast<CUmemGenericAllocationHandle>(handle)), "cuMemRelease");
  return 0;
}

// Rank 0: create the multicast object for `world` devices and export it.
int cb2rt_mc_create(int world, size_t bytes, uint64_t* handle_out, int* fd_out) {
  if (!driver()) return 1;
  if (!g_drv.p_cuMulticastCreate) {
    cb2rt::set_error("driver has no multicast support");
    return 3;
  }
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = world;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  DRV_OK(g_drv.p_cuMulticastCreate(&h, &mp), "cuMulticastCreate");
  int fd = -1;
  CUresult r = g_drv.p_cuMemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    g_drv.p_cuMemRelease(h);
    return fail(r, "cuMemExportToShareableHandle(multicast)");
  }
  *handle_out = static_cast<uint64_t>(h);
  *fd_out = fd;
  return 0;
}

int cb2rt_mc_add_device(uint64_t mc, int dev) {
  if (!driver() || ensure_ctx(dev)) return 1;
  CUdevice d;
  DRV_OK(g_drv.p_cuDeviceGet(&d, dev), "cuDeviceGet");
  DRV_OK(g_drv.p_cuMulticastAddDevice(static_cast<CUmemGenericAllocationHandle>(mc), d),
         "cuMulticastAddDevice");
  return 0;
}

// Bind this rank's physical allocation at offset 0 of the multicast object (after ALL ranks have
// added their device).
int cb2rt_mc_bind(uint64_t mc, uint64_t mem, size_t bytes) {
  if (!driver()) return 1;
  DRV_OK(g_drv.p_cuMulticastBindMem(static_cast<CUmemGenericAllocationHandle>(mc), 0,
                                    static_cast<CUmemGenericAllocationHandle>(mem), 0, bytes, 0),
         "cuMulticastBindMem");
  return 0;
}

int cb2rt_close_fd(int fd) { return close(fd); }

}  // extern "C"
