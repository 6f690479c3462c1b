// Symmetric-memory arena on the CUDA virtual-memory-management (VMM) driver API.
//
// Stands in for Horovod's fusion_buffer_manager.cc (one persistent buffer per device) - but
// the buffer is exportable: each rank cuMemCreate()s its arena with a POSIX-fd shareable
// handle, the fds travel to the peers (SCM_RIGHTS over a unix socket, done in
// parallel/peer.py), and every rank maps every peer's arena into its own address space.
// Where the box exposes NVLS, a multicast object is bound over all ranks' arenas so one
// multimem.ld_reduce / multimem.st reaches every copy through the switch.
//
// The driver is resolved with dlopen at first use so libhvd_core.so loads on a CPU-only
// host (the build box has only the libcuda stub).  C ABI; declared in include/hvd_b200.h.

#include <cuda.h>
#include <dlfcn.h>
#include <unistd.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>

void hvd_logf(int lvl, const char* fmt, ...);   // diag.cpp

namespace {

thread_local std::string g_symm_err;

struct Driver {
  void* lib = nullptr;
  bool ok = false;
#define HVD_DRV(name) decltype(&::name) p_##name = nullptr;
  HVD_DRV(cuInit)
  HVD_DRV(cuDeviceGet)
  HVD_DRV(cuDeviceGetAttribute)
  HVD_DRV(cuCtxGetCurrent)
  HVD_DRV(cuCtxSetCurrent)
  HVD_DRV(cuDevicePrimaryCtxRetain)
  HVD_DRV(cuGetErrorString)
  HVD_DRV(cuMemGetAllocationGranularity)
  HVD_DRV(cuMemCreate)
  HVD_DRV(cuMemRelease)
  HVD_DRV(cuMemExportToShareableHandle)
  HVD_DRV(cuMemImportFromShareableHandle)
  HVD_DRV(cuMemAddressReserve)
  HVD_DRV(cuMemAddressFree)
  HVD_DRV(cuMemMap)
  HVD_DRV(cuMemUnmap)
  HVD_DRV(cuMemSetAccess)
  HVD_DRV(cuMulticastCreate)
  HVD_DRV(cuMulticastAddDevice)
  HVD_DRV(cuMulticastBindMem)
  HVD_DRV(cuMulticastGetGranularity)
#undef HVD_DRV
};

Driver g_drv;
std::once_flag g_once;

void load_driver() {
  const char* names[] = {"libcuda.so.1", "libcuda.so"};
  for (const char* n : names) {
    g_drv.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (g_drv.lib) break;
  }
  if (!g_drv.lib) return;
  bool all = true;
  // cuda.h #defines some of these names to versioned symbols (e.g. cuCtxSetCurrent stays,
  // cuDevicePrimaryCtxRetain stays); stringify AFTER macro expansion so dlsym gets the real one.
#define HVD_STR2(x) #x
#define HVD_STR(x) HVD_STR2(x)
#define HVD_LOAD(name)                                                              \
  g_drv.p_##name = (decltype(g_drv.p_##name))dlsym(g_drv.lib, HVD_STR(name));       \
  if (!g_drv.p_##name) all = false;
  HVD_LOAD(cuInit)
  HVD_LOAD(cuDeviceGet)
  HVD_LOAD(cuDeviceGetAttribute)
  HVD_LOAD(cuCtxGetCurrent)
  HVD_LOAD(cuCtxSetCurrent)
  HVD_LOAD(cuDevicePrimaryCtxRetain)
  HVD_LOAD(cuGetErrorString)
  HVD_LOAD(cuMemGetAllocationGranularity)
  HVD_LOAD(cuMemCreate)
  HVD_LOAD(cuMemRelease)
  HVD_LOAD(cuMemExportToShareableHandle)
  HVD_LOAD(cuMemImportFromShareableHandle)
  HVD_LOAD(cuMemAddressReserve)
  HVD_LOAD(cuMemAddressFree)
  HVD_LOAD(cuMemMap)
  HVD_LOAD(cuMemUnmap)
  HVD_LOAD(cuMemSetAccess)
  HVD_LOAD(cuMulticastCreate)
  HVD_LOAD(cuMulticastAddDevice)
  HVD_LOAD(cuMulticastBindMem)
  HVD_LOAD(cuMulticastGetGranularity)
#undef HVD_LOAD
  g_drv.ok = all;
}

bool drv() {
  std::call_once(g_once, load_driver);
  if (!g_drv.ok) g_symm_err = "CUDA driver (libcuda.so.1) not available";
  return g_drv.ok;
}

int fail(const char* what, CUresult r) {
  const char* s = nullptr;
  if (g_drv.p_cuGetErrorString) g_drv.p_cuGetErrorString(r, &s);
  g_symm_err = std::string(what) + ": " + (s ? s : "unknown CUDA error") + " (" + std::to_string((int)r) + ")";
  hvd_logf(1 /*DEBUG: callers fall back to CUDA IPC and report at their own level*/, "symm: %s", g_symm_err.c_str());
  return -(int)r - 1000;
}

#define CK(call)                                   \
  do {                                             \
    CUresult _r = g_drv.p_##call;                  \
    if (_r != CUDA_SUCCESS) return fail(#call, _r); \
  } while (0)

int ensure_ctx(int device) {
  CUcontext cur = nullptr;
  CK(cuCtxGetCurrent(&cur));
  if (cur) return 0;
  CK(cuInit(0));
  CUdevice dev;
  CK(cuDeviceGet(&dev, device));
  CUcontext ctx;
  CK(cuDevicePrimaryCtxRetain(&ctx, dev));
  CK(cuCtxSetCurrent(ctx));
  return 0;
}

CUmemAllocationProp alloc_prop(int device) {
  CUmemAllocationProp p;
  std::memset(&p, 0, sizeof(p));
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = device;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

}  // namespace

extern "C" {

const char* hvd_symm_last_error() { return g_symm_err.c_str(); }

int hvd_symm_available() { return drv() ? 1 : 0; }

// Multicast (NVLS) support of `device`: 1 yes, 0 no, <0 error.
int hvd_symm_multicast_supported(int device) {
  if (!drv()) return -1;
  int rc = ensure_ctx(device);
  if (rc) return rc;
  CUdevice dev;
  CK(cuDeviceGet(&dev, device));
  int v = 0;
  CK(cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev));
  return v;
}

int hvd_symm_handle_type_supported(int device) {
  if (!drv()) return -1;
  int rc = ensure_ctx(device);
  if (rc) return rc;
  CUdevice dev;
  CK(cuDeviceGet(&dev, device));
  int v = 0;
  CK(cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev));
  return v;
}

// Round `bytes` up to the allocation (and, if requested, multicast) granularity.
long long hvd_symm_round_size(int device, long long bytes, int nranks_for_multicast) {
  if (!drv()) return -1;
  if (ensure_ctx(device)) return -1;
  CUmemAllocationProp p = alloc_prop(device);
  size_t g = 0;
  if (g_drv.p_cuMemGetAllocationGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || g == 0)
    g = 2u << 20;
  if (nranks_for_multicast > 1) {
    CUmulticastObjectProp mp;
    std::memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)nranks_for_multicast;
    mp.size = (size_t)bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (g_drv.p_cuMulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > g) g = mg;
  }
  return (long long)(((size_t)bytes + g - 1) / g * g);
}

// Allocate physical memory on `device`; returns the generic handle and an exportable fd.
int hvd_symm_create(int device, long long bytes, unsigned long long* handle, int* fd) {
  if (!drv()) return -1;
  int rc = ensure_ctx(device);
  if (rc) return rc;
  CUmemAllocationProp p = alloc_prop(device);
  CUmemGenericAllocationHandle h;
  CK(cuMemCreate(&h, (size_t)bytes, &p, 0));
  int f = -1;
  CUresult r = g_drv.p_cuMemExportToShareableHandle(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) { g_drv.p_cuMemRelease(h); return fail("cuMemExportToShareableHandle", r); }
  *handle = (unsigned long long)h;
  *fd = f;
  return 0;
}

int hvd_symm_import(int fd, unsigned long long* handle) {
  if (!drv()) return -1;
  CUmemGenericAllocationHandle h;
  CK(cuMemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *handle = (unsigned long long)h;
  return 0;
}

// Reserve VA, map `handle` and grant `device` read/write access.  Returns the device pointer.
int hvd_symm_map(int device, unsigned long long handle, long long bytes, unsigned long long* ptr) {
  if (!drv()) return -1;
  CUdeviceptr va = 0;
  CK(cuMemAddressReserve(&va, (size_t)bytes, 0, 0, 0));
  CUresult r = g_drv.p_cuMemMap(va, (size_t)bytes, 0, (CUmemGenericAllocationHandle)handle, 0);
  if (r != CUDA_SUCCESS) { g_drv.p_cuMemAddressFree(va, (size_t)bytes); return fail("cuMemMap", r); }
  CUmemAccessDesc d;
  std::memset(&d, 0, sizeof(d));
  d.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  d.location.id = device;
  d.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = g_drv.p_cuMemSetAccess(va, (size_t)bytes, &d, 1);
  if (r != CUDA_SUCCESS) {
    g_drv.p_cuMemUnmap(va, (size_t)bytes);
    g_drv.p_cuMemAddressFree(va, (size_t)bytes);
    return fail("cuMemSetAccess", r);
  }
  *ptr = (unsigned long long)va;
  return 0;
}

int hvd_symm_unmap(unsigned long long ptr, long long bytes) {
  if (!drv()) return -1;
  CK(cuMemUnmap((CUdeviceptr)ptr, (size_t)bytes));
  CK(cuMemAddressFree((CUdeviceptr)ptr, (size_t)bytes));
  return 0;
}

int hvd_symm_release(unsigned long long handle) {
  if (!drv()) return -1;
  CK(cuMemRelease((CUmemGenericAllocationHandle)handle));
  return 0;
}

// ---- multicast (NVLS) -------------------------------------------------------------------------
int hvd_symm_mc_create(int ndev, long long bytes, unsigned long long* mc, int* fd) {
  if (!drv()) return -1;
  CUmulticastObjectProp mp;
  std::memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)ndev;
  mp.size = (size_t)bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  CK(cuMulticastCreate(&h, &mp));
  int f = -1;
  CUresult r = g_drv.p_cuMemExportToShareableHandle(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) { g_drv.p_cuMemRelease(h); return fail("cuMemExportToShareableHandle(mc)", r); }
  *mc = (unsigned long long)h;
  *fd = f;
  return 0;
}

int hvd_symm_mc_add_device(unsigned long long mc, int device) {
  if (!drv()) return -1;
  CUdevice dev;
  CK(cuDeviceGet(&dev, device));
  CK(cuMulticastAddDevice((CUmemGenericAllocationHandle)mc, dev));
  return 0;
}

// Bind this rank's physical allocation at offset 0 of the multicast object.  Every rank must
// have called mc_add_device before ANY rank binds (callers barrier in between).
int hvd_symm_mc_bind(unsigned long long mc, unsigned long long mem, long long bytes) {
  if (!drv()) return -1;
  CK(cuMulticastBindMem((CUmemGenericAllocationHandle)mc, 0, (CUmemGenericAllocationHandle)mem, 0, (size_t)bytes, 0));
  return 0;
}

int hvd_symm_close_fd(int fd) { return close(fd); }

}  // extern "C"
