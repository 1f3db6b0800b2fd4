// vmm.cpp -- peer-mappable device memory through the CUDA virtual-memory-management API.
//
// The symmetric buffers of a plan (gradient buckets, mirrored weights, flag words) must
// be loadable/storable by every GPU of the box: one address space for a local context
// (MirroredStrategy's shape, imagenet-resnet50-mirror.py:21), POSIX-fd shareable for rank
// contexts (Horovod / MWMS shape, imagenet-resnet50-hvd.py:16, -multiworkers.py:24-25).
// cuMemCreate allocations are the one kind of memory that serves both (and can later be
// bound to an NVSwitch multicast object).  The driver entry points are resolved at run
// time through the runtime, so the library carries no link-time dependency on libcuda and
// still loads (for the host-only layout/fd code) on a machine without a driver.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>

#include "internal.h"

namespace b2dp {

namespace {

struct Driver {
  bool ok = false;
  int rc = B2DP_ERR_CUDA;
  std::string why;
  CUresult (*GetErrorString)(CUresult, const char**) = nullptr;
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*,
                                          CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*,
                        unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
                                         unsigned long long) = nullptr;
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*,
                                           CUmemAllocationHandleType) = nullptr;
  // NVSwitch multicast (optional: absent on drivers/boxes without NVLS)
  bool mc_ok = false;
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                               unsigned long long) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
  CUresult (*DeviceGet)(CUdevice*, int) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
};

Driver g_drv;
std::once_flag g_once;

template <typename F>
bool resolve(const char* name, F* fn, std::string* why) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
    *why = std::string("cannot resolve driver symbol ") + name + ": " +
           (e != cudaSuccess ? cudaGetErrorString(e) : "not found");
    (void)cudaGetLastError();
    return false;
  }
  *fn = reinterpret_cast<F>(p);
  return true;
}

void init_once() {
  Driver& d = g_drv;
  std::string why;
  bool ok = resolve("cuGetErrorString", &d.GetErrorString, &why) &&
            resolve("cuMemGetAllocationGranularity", &d.MemGetAllocationGranularity, &why) &&
            resolve("cuMemCreate", &d.MemCreate, &why) && resolve("cuMemRelease", &d.MemRelease, &why) &&
            resolve("cuMemAddressReserve", &d.MemAddressReserve, &why) &&
            resolve("cuMemAddressFree", &d.MemAddressFree, &why) && resolve("cuMemMap", &d.MemMap, &why) &&
            resolve("cuMemUnmap", &d.MemUnmap, &why) && resolve("cuMemSetAccess", &d.MemSetAccess, &why) &&
            resolve("cuMemExportToShareableHandle", &d.MemExportToShareableHandle, &why) &&
            resolve("cuMemImportFromShareableHandle", &d.MemImportFromShareableHandle, &why);
  d.ok = ok;
  d.why = why;
  std::string why_mc;
  d.mc_ok = ok && resolve("cuMulticastCreate", &d.MulticastCreate, &why_mc) &&
            resolve("cuMulticastAddDevice", &d.MulticastAddDevice, &why_mc) &&
            resolve("cuMulticastBindMem", &d.MulticastBindMem, &why_mc) &&
            resolve("cuMulticastGetGranularity", &d.MulticastGetGranularity, &why_mc) &&
            resolve("cuDeviceGet", &d.DeviceGet, &why_mc) &&
            resolve("cuDeviceGetAttribute", &d.DeviceGetAttribute, &why_mc);
}

const char* cu_str(CUresult r) {
  const char* s = nullptr;
  if (g_drv.GetErrorString && g_drv.GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
  return "unknown CUresult";
}

#define CU_TRY(call, what)                                                        \
  do {                                                                            \
    CUresult _r = (call);                                                         \
    if (_r != CUDA_SUCCESS) return fail(B2DP_ERR_CUDA, "%s: %s (%d)", what, cu_str(_r), (int)_r); \
  } while (0)

CUmemAllocationProp make_prop(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof prop);
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

int map_and_grant(CUmemGenericAllocationHandle h, size_t size, size_t gran, const int* devs, int ndev,
                  CUdeviceptr* va_out) {
  CUdeviceptr va = 0;
  CU_TRY(g_drv.MemAddressReserve(&va, size, gran, 0, 0), "cuMemAddressReserve");
  CUresult r = g_drv.MemMap(va, size, 0, h, 0);
  if (r != CUDA_SUCCESS) {
    g_drv.MemAddressFree(va, size);
    return fail(B2DP_ERR_CUDA, "cuMemMap: %s (%d)", cu_str(r), (int)r);
  }
  std::vector<CUmemAccessDesc> acc;
  for (int i = 0; i < ndev; ++i) {
    bool seen = false;
    for (auto& a : acc) seen |= (a.location.id == devs[i]);
    if (seen) continue;
    CUmemAccessDesc a;
    memset(&a, 0, sizeof a);
    a.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    a.location.id = devs[i];
    a.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    acc.push_back(a);
  }
  r = g_drv.MemSetAccess(va, size, acc.data(), acc.size());
  if (r != CUDA_SUCCESS) {
    g_drv.MemUnmap(va, size);
    g_drv.MemAddressFree(va, size);
    return fail(B2DP_ERR_CUDA, "cuMemSetAccess (is peer access between the GPUs available?): %s (%d)",
                cu_str(r), (int)r);
  }
  *va_out = va;
  return B2DP_OK;
}

}  // namespace

int vmm_init() {
  std::call_once(g_once, init_once);
  if (!g_drv.ok) return fail(B2DP_ERR_CUDA, "%s", g_drv.why.c_str());
  return B2DP_OK;
}

int vmm_granularity(int device, size_t* gran) {
  if (int rc = vmm_init()) return rc;
  CUmemAllocationProp prop = make_prop(device);
  CU_TRY(g_drv.MemGetAllocationGranularity(gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED),
         "cuMemGetAllocationGranularity");
  return B2DP_OK;
}

int vmm_alloc(int device, size_t size, const int* access_devices, int n_access, VmmBlock* out) {
  if (int rc = vmm_init()) return rc;
  size_t gran = 0;
  if (int rc = vmm_granularity(device, &gran)) return rc;
  size = (size + gran - 1) / gran * gran;
  CUmemAllocationProp prop = make_prop(device);
  CUmemGenericAllocationHandle h = 0;
  CU_TRY(g_drv.MemCreate(&h, size, &prop, 0), "cuMemCreate");
  CUdeviceptr va = 0;
  int rc = map_and_grant(h, size, gran, access_devices, n_access, &va);
  if (rc != B2DP_OK) {
    g_drv.MemRelease(h);
    return rc;
  }
  out->handle = h;
  out->va = va;
  out->size = size;
  out->owner_device = device;
  out->mapped = true;
  return B2DP_OK;
}

int vmm_export_fd(const VmmBlock& b, int* fd) {
  if (int rc = vmm_init()) return rc;
  int f = -1;
  CU_TRY(g_drv.MemExportToShareableHandle(&f, b.handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0),
         "cuMemExportToShareableHandle");
  *fd = f;
  return B2DP_OK;
}

int vmm_import_fd(int fd, size_t size, int device, VmmBlock* out) {
  if (int rc = vmm_init()) return rc;
  size_t gran = 0;
  if (int rc = vmm_granularity(device, &gran)) return rc;
  size = (size + gran - 1) / gran * gran;
  CUmemGenericAllocationHandle h = 0;
  CU_TRY(g_drv.MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
         "cuMemImportFromShareableHandle");
  CUdeviceptr va = 0;
  int rc = map_and_grant(h, size, gran, &device, 1, &va);
  if (rc != B2DP_OK) {
    g_drv.MemRelease(h);
    return rc;
  }
  out->handle = h;
  out->va = va;
  out->size = size;
  out->owner_device = -1;
  out->mapped = true;
  return B2DP_OK;
}

// ---- NVSwitch multicast objects (Path B) ------------------------------------------------
namespace {
CUmulticastObjectProp mc_prop(int ndev, size_t size) {
  CUmulticastObjectProp prop;
  memset(&prop, 0, sizeof prop);
  prop.numDevices = (unsigned)ndev;
  prop.size = size;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}
}  // namespace

int mc_supported(int device) {
  if (int rc = vmm_init()) return rc;
  if (!g_drv.mc_ok) return fail(B2DP_ERR_CUDA, "this driver has no cuMulticast* entry points");
  CUdevice d;
  CU_TRY(g_drv.DeviceGet(&d, device), "cuDeviceGet");
  int v = 0;
  CU_TRY(g_drv.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d), "cuDeviceGetAttribute");
  if (!v) return fail(B2DP_ERR_CUDA, "device %d does not support NVSwitch multicast", device);
  return B2DP_OK;
}

int mc_granularity(int ndev, size_t size, size_t* gran) {
  if (int rc = vmm_init()) return rc;
  if (!g_drv.mc_ok) return fail(B2DP_ERR_CUDA, "this driver has no cuMulticast* entry points");
  CUmulticastObjectProp prop = mc_prop(ndev, size);
  CU_TRY(g_drv.MulticastGetGranularity(gran, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED), "cuMulticastGetGranularity");
  return B2DP_OK;
}

int mc_create(int ndev, size_t size, unsigned long long* handle) {
  if (int rc = vmm_init()) return rc;
  if (!g_drv.mc_ok) return fail(B2DP_ERR_CUDA, "this driver has no cuMulticast* entry points");
  CUmulticastObjectProp prop = mc_prop(ndev, size);
  CUmemGenericAllocationHandle h = 0;
  CU_TRY(g_drv.MulticastCreate(&h, &prop), "cuMulticastCreate");
  *handle = h;
  return B2DP_OK;
}

int mc_export_fd(unsigned long long handle, int* fd) {
  int f = -1;
  CU_TRY(g_drv.MemExportToShareableHandle(&f, handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0),
         "cuMemExportToShareableHandle(multicast)");
  *fd = f;
  return B2DP_OK;
}

int mc_import_fd(int fd, unsigned long long* handle) {
  if (int rc = vmm_init()) return rc;
  CUmemGenericAllocationHandle h = 0;
  CU_TRY(g_drv.MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
         "cuMemImportFromShareableHandle(multicast)");
  *handle = h;
  return B2DP_OK;
}

int mc_add_device(unsigned long long handle, int device) {
  CUdevice d;
  CU_TRY(g_drv.DeviceGet(&d, device), "cuDeviceGet");
  CU_TRY(g_drv.MulticastAddDevice(handle, d), "cuMulticastAddDevice");
  return B2DP_OK;
}

int mc_bind(unsigned long long handle, const VmmBlock& mem, size_t size) {
  CU_TRY(g_drv.MulticastBindMem(handle, 0, mem.handle, 0, size, 0), "cuMulticastBindMem");
  return B2DP_OK;
}

int mc_map(unsigned long long handle, size_t size, const int* devs, int ndev, unsigned long long* va_out) {
  size_t gran = 0;
  if (int rc = vmm_granularity(devs[0], &gran)) return rc;
  CUdeviceptr va = 0;
  if (int rc = map_and_grant(handle, size, gran, devs, ndev, &va)) return rc;
  *va_out = va;
  return B2DP_OK;
}

void mc_release(unsigned long long handle, unsigned long long va, size_t size) {
  if (!g_drv.ok) return;
  if (va) {
    g_drv.MemUnmap(va, size);
    g_drv.MemAddressFree(va, size);
  }
  if (handle) g_drv.MemRelease(handle);
}

void vmm_free(VmmBlock* b) {
  if (!b || !b->mapped || !g_drv.ok) return;
  g_drv.MemUnmap(b->va, b->size);
  g_drv.MemAddressFree(b->va, b->size);
  g_drv.MemRelease(b->handle);
  b->mapped = false;
}

}  // namespace b2dp
