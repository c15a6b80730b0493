// vmm_arena.cu — the symmetric arena on the CUDA virtual-memory-management API, with the NVLS multicast object bound by us.
//
// One allocation per rank (cuMemCreate, exportable as a POSIX file descriptor), every rank maps every peer's allocation
// into its own address space (cuMemImportFromShareableHandle + cuMemAddressReserve + cuMemMap + cuMemSetAccess), and —
// when the devices support multicast — rank 0 creates a multicast object over all devices (cuMulticastCreate), every rank
// adds its device and binds its allocation (cuMulticastAddDevice / cuMulticastBindMem) and maps the object: stores to that
// mapping land in every GPU's copy, `multimem.ld_reduce` through it is reduced in the switch. The descriptors travel
// between the ranks' processes over unix sockets (SCM_RIGHTS), driven from determined_b200/parallel/vmm.py; this file is
// the device-facing half. It replaces what the first version leaned on torch's symmetric-memory module for (kept as the
// fallback) and what Horovod/NCCL hides behind its communicator in the reference
// (harness/determined/horovod.py, _pytorch_context.py:152-204).
//
// The driver entry points are resolved at run time (cudaGetDriverEntryPoint): the library carries no link-time
// dependency on libcuda, so it still loads on a box without a driver (the CPU test tier).
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <string>

#include "host_utils.h"

namespace det {
namespace {

struct Driver {
  bool ok = false;
  std::string why;
  CUresult (*GetErrorString)(CUresult, const char**) = nullptr;
  CUresult (*DeviceGet)(CUdevice*, int) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
};

template <typename F>
bool resolve(const char* name, F& fn, std::string& why) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || p == nullptr) {
    cudaGetLastError();
    why = std::string("driver entry point ") + name + " is not available";
    return false;
  }
  fn = reinterpret_cast<F>(p);
  return true;
}

Driver& driver() {
  static Driver d = [] {
    Driver x;
    bool ok = resolve("cuGetErrorString", x.GetErrorString, x.why) && resolve("cuDeviceGet", x.DeviceGet, x.why) &&
              resolve("cuDeviceGetAttribute", x.DeviceGetAttribute, x.why) &&
              resolve("cuMemGetAllocationGranularity", x.MemGetAllocationGranularity, x.why) && resolve("cuMemCreate", x.MemCreate, x.why) &&
              resolve("cuMemRelease", x.MemRelease, x.why) && resolve("cuMemExportToShareableHandle", x.MemExportToShareableHandle, x.why) &&
              resolve("cuMemImportFromShareableHandle", x.MemImportFromShareableHandle, x.why) &&
              resolve("cuMemAddressReserve", x.MemAddressReserve, x.why) && resolve("cuMemAddressFree", x.MemAddressFree, x.why) &&
              resolve("cuMemMap", x.MemMap, x.why) && resolve("cuMemUnmap", x.MemUnmap, x.why) && resolve("cuMemSetAccess", x.MemSetAccess, x.why) &&
              resolve("cuMulticastCreate", x.MulticastCreate, x.why) && resolve("cuMulticastAddDevice", x.MulticastAddDevice, x.why) &&
              resolve("cuMulticastBindMem", x.MulticastBindMem, x.why) && resolve("cuMulticastGetGranularity", x.MulticastGetGranularity, x.why);
    x.ok = ok;
    return x;
  }();
  return d;
}

int fail(const char* what, CUresult r) {
  const char* s = nullptr;
  if (driver().GetErrorString) driver().GetErrorString(r, &s);
  set_last_error(std::string(what) + ": " + (s ? s : "unknown driver error") + " (" + std::to_string(static_cast<int>(r)) + ")");
  return 1;
}

#define DET_CU(call, what)                       \
  do {                                           \
    CUresult _r = (call);                        \
    if (_r != CUDA_SUCCESS) return fail(what, _r); \
  } while (0)

int need_driver() {
  if (driver().ok) return 0;
  set_last_error("CUDA driver API unavailable: " + driver().why);
  return 1;
}

CUmemAllocationProp mem_prop(int device) {
  CUmemAllocationProp p = {};
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = device;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

CUmulticastObjectProp mc_prop(int ndev, size_t bytes) {
  CUmulticastObjectProp p = {};
  p.numDevices = static_cast<unsigned int>(ndev);
  p.size = bytes;
  p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  p.flags = 0;
  return p;
}

}  // namespace
}  // namespace det

using namespace det;

// 1 when `device` can take part in a multicast object (NVLS), else 0.
extern "C" int det_vmm_multicast_supported(int device, int* out) {
  *out = 0;
  if (need_driver()) return 1;
  CUdevice dev;
  DET_CU(driver().DeviceGet(&dev, device), "cuDeviceGet");
  int v = 0;
  DET_CU(driver().DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev), "cuDeviceGetAttribute(MULTICAST_SUPPORTED)");
  *out = v;
  return 0;
}

// Size granule for an arena region over `ndev` devices: the larger of the allocation granularity and, with multicast, the
// multicast object's minimum granularity. Callers round region sizes up to a multiple of it.
extern "C" int det_vmm_granularity(int device, int ndev, int with_multicast, long long* out) {
  if (need_driver()) return 1;
  const CUmemAllocationProp p = mem_prop(device);
  size_t g = 0;
  DET_CU(driver().MemGetAllocationGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "cuMemGetAllocationGranularity");
  if (with_multicast) {
    const CUmulticastObjectProp mp = mc_prop(ndev, g);
    size_t mg = 0;
    // the MINIMUM granule (2 MB on Hopper / Blackwell): the recommended one is hundreds of MB, and the arena also holds
    // regions of a few KB (flag tables, metric mailboxes)
    DET_CU(driver().MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_MINIMUM), "cuMulticastGetGranularity");
    if (mg > g) g = mg;
  }
  *out = static_cast<long long>(g);
  return 0;
}

// Physical allocation on `device`, exportable: returns the allocation handle and a file descriptor that names it.
extern "C" int det_vmm_create(int device, long long bytes, unsigned long long* handle, int* fd) {
  if (need_driver()) return 1;
  const CUmemAllocationProp p = mem_prop(device);
  CUmemGenericAllocationHandle h;
  DET_CU(driver().MemCreate(&h, static_cast<size_t>(bytes), &p, 0), "cuMemCreate");
  int f = -1;
  DET_CU(driver().MemExportToShareableHandle(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "cuMemExportToShareableHandle");
  *handle = static_cast<unsigned long long>(h);
  *fd = f;
  return 0;
}

// A peer's allocation (or the multicast object) from its file descriptor.
extern "C" int det_vmm_import(int fd, unsigned long long* handle) {
  if (need_driver()) return 1;
  CUmemGenericAllocationHandle h;
  DET_CU(driver().MemImportFromShareableHandle(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
         "cuMemImportFromShareableHandle");
  *handle = static_cast<unsigned long long>(h);
  return 0;
}

// Map an allocation (local, imported, or a bound multicast object) into this process and open it to `device` for
// reads and writes. The address range belongs to the caller until det_vmm_unmap.
extern "C" int det_vmm_map(unsigned long long handle, long long bytes, long long granularity, int device, void** out) {
  if (need_driver()) return 1;
  CUdeviceptr va = 0;
  DET_CU(driver().MemAddressReserve(&va, static_cast<size_t>(bytes), static_cast<size_t>(granularity), 0, 0), "cuMemAddressReserve");
  CUresult r = driver().MemMap(va, static_cast<size_t>(bytes), 0, static_cast<CUmemGenericAllocationHandle>(handle), 0);
  if (r != CUDA_SUCCESS) {
    driver().MemAddressFree(va, static_cast<size_t>(bytes));
    return fail("cuMemMap", r);
  }
  CUmemAccessDesc acc = {};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = driver().MemSetAccess(va, static_cast<size_t>(bytes), &acc, 1);
  if (r != CUDA_SUCCESS) {
    driver().MemUnmap(va, static_cast<size_t>(bytes));
    driver().MemAddressFree(va, static_cast<size_t>(bytes));
    return fail("cuMemSetAccess", r);
  }
  *out = reinterpret_cast<void*>(va);
  return 0;
}

extern "C" int det_vmm_unmap(void* ptr, long long bytes) {
  if (need_driver()) return 1;
  const CUdeviceptr va = reinterpret_cast<CUdeviceptr>(ptr);
  DET_CU(driver().MemUnmap(va, static_cast<size_t>(bytes)), "cuMemUnmap");
  DET_CU(driver().MemAddressFree(va, static_cast<size_t>(bytes)), "cuMemAddressFree");
  return 0;
}

extern "C" int det_vmm_release(unsigned long long handle) {
  if (need_driver()) return 1;
  DET_CU(driver().MemRelease(static_cast<CUmemGenericAllocationHandle>(handle)), "cuMemRelease");
  return 0;
}

// Rank 0: the multicast object over `ndev` devices and `bytes` per device; fd names it for the other ranks.
extern "C" int det_vmm_multicast_create(int ndev, long long bytes, unsigned long long* handle, int* fd) {
  if (need_driver()) return 1;
  const CUmulticastObjectProp p = mc_prop(ndev, static_cast<size_t>(bytes));
  CUmemGenericAllocationHandle h;
  DET_CU(driver().MulticastCreate(&h, &p), "cuMulticastCreate");
  int f = -1;
  DET_CU(driver().MemExportToShareableHandle(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "cuMemExportToShareableHandle(multicast)");
  *handle = static_cast<unsigned long long>(h);
  *fd = f;
  return 0;
}

// Every rank, once: join the object with its device. All devices must have joined before anyone binds.
extern "C" int det_vmm_multicast_add_device(unsigned long long mc, int device) {
  if (need_driver()) return 1;
  CUdevice dev;
  DET_CU(driver().DeviceGet(&dev, device), "cuDeviceGet");
  DET_CU(driver().MulticastAddDevice(static_cast<CUmemGenericAllocationHandle>(mc), dev), "cuMulticastAddDevice");
  return 0;
}

// Every rank: back the object's range [0, bytes) on its device with its own allocation.
extern "C" int det_vmm_multicast_bind(unsigned long long mc, unsigned long long mem, long long bytes) {
  if (need_driver()) return 1;
  DET_CU(driver().MulticastBindMem(static_cast<CUmemGenericAllocationHandle>(mc), 0, static_cast<CUmemGenericAllocationHandle>(mem), 0,
                                   static_cast<size_t>(bytes), 0),
         "cuMulticastBindMem");
  return 0;
}
