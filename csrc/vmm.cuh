// Symmetric memory through the CUDA virtual-memory-management API, so that a buffer can be bound to an NVSwitch
// multicast object (NVLS): physical memory from cuMemCreate, shared between the ranks' processes as POSIX file
// descriptors (the descriptors travel over a unix socket, done in Python: parallel/comm.py), mapped with
// cuMemMap / cuMemSetAccess; a multicast object created by rank 0, every device added, every rank's allocation bound,
// and the multicast handle mapped into the address space as a second view of the same range.  Loads through that view
// (multimem.ld_reduce) return the sum over all ranks computed inside the switch; stores (multimem.st) land on all ranks.
//
// Driver entry points are fetched at run time (cudaGetDriverEntryPoint), so the library still links without libcuda.
// Everything here is host code.  Stands in for the transport layer under the reference's torch.distributed calls
// (c10d ProcessGroupGloo, reference main_gather.py:107); cudaIpc (dpb_symm_*) stays as the fallback where the VMM path
// or multicast is not available.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <mutex>
#include "errors.cuh"

namespace dpb {

struct VmmApi {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*DeviceGet)(CUdevice*, int);
  bool ok;
};

inline const VmmApi& vmm_api() {
  static VmmApi api{};
  static std::once_flag once;
  std::call_once(once, [] {
    bool ok = true;
    auto get = [&](const char* name, void** fn) {
      cudaDriverEntryPointQueryResult q;
      if (cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
          *fn == nullptr)
        ok = false;
    };
    get("cuMemCreate", (void**)&api.MemCreate);
    get("cuMemRelease", (void**)&api.MemRelease);
    get("cuMemAddressReserve", (void**)&api.MemAddressReserve);
    get("cuMemAddressFree", (void**)&api.MemAddressFree);
    get("cuMemMap", (void**)&api.MemMap);
    get("cuMemUnmap", (void**)&api.MemUnmap);
    get("cuMemSetAccess", (void**)&api.MemSetAccess);
    get("cuMemExportToShareableHandle", (void**)&api.MemExportToShareableHandle);
    get("cuMemImportFromShareableHandle", (void**)&api.MemImportFromShareableHandle);
    get("cuMemGetAllocationGranularity", (void**)&api.MemGetAllocationGranularity);
    get("cuMulticastCreate", (void**)&api.MulticastCreate);
    get("cuMulticastAddDevice", (void**)&api.MulticastAddDevice);
    get("cuMulticastBindMem", (void**)&api.MulticastBindMem);
    get("cuMulticastUnbind", (void**)&api.MulticastUnbind);
    get("cuMulticastGetGranularity", (void**)&api.MulticastGetGranularity);
    get("cuDeviceGetAttribute", (void**)&api.DeviceGetAttribute);
    get("cuDeviceGet", (void**)&api.DeviceGet);
    api.ok = ok;
  });
  return api;
}

#define DPB_CU(expr)                                                                                   \
  do {                                                                                                 \
    CUresult _r = (expr);                                                                              \
    if (_r != CUDA_SUCCESS) return ::dpb::set_error(-50, "%s failed: CUresult %d (%s:%d)", #expr, (int)_r, __FILE__, __LINE__); \
  } while (0)

inline int vmm_current_device(CUdevice* dev, int* ordinal) {
  int d = 0;
  DPB_CUDA(cudaGetDevice(&d));
  DPB_CUDA(cudaFree(nullptr));  // make sure the primary context exists
  const VmmApi& a = vmm_api();
  if (!a.ok) return set_error(-51, "CUDA driver entry points for virtual memory management are unavailable");
  DPB_CU(a.DeviceGet(dev, d));
  if (ordinal) *ordinal = d;
  return 0;
}

inline CUmemAllocationProp vmm_prop(int ordinal) {
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = ordinal;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

// out3: [0] VMM + POSIX-fd handles supported, [1] multicast supported, [2] allocation granularity in bytes (the larger
// of the memory and the multicast granularity for `ndev` devices)
inline int vmm_probe(int ndev, long long* out3) {
  out3[0] = out3[1] = out3[2] = 0;
  CUdevice dev;
  int ord = 0;
  if (int rc = vmm_current_device(&dev, &ord)) return rc;
  const VmmApi& a = vmm_api();
  int v = 0, fd = 0, mc = 0;
  DPB_CU(a.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev));
  DPB_CU(a.DeviceGetAttribute(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev));
  if (a.DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) mc = 0;
  out3[0] = (v && fd) ? 1 : 0;
  out3[1] = mc ? 1 : 0;
  if (out3[0]) {
    size_t gran = 0;
    CUmemAllocationProp prop = vmm_prop(ord);
    DPB_CU(a.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    if (mc && ndev > 1) {
      CUmulticastObjectProp mp{};
      mp.numDevices = (unsigned)ndev;
      mp.size = gran;
      mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      size_t mg = 0;
      if (a.MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran) gran = mg;
    }
    out3[2] = (long long)gran;
  }
  return 0;
}

inline int vmm_map(CUmemGenericAllocationHandle h, size_t bytes, size_t align, int ordinal, void** ptr) {
  const VmmApi& a = vmm_api();
  CUdeviceptr va = 0;
  DPB_CU(a.MemAddressReserve(&va, bytes, align, 0, 0));
  CUresult r = a.MemMap(va, bytes, 0, h, 0);
  if (r != CUDA_SUCCESS) {
    a.MemAddressFree(va, bytes);
    return set_error(-50, "cuMemMap failed: CUresult %d", (int)r);
  }
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = ordinal;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = a.MemSetAccess(va, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) {
    a.MemUnmap(va, bytes);
    a.MemAddressFree(va, bytes);
    return set_error(-50, "cuMemSetAccess failed: CUresult %d", (int)r);
  }
  *ptr = reinterpret_cast<void*>(va);
  return 0;
}

// Physical memory on the current device (bytes must be a multiple of the granularity), mapped, zeroed, exported.
inline int vmm_alloc(size_t bytes, size_t gran, void** ptr, unsigned long long* handle, int* fd) {
  CUdevice dev;
  int ord = 0;
  if (int rc = vmm_current_device(&dev, &ord)) return rc;
  const VmmApi& a = vmm_api();
  CUmemAllocationProp prop = vmm_prop(ord);
  CUmemGenericAllocationHandle h = 0;
  DPB_CU(a.MemCreate(&h, bytes, &prop, 0));
  if (int rc = vmm_map(h, bytes, gran, ord, ptr)) {
    a.MemRelease(h);
    return rc;
  }
  DPB_CUDA(cudaMemset(*ptr, 0, bytes));
  DPB_CUDA(cudaDeviceSynchronize());
  int f = -1;
  DPB_CU(a.MemExportToShareableHandle(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle = (unsigned long long)h;
  *fd = f;
  return 0;
}

// A peer's allocation (or a multicast object), received as a file descriptor, into a handle; `map` also maps it.
inline int vmm_import(int fd, size_t bytes, size_t gran, int map, void** ptr, unsigned long long* handle) {
  CUdevice dev;
  int ord = 0;
  if (int rc = vmm_current_device(&dev, &ord)) return rc;
  const VmmApi& a = vmm_api();
  CUmemGenericAllocationHandle h = 0;
  DPB_CU(a.MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  if (map) {
    if (int rc = vmm_map(h, bytes, gran, ord, ptr)) {
      a.MemRelease(h);
      return rc;
    }
  }
  *handle = (unsigned long long)h;
  return 0;
}

inline int vmm_unmap(void* ptr, size_t bytes, unsigned long long handle) {
  const VmmApi& a = vmm_api();
  if (!a.ok) return 0;
  if (ptr != nullptr) {
    a.MemUnmap((CUdeviceptr)(uintptr_t)ptr, bytes);
    a.MemAddressFree((CUdeviceptr)(uintptr_t)ptr, bytes);
  }
  if (handle != 0) a.MemRelease((CUmemGenericAllocationHandle)handle);
  return 0;
}

inline int mc_create(int ndev, size_t bytes, unsigned long long* handle, int* fd) {
  const VmmApi& a = vmm_api();
  if (!a.ok) return set_error(-51, "virtual memory management entry points unavailable");
  CUmulticastObjectProp mp{};
  mp.numDevices = (unsigned)ndev;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h = 0;
  DPB_CU(a.MulticastCreate(&h, &mp));
  int f = -1;
  DPB_CU(a.MemExportToShareableHandle(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle = (unsigned long long)h;
  *fd = f;
  return 0;
}

inline int mc_add_device(unsigned long long mc) {
  CUdevice dev;
  if (int rc = vmm_current_device(&dev, nullptr)) return rc;
  DPB_CU(vmm_api().MulticastAddDevice((CUmemGenericAllocationHandle)mc, dev));
  return 0;
}

// Every device has been added (the caller synchronised the ranks): bind this rank's memory and map the multicast view.
inline int mc_bind_map(unsigned long long mc, unsigned long long mem, size_t bytes, size_t gran, void** mc_ptr) {
  CUdevice dev;
  int ord = 0;
  if (int rc = vmm_current_device(&dev, &ord)) return rc;
  const VmmApi& a = vmm_api();
  DPB_CU(a.MulticastBindMem((CUmemGenericAllocationHandle)mc, 0, (CUmemGenericAllocationHandle)mem, 0, bytes, 0));
  return vmm_map((CUmemGenericAllocationHandle)mc, bytes, gran, ord, mc_ptr);
}

inline int mc_unbind(unsigned long long mc, size_t bytes) {
  CUdevice dev;
  if (int rc = vmm_current_device(&dev, nullptr)) return rc;
  vmm_api().MulticastUnbind((CUmemGenericAllocationHandle)mc, dev, 0, bytes);
  return 0;
}

}  // namespace dpb
