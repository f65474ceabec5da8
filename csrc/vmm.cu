// H2: symmetric-memory allocator on the CUDA virtual-memory-management driver API.
//
// One physical allocation per rank (cuMemCreate, exportable as a POSIX file descriptor), mapped into the address
// space of every rank of the node (cuMemImportFromShareableHandle + cuMemMap + cuMemSetAccess), and — where the
// fabric offers it — bound into one NVSwitch multicast object (cuMulticastCreate / AddDevice / BindMem) whose own
// mapping is the address `multimem.*` instructions take. File descriptors travel between the processes over unix
// sockets (parallel/symm.py); this file only talks to the driver.
//
// The driver entry points are looked up through cudaGetDriverEntryPoint, so the library still does not link libcuda.
// Kernels read peers' mappings with plain ld/st and the multicast mapping with multimem.ld_reduce / multimem.st;
// no tensor map and no bulk copy is ever built over either ("Peer-memory rule").
// The reference has no communication layer.
#include <cstdint>
#include <cstring>
#include <mutex>

#include "common.cuh"
#include "../include/b200zero.h"

namespace bz {

struct VmmApi {
  CUresult (*memCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*memRelease)(CUmemGenericAllocationHandle);
  CUresult (*memAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*memAddressFree)(CUdeviceptr, size_t);
  CUresult (*memMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*memUnmap)(CUdeviceptr, size_t);
  CUresult (*memSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*memExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*memImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*memGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*mcCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*mcAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*mcBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                        unsigned long long);
  CUresult (*mcUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*mcGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*deviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  bool ok = false;
  bool mc_ok = false;
};

template <class F>
static bool entry(const char* name, F* out) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
      p == nullptr) {
    cudaGetLastError();
    return false;
  }
  *out = reinterpret_cast<F>(p);
  return true;
}

static VmmApi& api() {
  static VmmApi a;
  static std::once_flag once;
  std::call_once(once, [] {
    a.ok = entry("cuMemCreate", &a.memCreate) && entry("cuMemRelease", &a.memRelease) &&
           entry("cuMemAddressReserve", &a.memAddressReserve) && entry("cuMemAddressFree", &a.memAddressFree) &&
           entry("cuMemMap", &a.memMap) && entry("cuMemUnmap", &a.memUnmap) && entry("cuMemSetAccess", &a.memSetAccess) &&
           entry("cuMemExportToShareableHandle", &a.memExport) &&
           entry("cuMemImportFromShareableHandle", &a.memImport) &&
           entry("cuMemGetAllocationGranularity", &a.memGranularity) &&
           entry("cuDeviceGetAttribute", &a.deviceGetAttribute);
    a.mc_ok = a.ok && entry("cuMulticastCreate", &a.mcCreate) && entry("cuMulticastAddDevice", &a.mcAddDevice) &&
              entry("cuMulticastBindMem", &a.mcBindMem) && entry("cuMulticastUnbind", &a.mcUnbind) &&
              entry("cuMulticastGetGranularity", &a.mcGranularity);
  });
  return a;
}

#define BZ_CU_CHECK(expr)                                                                          \
  do {                                                                                             \
    CUresult _r = (expr);                                                                          \
    if (_r != CUDA_SUCCESS) {                                                                      \
      bz::set_last_error("%s failed with CUresult %d (%s:%d)", #expr, (int)_r, __FILE__, __LINE__); \
      return 1;                                                                                    \
    }                                                                                              \
  } while (0)

static int current_device(int* dev) {
  BZ_CUDA_CHECK(cudaFree(0));          // make sure the primary context exists and is current
  BZ_CUDA_CHECK(cudaGetDevice(dev));
  return 0;
}

static CUmemAllocationProp device_prop(int dev) {
  CUmemAllocationProp p;
  memset(&p, 0, sizeof(p));
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = dev;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

static int map_rw(VmmApi& a, CUmemGenericAllocationHandle h, size_t size, int dev, void** out) {
  CUdeviceptr va = 0;
  BZ_CU_CHECK(a.memAddressReserve(&va, size, 0, 0, 0));
  if (a.memMap(va, size, 0, h, 0) != CUDA_SUCCESS) {
    a.memAddressFree(va, size);
    set_last_error("cuMemMap failed for %zu bytes", size);
    return 1;
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  if (a.memSetAccess(va, size, &acc, 1) != CUDA_SUCCESS) {
    a.memUnmap(va, size);
    a.memAddressFree(va, size);
    set_last_error("cuMemSetAccess failed (device %d cannot reach this allocation)", dev);
    return 1;
  }
  *out = reinterpret_cast<void*>(va);
  return 0;
}

}  // namespace bz

using namespace bz;

extern "C" {

// 1 when this device and driver can allocate exportable VMM memory; *multicast = 1 when NVSwitch multicast objects
// are available too.
int bz_vmm_supported(int* vmm, int* multicast) {
  VmmApi& a = api();
  *vmm = 0;
  *multicast = 0;
  if (!a.ok) return 0;
  int dev = 0;
  if (current_device(&dev)) return 1;
  int v = 0;
  if (a.deviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev) == CUDA_SUCCESS) *vmm = v;
  v = 0;
  if (a.mc_ok && a.deviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) == CUDA_SUCCESS) *multicast = v;
  return 0;
}

// Size that an allocation of `bytes` is rounded up to: a multiple of the allocation granularity and, when
// `world` > 1 and multicast is available, of the multicast granularity for `world` devices as well.
int bz_vmm_round_size(unsigned long long bytes, int world, unsigned long long* out) {
  VmmApi& a = api();
  BZ_REQUIRE(a.ok, "vmm: driver entry points unavailable");
  int dev = 0;
  if (current_device(&dev)) return 1;
  CUmemAllocationProp p = device_prop(dev);
  size_t g = 0;
  BZ_CU_CHECK(a.memGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  int mc = 0;
  if (a.mc_ok && world > 1) a.deviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  if (mc) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)world;
    mp.size = bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (a.mcGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > g) g = mg;
  }
  if (g == 0) g = 2u << 20;
  *out = (bytes + g - 1) / g * g;
  return 0;
}

// Allocate `size` bytes (already rounded) on the current device, map them read/write, zero them, and export the
// allocation as a file descriptor for the other ranks.
int bz_vmm_alloc(unsigned long long size, void** ptr, unsigned long long* handle, int* fd) {
  VmmApi& a = api();
  BZ_REQUIRE(a.ok, "vmm: driver entry points unavailable");
  int dev = 0;
  if (current_device(&dev)) return 1;
  CUmemAllocationProp p = device_prop(dev);
  CUmemGenericAllocationHandle h;
  BZ_CU_CHECK(a.memCreate(&h, size, &p, 0));
  if (map_rw(a, h, size, dev, ptr)) {
    a.memRelease(h);
    return 1;
  }
  int f = -1;
  if (a.memExport(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) != CUDA_SUCCESS) {
    set_last_error("cuMemExportToShareableHandle failed");
    return 1;
  }
  BZ_CUDA_CHECK(cudaMemset(*ptr, 0, size));
  BZ_CUDA_CHECK(cudaDeviceSynchronize());
  *handle = (unsigned long long)h;
  *fd = f;
  return 0;
}

// Map a peer's allocation (received as a file descriptor) into this process, readable and writable from the
// current device.
int bz_vmm_import(int fd, unsigned long long size, void** ptr, unsigned long long* handle) {
  VmmApi& a = api();
  BZ_REQUIRE(a.ok, "vmm: driver entry points unavailable");
  int dev = 0;
  if (current_device(&dev)) return 1;
  CUmemGenericAllocationHandle h;
  BZ_CU_CHECK(a.memImport(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  if (map_rw(a, h, size, dev, ptr)) {
    a.memRelease(h);
    return 1;
  }
  *handle = (unsigned long long)h;
  return 0;
}

int bz_vmm_unmap(void* ptr, unsigned long long size, unsigned long long handle) {
  VmmApi& a = api();
  BZ_REQUIRE(a.ok, "vmm: driver entry points unavailable");
  if (ptr != nullptr) {
    BZ_CU_CHECK(a.memUnmap((CUdeviceptr)ptr, size));
    BZ_CU_CHECK(a.memAddressFree((CUdeviceptr)ptr, size));
  }
  if (handle != 0) BZ_CU_CHECK(a.memRelease((CUmemGenericAllocationHandle)handle));
  return 0;
}

// Rank 0: create the multicast object for `world` devices and export it.
int bz_mc_create(int world, unsigned long long size, unsigned long long* mc_handle, int* fd) {
  VmmApi& a = api();
  BZ_REQUIRE(a.mc_ok, "multicast: driver entry points unavailable");
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)world;
  mp.size = size;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  BZ_CU_CHECK(a.mcCreate(&h, &mp));
  int f = -1;
  BZ_CU_CHECK(a.memExport(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *mc_handle = (unsigned long long)h;
  *fd = f;
  return 0;
}

int bz_mc_import(int fd, unsigned long long* mc_handle) {
  VmmApi& a = api();
  BZ_REQUIRE(a.mc_ok, "multicast: driver entry points unavailable");
  CUmemGenericAllocationHandle h;
  BZ_CU_CHECK(a.memImport(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *mc_handle = (unsigned long long)h;
  return 0;
}

// Every rank, once: add the current device to the object. All ranks must have done this before anyone binds.
int bz_mc_add_device(unsigned long long mc_handle) {
  VmmApi& a = api();
  BZ_REQUIRE(a.mc_ok, "multicast: driver entry points unavailable");
  int dev = 0;
  if (current_device(&dev)) return 1;
  BZ_CU_CHECK(a.mcAddDevice((CUmemGenericAllocationHandle)mc_handle, dev));
  return 0;
}

// Every rank, after all devices were added: bind this rank's allocation at offset 0 of the object and map the
// object itself; *mc_ptr is the address multimem instructions use.
int bz_mc_bind_and_map(unsigned long long mc_handle, unsigned long long mem_handle, unsigned long long size,
                       void** mc_ptr) {
  VmmApi& a = api();
  BZ_REQUIRE(a.mc_ok, "multicast: driver entry points unavailable");
  int dev = 0;
  if (current_device(&dev)) return 1;
  BZ_CU_CHECK(a.mcBindMem((CUmemGenericAllocationHandle)mc_handle, 0, (CUmemGenericAllocationHandle)mem_handle, 0, size, 0));
  return map_rw(a, (CUmemGenericAllocationHandle)mc_handle, size, dev, mc_ptr);
}

int bz_mc_unmap(void* mc_ptr, unsigned long long size, unsigned long long mc_handle) {
  VmmApi& a = api();
  BZ_REQUIRE(a.mc_ok, "multicast: driver entry points unavailable");
  int dev = 0;
  if (current_device(&dev)) return 1;
  if (mc_ptr != nullptr) {
    BZ_CU_CHECK(a.memUnmap((CUdeviceptr)mc_ptr, size));
    BZ_CU_CHECK(a.memAddressFree((CUdeviceptr)mc_ptr, size));
  }
  if (mc_handle != 0) {
    a.mcUnbind((CUmemGenericAllocationHandle)mc_handle, dev, 0, size);
    BZ_CU_CHECK(a.memRelease((CUmemGenericAllocationHandle)mc_handle));
  }
  return 0;
}

}  // extern "C"
