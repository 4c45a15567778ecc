// Symmetric-memory allocator with NVLink-SHARP multicast binding, on the CUDA virtual-memory-management API
// (SURVEY work-list X-a).  Replaces, for the peer-memory communicator, what the reference gets from TensorFlow's gRPC
// runtime: a way for every worker to address every other worker's variables and gradients (main.py:80-93).
//
//   dtm_vmm_create      physical allocation on one GPU (cuMemCreate, shareable as a POSIX file descriptor), mapped into a
//                       reserved virtual range with read/write access for that GPU, zero-filled
//   dtm_vmm_export_fd   file descriptor of an allocation or of a multicast object, to be passed to the peer PROCESSES
//                       (SCM_RIGHTS over a unix socket - parallel/peer_comm.py); a single-process job needs no export
//   dtm_vmm_import_fd   maps a peer's allocation into this process (peer loads / stores over NVLink)
//   dtm_vmm_grant       gives one more GPU of THIS process access to a mapped range (single-process tower mode)
//   dtm_mc_*            multicast object: created once for N devices, every device is added, every device binds its
//                       allocation, and the object is mapped like memory - the resulting address reaches all N copies:
//                       multimem.ld_reduce sums them inside the NVSwitch, multimem.st writes all of them
//
// The driver entry points are resolved at run time (cudaGetDriverEntryPoint): the library keeps no link-time dependency on
// libcuda.so and still loads on a machine without a driver.
#include <mutex>
#include <stdio.h>
#include <string.h>
#include <unistd.h>
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"

namespace dtm {

struct VmmApi {
  CUresult (*memCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*memRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*memGetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*memAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*memAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*memMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*memUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*memSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*memExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*memImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
  CUresult (*mcCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*mcAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*mcBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
  CUresult (*mcGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
  CUresult (*deviceGet)(CUdevice*, int) = nullptr;
  CUresult (*deviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  CUresult (*getErrorString)(CUresult, const char**) = nullptr;
  bool ok = false, mc_ok = false;
};

static const VmmApi& vmm_api() {
  static VmmApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    auto get = [](const char* name, void** fn) {
      cudaDriverEntryPointQueryResult q;
      void* p = nullptr;
      if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
      *fn = p;
      return p != nullptr;
    };
    bool ok = true;
    ok &= get("cuMemCreate", reinterpret_cast<void**>(&api.memCreate));
    ok &= get("cuMemRelease", reinterpret_cast<void**>(&api.memRelease));
    ok &= get("cuMemGetAllocationGranularity", reinterpret_cast<void**>(&api.memGetGranularity));
    ok &= get("cuMemAddressReserve", reinterpret_cast<void**>(&api.memAddressReserve));
    ok &= get("cuMemAddressFree", reinterpret_cast<void**>(&api.memAddressFree));
    ok &= get("cuMemMap", reinterpret_cast<void**>(&api.memMap));
    ok &= get("cuMemUnmap", reinterpret_cast<void**>(&api.memUnmap));
    ok &= get("cuMemSetAccess", reinterpret_cast<void**>(&api.memSetAccess));
    ok &= get("cuMemExportToShareableHandle", reinterpret_cast<void**>(&api.memExport));
    ok &= get("cuMemImportFromShareableHandle", reinterpret_cast<void**>(&api.memImport));
    ok &= get("cuDeviceGet", reinterpret_cast<void**>(&api.deviceGet));
    ok &= get("cuDeviceGetAttribute", reinterpret_cast<void**>(&api.deviceGetAttribute));
    ok &= get("cuGetErrorString", reinterpret_cast<void**>(&api.getErrorString));
    api.ok = ok;
    bool mc = ok;
    mc &= get("cuMulticastCreate", reinterpret_cast<void**>(&api.mcCreate));
    mc &= get("cuMulticastAddDevice", reinterpret_cast<void**>(&api.mcAddDevice));
    mc &= get("cuMulticastBindMem", reinterpret_cast<void**>(&api.mcBindMem));
    mc &= get("cuMulticastGetGranularity", reinterpret_cast<void**>(&api.mcGetGranularity));
    api.mc_ok = mc;
  });
  return api;
}

static int fail_cu(CUresult r, const char* what) {
  const char* msg = nullptr;
  const VmmApi& a = vmm_api();
  if (a.getErrorString) a.getErrorString(r, &msg);
  char buf[256];
  snprintf(buf, sizeof(buf), "%s failed: CUresult %d (%s)", what, static_cast<int>(r), msg ? msg : "?");
  return dtm_fail(buf);
}
#define DTM_CU_OK(expr)                                   \
  do {                                                    \
    CUresult _r = (expr);                                 \
    if (_r != CUDA_SUCCESS) return fail_cu(_r, #expr);    \
  } while (0)

static CUmemAllocationProp alloc_prop(int device) {
  CUmemAllocationProp p;
  memset(&p, 0, sizeof(p));
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = device;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

static int map_rw(CUmemGenericAllocationHandle h, size_t bytes, int device, void** va_out) {
  const VmmApi& a = vmm_api();
  CUdeviceptr va = 0;
  DTM_CU_OK(a.memAddressReserve(&va, bytes, 0, 0, 0));
  CUresult r = a.memMap(va, bytes, 0, h, 0);
  if (r != CUDA_SUCCESS) { a.memAddressFree(va, bytes); return fail_cu(r, "cuMemMap"); }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = a.memSetAccess(va, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) { a.memUnmap(va, bytes); a.memAddressFree(va, bytes); return fail_cu(r, "cuMemSetAccess"); }
  *va_out = reinterpret_cast<void*>(va);
  return 0;
}

}  // namespace dtm

extern "C" int dtm_vmm_supported(int device, int* vmm, int* multicast, size_t* granularity) {
  using namespace dtm;
  const VmmApi& a = vmm_api();
  if (vmm) *vmm = 0;
  if (multicast) *multicast = 0;
  if (granularity) *granularity = 0;
  if (!a.ok) return 0;
  DTM_CUDA_OK(cudaFree(nullptr));     // make sure the primary context exists before driver calls
  CUdevice dev;
  DTM_CU_OK(a.deviceGet(&dev, device));
  int v = 0, fd = 0, mc = 0;
  a.deviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev);
  a.deviceGetAttribute(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
  if (a.mc_ok) a.deviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  if (vmm) *vmm = (v && fd) ? 1 : 0;
  if (multicast) *multicast = (v && fd && mc) ? 1 : 0;
  if (granularity && v) {
    CUmemAllocationProp p = alloc_prop(device);
    size_t g = 0;
    DTM_CU_OK(a.memGetGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    if (a.mc_ok && mc) {
      CUmulticastObjectProp mp;
      memset(&mp, 0, sizeof(mp));
      mp.numDevices = 2;
      mp.size = g;
      mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      size_t mg = 0;
      if (a.mcGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > g) g = mg;
    }
    *granularity = g;
  }
  return 0;
}

extern "C" int dtm_vmm_create(size_t bytes, int device, unsigned long long* handle, void** va) {
  using namespace dtm;
  const VmmApi& a = vmm_api();
  if (!a.ok) return dtm_fail("vmm: driver entry points unavailable");
  if (!handle || !va || bytes == 0) return dtm_fail("vmm_create: bad arguments");
  CUmemAllocationProp p = alloc_prop(device);
  CUmemGenericAllocationHandle h;
  DTM_CU_OK(a.memCreate(&h, bytes, &p, 0));
  if (int rc = map_rw(h, bytes, device, va)) { a.memRelease(h); return rc; }
  int cur = 0;
  DTM_CUDA_OK(cudaGetDevice(&cur));
  DTM_CUDA_OK(cudaSetDevice(device));
  DTM_CUDA_OK(cudaMemset(*va, 0, bytes));
  DTM_CUDA_OK(cudaDeviceSynchronize());
  DTM_CUDA_OK(cudaSetDevice(cur));
  *handle = static_cast<unsigned long long>(h);
  return 0;
}

extern "C" int dtm_vmm_export_fd(unsigned long long handle, int* fd) {
  using namespace dtm;
  const VmmApi& a = vmm_api();
  if (!a.ok || !fd) return dtm_fail("vmm_export_fd: unavailable or null argument");
  int out = -1;
  DTM_CU_OK(a.memExport(&out, static_cast<CUmemGenericAllocationHandle>(handle), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd = out;
  return 0;
}

extern "C" int dtm_vmm_import_fd(int fd, unsigned long long* handle) {
  using namespace dtm;
  const VmmApi& a = vmm_api();
  if (!a.ok || !handle) return dtm_fail("vmm_import_fd: unavailable or null argument");
  CUmemGenericAllocationHandle h;
  DTM_CU_OK(a.memImport(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *handle = static_cast<unsigned long long>(h);
  return 0;
}

extern "C" int dtm_vmm_map(unsigned long long handle, size_t bytes, int device, void** va) {
  using namespace dtm;
  if (!vmm_api().ok || !va) return dtm_fail("vmm_map: unavailable or null argument");
  return map_rw(static_cast<CUmemGenericAllocationHandle>(handle), bytes, device, va);
}

extern "C" int dtm_vmm_grant(void* va, size_t bytes, int device) {
  using namespace dtm;
  const VmmApi& a = vmm_api();
  if (!a.ok) return dtm_fail("vmm_grant: unavailable");
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DTM_CU_OK(a.memSetAccess(reinterpret_cast<CUdeviceptr>(va), bytes, &acc, 1));
  return 0;
}

extern "C" int dtm_vmm_unmap(void* va, size_t bytes) {
  using namespace dtm;
  const VmmApi& a = vmm_api();
  if (!a.ok) return dtm_fail("vmm_unmap: unavailable");
  DTM_CU_OK(a.memUnmap(reinterpret_cast<CUdeviceptr>(va), bytes));
  DTM_CU_OK(a.memAddressFree(reinterpret_cast<CUdeviceptr>(va), bytes));
  return 0;
}

extern "C" int dtm_vmm_release(unsigned long long handle) {
  using namespace dtm;
  const VmmApi& a = vmm_api();
  if (!a.ok) return dtm_fail("vmm_release: unavailable");
  DTM_CU_OK(a.memRelease(static_cast<CUmemGenericAllocationHandle>(handle)));
  return 0;
}

extern "C" int dtm_mc_create(int num_devices, size_t bytes, unsigned long long* mc_handle) {
  using namespace dtm;
  const VmmApi& a = vmm_api();
  if (!a.mc_ok || !mc_handle) return dtm_fail("mc_create: multicast entry points unavailable");
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = static_cast<unsigned int>(num_devices);
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  DTM_CU_OK(a.mcCreate(&h, &mp));
  *mc_handle = static_cast<unsigned long long>(h);
  return 0;
}

extern "C" int dtm_mc_add_device(unsigned long long mc_handle, int device) {
  using namespace dtm;
  const VmmApi& a = vmm_api();
  if (!a.mc_ok) return dtm_fail("mc_add_device: unavailable");
  CUdevice dev;
  DTM_CU_OK(a.deviceGet(&dev, device));
  DTM_CU_OK(a.mcAddDevice(static_cast<CUmemGenericAllocationHandle>(mc_handle), dev));
  return 0;
}

// Only after EVERY device of the object has been added (the caller puts a barrier in between).
extern "C" int dtm_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, size_t bytes) {
  using namespace dtm;
  const VmmApi& a = vmm_api();
  if (!a.mc_ok) return dtm_fail("mc_bind: unavailable");
  DTM_CU_OK(a.mcBindMem(static_cast<CUmemGenericAllocationHandle>(mc_handle), 0, static_cast<CUmemGenericAllocationHandle>(mem_handle), 0,
                        bytes, 0));
  return 0;
}
