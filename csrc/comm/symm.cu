// Symmetric-heap segments: device memory that every rank of the node maps.
//
// Two back ends:
//   * VMM (preferred): the segment is a cuMemCreate allocation exported as a POSIX file descriptor; the fd travels
//     to the peers over a Unix socket (SCM_RIGHTS, parallel/symm.py) and each peer imports and maps it.  The same
//     physical pages of all ranks are also bound to ONE multicast object (cuMulticastCreate / AddDevice / BindMem),
//     whose mapping is the NVLS address the multimem.* kernels in collectives.cu use: a load from it is reduced
//     inside the NVSwitch, a store to it lands in every rank's copy.
//   * CUDA IPC (fallback when the driver reports no VMM / fd-handle support): cudaMalloc + cudaIpcGetMemHandle,
//     the 64-byte handle travels over the bootstrap process group.  No multicast mapping on this path.
//
// Stands in for what ProcessGroupNCCL sets up behind init_process_group in the
// reference (/root/reference/src/distributed_trainer.py:62).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstring>
#include <mutex>

#include "../../include/dtb200.h"
#include "../common/host.h"

// ---- CUDA IPC back end ----------------------------------------------------------------------------------------------
extern "C" int dtb_symm_alloc(size_t bytes, void** ptr_out, void* handle_out) {
  DTB_REQUIRE(bytes > 0 && ptr_out && handle_out, "dtb_symm_alloc: bad arguments");
  void* p = nullptr;
  DTB_CUDA_OK(cudaMalloc(&p, bytes));
  DTB_CUDA_OK(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    dtb::set_error("cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    return 1;
  }
  static_assert(sizeof(h) == 64, "IPC handle is 64 bytes");
  memcpy(handle_out, &h, sizeof(h));
  *ptr_out = p;
  DTB_CUDA_OK(cudaDeviceSynchronize());
  return 0;
}

extern "C" int dtb_symm_open(const void* handle, void** ptr_out) {
  DTB_REQUIRE(handle && ptr_out, "dtb_symm_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  DTB_CUDA_OK(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

extern "C" int dtb_symm_close(void* ptr) {
  if (ptr) DTB_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

extern "C" int dtb_symm_free(void* ptr) {
  if (ptr) DTB_CUDA_OK(cudaFree(ptr));
  return 0;
}

// ---- VMM + multicast back end ---------------------------------------------------------------------------------------
// The driver API is reached through cudaGetDriverEntryPoint so the library keeps loading on a box without libcuda.
namespace {

struct DriverApi {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                               unsigned long long);
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*GetErrorString)(CUresult, const char**);
  bool ok = false;
};

template <typename F>
bool fetch(const char* name, F* out) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p)
    return false;
  *out = reinterpret_cast<F>(p);
  return true;
}

const DriverApi& api() {
  static DriverApi d;
  static std::once_flag once;
  std::call_once(once, [] {
    bool ok = true;
    ok &= fetch("cuMemCreate", &d.MemCreate);
    ok &= fetch("cuMemRelease", &d.MemRelease);
    ok &= fetch("cuMemExportToShareableHandle", &d.MemExport);
    ok &= fetch("cuMemImportFromShareableHandle", &d.MemImport);
    ok &= fetch("cuMemAddressReserve", &d.MemAddressReserve);
    ok &= fetch("cuMemAddressFree", &d.MemAddressFree);
    ok &= fetch("cuMemMap", &d.MemMap);
    ok &= fetch("cuMemUnmap", &d.MemUnmap);
    ok &= fetch("cuMemSetAccess", &d.MemSetAccess);
    ok &= fetch("cuMemGetAllocationGranularity", &d.MemGetAllocationGranularity);
    ok &= fetch("cuMulticastCreate", &d.MulticastCreate);
    ok &= fetch("cuMulticastAddDevice", &d.MulticastAddDevice);
    ok &= fetch("cuMulticastBindMem", &d.MulticastBindMem);
    ok &= fetch("cuMulticastUnbind", &d.MulticastUnbind);
    ok &= fetch("cuMulticastGetGranularity", &d.MulticastGetGranularity);
    ok &= fetch("cuDeviceGetAttribute", &d.DeviceGetAttribute);
    ok &= fetch("cuDeviceGet", &d.DeviceGet);
    ok &= fetch("cuGetErrorString", &d.GetErrorString);
    d.ok = ok;
  });
  return d;
}

const char* cu_text(CUresult r) {
  const char* s = nullptr;
  if (api().GetErrorString && api().GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
  return "unknown driver error";
}

int current_cu_device(CUdevice* dev, int* ordinal) {
  int ord = 0;
  DTB_CUDA_OK(cudaGetDevice(&ord));
  DTB_CUDA_OK(cudaFree(nullptr));     // make sure the primary context exists
  CUresult r = api().DeviceGet(dev, ord);
  if (r != CUDA_SUCCESS) {
    dtb::set_error("cuDeviceGet(%d) failed: %s", ord, cu_text(r));
    return 1;
  }
  if (ordinal) *ordinal = ord;
  return 0;
}

CUmemAllocationProp local_prop(int ordinal) {
  CUmemAllocationProp prop = {};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = ordinal;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

}  // namespace

#define DTB_CU_OK(expr)                                                                     \
  do {                                                                                      \
    CUresult _r = (expr);                                                                   \
    if (_r != CUDA_SUCCESS) {                                                               \
      dtb::set_error("%s failed: %s (%s:%d)", #expr, cu_text(_r), __FILE__, __LINE__);      \
      return 1;                                                                             \
    }                                                                                       \
  } while (0)

extern "C" int dtb_symm_vmm_probe(int* vmm_ok, int* multicast_ok) {
  DTB_REQUIRE(vmm_ok && multicast_ok, "dtb_symm_vmm_probe: bad arguments");
  *vmm_ok = *multicast_ok = 0;
  if (!api().ok) return 0;
  CUdevice dev;
  int ord = 0;
  if (current_cu_device(&dev, &ord)) return 1;
  int v = 0, fd = 0, mc = 0;
  api().DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev);
  api().DeviceGetAttribute(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
  api().DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  *vmm_ok = (v && fd) ? 1 : 0;
  *multicast_ok = (v && fd && mc) ? 1 : 0;
  return 0;
}

extern "C" int dtb_symm_vmm_granularity(int world, int multicast, size_t* gran_out) {
  DTB_REQUIRE(gran_out && world >= 1, "dtb_symm_vmm_granularity: bad arguments");
  DTB_REQUIRE(api().ok, "dtb_symm_vmm_granularity: driver VMM entry points unavailable");
  CUdevice dev;
  int ord = 0;
  if (current_cu_device(&dev, &ord)) return 1;
  CUmemAllocationProp prop = local_prop(ord);
  size_t g = 0;
  DTB_CU_OK(api().MemGetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (multicast) {
    CUmulticastObjectProp mp = {};
    mp.numDevices = static_cast<unsigned>(world);
    mp.size = g;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    DTB_CU_OK(api().MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    if (mg > g) g = mg;
  }
  *gran_out = g;
  return 0;
}

extern "C" int dtb_symm_vmm_create(size_t bytes, unsigned long long* handle_out, int* fd_out) {
  DTB_REQUIRE(bytes > 0 && handle_out && fd_out, "dtb_symm_vmm_create: bad arguments");
  DTB_REQUIRE(api().ok, "dtb_symm_vmm_create: driver VMM entry points unavailable");
  CUdevice dev;
  int ord = 0;
  if (current_cu_device(&dev, &ord)) return 1;
  CUmemAllocationProp prop = local_prop(ord);
  CUmemGenericAllocationHandle h = 0;
  DTB_CU_OK(api().MemCreate(&h, bytes, &prop, 0));
  int fd = -1;
  CUresult r = api().MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    api().MemRelease(h);
    dtb::set_error("cuMemExportToShareableHandle failed: %s", cu_text(r));
    return 1;
  }
  *handle_out = h;
  *fd_out = fd;
  return 0;
}

extern "C" int dtb_symm_vmm_import(int fd, unsigned long long* handle_out) {
  DTB_REQUIRE(fd >= 0 && handle_out, "dtb_symm_vmm_import: bad arguments");
  DTB_REQUIRE(api().ok, "dtb_symm_vmm_import: driver VMM entry points unavailable");
  CUmemGenericAllocationHandle h = 0;
  DTB_CU_OK(api().MemImport(&h, reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                            CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *handle_out = h;
  return 0;
}

extern "C" int dtb_symm_vmm_map(unsigned long long handle, size_t bytes, size_t alignment, void** ptr_out) {
  DTB_REQUIRE(handle && bytes > 0 && ptr_out, "dtb_symm_vmm_map: bad arguments");
  DTB_REQUIRE(api().ok, "dtb_symm_vmm_map: driver VMM entry points unavailable");
  CUdevice dev;
  int ord = 0;
  if (current_cu_device(&dev, &ord)) return 1;
  CUdeviceptr va = 0;
  DTB_CU_OK(api().MemAddressReserve(&va, bytes, alignment, 0, 0));
  CUresult r = api().MemMap(va, bytes, 0, handle, 0);
  if (r != CUDA_SUCCESS) {
    api().MemAddressFree(va, bytes);
    dtb::set_error("cuMemMap failed: %s", cu_text(r));
    return 1;
  }
  CUmemAccessDesc acc = {};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = ord;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = api().MemSetAccess(va, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) {
    api().MemUnmap(va, bytes);
    api().MemAddressFree(va, bytes);
    dtb::set_error("cuMemSetAccess failed: %s (is peer access between the GPUs available?)", cu_text(r));
    return 1;
  }
  *ptr_out = reinterpret_cast<void*>(va);
  return 0;
}

extern "C" int dtb_symm_vmm_unmap(void* ptr, size_t bytes) {
  if (!ptr) return 0;
  DTB_REQUIRE(api().ok, "dtb_symm_vmm_unmap: driver VMM entry points unavailable");
  DTB_CU_OK(api().MemUnmap(reinterpret_cast<CUdeviceptr>(ptr), bytes));
  DTB_CU_OK(api().MemAddressFree(reinterpret_cast<CUdeviceptr>(ptr), bytes));
  return 0;
}

extern "C" int dtb_symm_vmm_release(unsigned long long handle) {
  if (!handle) return 0;
  DTB_REQUIRE(api().ok, "dtb_symm_vmm_release: driver VMM entry points unavailable");
  DTB_CU_OK(api().MemRelease(handle));
  return 0;
}

extern "C" int dtb_symm_mc_create(size_t bytes, int world, unsigned long long* handle_out, int* fd_out) {
  DTB_REQUIRE(bytes > 0 && world >= 2 && handle_out && fd_out, "dtb_symm_mc_create: bad arguments");
  DTB_REQUIRE(api().ok, "dtb_symm_mc_create: driver VMM entry points unavailable");
  CUmulticastObjectProp mp = {};
  mp.numDevices = static_cast<unsigned>(world);
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h = 0;
  DTB_CU_OK(api().MulticastCreate(&h, &mp));
  int fd = -1;
  CUresult r = api().MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    api().MemRelease(h);
    dtb::set_error("cuMemExportToShareableHandle(multicast) failed: %s", cu_text(r));
    return 1;
  }
  *handle_out = h;
  *fd_out = fd;
  return 0;
}

extern "C" int dtb_symm_mc_add_device(unsigned long long mc_handle) {
  DTB_REQUIRE(mc_handle, "dtb_symm_mc_add_device: bad arguments");
  DTB_REQUIRE(api().ok, "dtb_symm_mc_add_device: driver VMM entry points unavailable");
  CUdevice dev;
  if (current_cu_device(&dev, nullptr)) return 1;
  DTB_CU_OK(api().MulticastAddDevice(mc_handle, dev));
  return 0;
}

extern "C" int dtb_symm_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, size_t bytes) {
  DTB_REQUIRE(mc_handle && mem_handle && bytes > 0, "dtb_symm_mc_bind: bad arguments");
  DTB_REQUIRE(api().ok, "dtb_symm_mc_bind: driver VMM entry points unavailable");
  DTB_CU_OK(api().MulticastBindMem(mc_handle, 0, mem_handle, 0, bytes, 0));
  return 0;
}

extern "C" int dtb_symm_mc_unbind(unsigned long long mc_handle, size_t bytes) {
  if (!mc_handle) return 0;
  DTB_REQUIRE(api().ok, "dtb_symm_mc_unbind: driver VMM entry points unavailable");
  CUdevice dev;
  if (current_cu_device(&dev, nullptr)) return 1;
  DTB_CU_OK(api().MulticastUnbind(mc_handle, dev, 0, bytes));
  return 0;
}

// Plain cudaMemcpyAsync between any two mapped addresses (local or peer): the copy-engine rate bench/allreduce_sweep.py
// records as the NVLink denominator.
extern "C" int dtb_debug_memcpy_async(void* dst, const void* src, size_t bytes, void* stream) {
  DTB_CUDA_OK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, static_cast<cudaStream_t>(stream)));
  return 0;
}
