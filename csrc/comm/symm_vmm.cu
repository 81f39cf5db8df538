// Symmetric heap on the CUDA virtual-memory-management API (component C1, large-page variant).
//
// Why not only cudaIpc*: a slab opened with cudaIpcOpenMemHandle is mapped into the peer with small
// pages.  Streaming access does not care, but the embedding gather reads random 512-byte rows of a
// multi-gigabyte shard, and then every row is a TLB miss on the reader: measured on 2xB200, random rows
// of a 6.4 GB peer shard ran at ~40 GB/s through an IPC mapping against ~735 GB/s for the same kernel
// on a <= 2 GB shard (benchmarks/ipc_probe.py).  Memory created with cuMemCreate is exported as a file
// descriptor and mapped by the peer with cuMemMap at the allocation's own (2 MiB) granularity, which
// keeps the peer's TLB reach the same as for local memory.
//   owner : zoo_vmm_alloc  -> device pointer + POSIX fd (sent to the peers over a unix socket, SCM_RIGHTS)
//   peer  : zoo_vmm_import -> device pointer of the same physical memory, read/write over NVLink
// Driver entry points are resolved at run time (no link-time dependency on libcuda).
#include "../common/zoo_common.cuh"
#include <unistd.h>

namespace {

template <typename F> F drv(const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
  return reinterpret_cast<F>(p);
}

typedef CUresult (*PFN_gran)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
typedef CUresult (*PFN_create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
typedef CUresult (*PFN_reserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
typedef CUresult (*PFN_map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
typedef CUresult (*PFN_access)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
typedef CUresult (*PFN_export)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
typedef CUresult (*PFN_import)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
typedef CUresult (*PFN_unmap)(CUdeviceptr, size_t);
typedef CUresult (*PFN_addrfree)(CUdeviceptr, size_t);
typedef CUresult (*PFN_release)(CUmemGenericAllocationHandle);

int fail(const char* what, CUresult r, int line) {
  char msg[160];
  snprintf(msg, sizeof(msg), "%s failed (CUresult %d)", what, (int)r);
  zoo_set_last_error(msg, __FILE__, line);
  return ZOO_ERR_CUDA;
}

int map_handle(CUmemGenericAllocationHandle h, size_t size, size_t gran, int dev, void** out) {
  auto reserve = drv<PFN_reserve>("cuMemAddressReserve");
  auto map = drv<PFN_map>("cuMemMap");
  auto access = drv<PFN_access>("cuMemSetAccess");
  if (!reserve || !map || !access) return fail("resolving cuMem* entry points", CUDA_ERROR_NOT_SUPPORTED, __LINE__);
  CUdeviceptr va = 0;
  CUresult r = reserve(&va, size, gran, 0, 0);
  if (r != CUDA_SUCCESS) return fail("cuMemAddressReserve", r, __LINE__);
  r = map(va, size, 0, h, 0);
  if (r != CUDA_SUCCESS) return fail("cuMemMap", r, __LINE__);
  CUmemAccessDesc d{};
  d.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  d.location.id = dev;
  d.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = access(va, size, &d, 1);
  if (r != CUDA_SUCCESS) return fail("cuMemSetAccess", r, __LINE__);
  *out = reinterpret_cast<void*>(va);
  return ZOO_OK;
}

CUmemAllocationProp prop_for(int dev) {
  CUmemAllocationProp p{};
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = dev;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

}  // namespace

// Allocate `bytes` (rounded up to the allocation granularity) on the current device, zero it, and export it.
//   ptr: device pointer; fd: shareable file descriptor (close it after the peers imported it);
//   mapped_bytes: size of the mapping (what zoo_vmm_import / zoo_vmm_free must be given).
ZOO_API int zoo_vmm_alloc(long long bytes, void** ptr, int* fd, long long* mapped_bytes) {
  if (!ptr || !fd || !mapped_bytes || bytes <= 0) return ZOO_ERR_ARG;
  int dev = 0;
  ZOO_CUDA_CHECK(cudaGetDevice(&dev));
  ZOO_CUDA_CHECK(cudaFree(0));
  auto gran_fn = drv<PFN_gran>("cuMemGetAllocationGranularity");
  auto create = drv<PFN_create>("cuMemCreate");
  auto exp = drv<PFN_export>("cuMemExportToShareableHandle");
  auto release = drv<PFN_release>("cuMemRelease");
  if (!gran_fn || !create || !exp || !release) return fail("resolving cuMem* entry points", CUDA_ERROR_NOT_SUPPORTED, __LINE__);
  CUmemAllocationProp prop = prop_for(dev);
  size_t gran = 0;
  CUresult r = gran_fn(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  if (r != CUDA_SUCCESS || gran == 0) return fail("cuMemGetAllocationGranularity", r, __LINE__);
  const size_t size = ((size_t)bytes + gran - 1) / gran * gran;
  CUmemGenericAllocationHandle h;
  r = create(&h, size, &prop, 0);
  if (r != CUDA_SUCCESS) return fail("cuMemCreate", r, __LINE__);
  int rc = map_handle(h, size, gran, dev, ptr);
  if (rc) { release(h); return rc; }
  int out_fd = -1;
  r = exp(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) return fail("cuMemExportToShareableHandle", r, __LINE__);
  release(h);                       // the mapping keeps the memory alive
  ZOO_CUDA_CHECK(cudaMemset(*ptr, 0, size));
  ZOO_CUDA_CHECK(cudaDeviceSynchronize());
  *fd = out_fd;
  *mapped_bytes = (long long)size;
  return ZOO_OK;
}

// Map a peer's allocation (received as a file descriptor) for read/write access from the current device.
ZOO_API int zoo_vmm_import(int fd, long long mapped_bytes, void** ptr) {
  if (!ptr || fd < 0 || mapped_bytes <= 0) return ZOO_ERR_ARG;
  int dev = 0;
  ZOO_CUDA_CHECK(cudaGetDevice(&dev));
  ZOO_CUDA_CHECK(cudaFree(0));
  auto imp = drv<PFN_import>("cuMemImportFromShareableHandle");
  auto gran_fn = drv<PFN_gran>("cuMemGetAllocationGranularity");
  auto release = drv<PFN_release>("cuMemRelease");
  if (!imp || !gran_fn || !release) return fail("resolving cuMem* entry points", CUDA_ERROR_NOT_SUPPORTED, __LINE__);
  CUmemGenericAllocationHandle h;
  CUresult r = imp(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) return fail("cuMemImportFromShareableHandle", r, __LINE__);
  CUmemAllocationProp prop = prop_for(dev);
  size_t gran = 0;
  r = gran_fn(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  if (r != CUDA_SUCCESS || gran == 0) return fail("cuMemGetAllocationGranularity", r, __LINE__);
  int rc = map_handle(h, (size_t)mapped_bytes, gran, dev, ptr);
  release(h);
  return rc;
}

ZOO_API int zoo_vmm_free(void* ptr, long long mapped_bytes) {
  if (!ptr || mapped_bytes <= 0) return ZOO_ERR_ARG;
  auto unmap = drv<PFN_unmap>("cuMemUnmap");
  auto afree = drv<PFN_addrfree>("cuMemAddressFree");
  if (!unmap || !afree) return fail("resolving cuMem* entry points", CUDA_ERROR_NOT_SUPPORTED, __LINE__);
  CUresult r = unmap((CUdeviceptr)(uintptr_t)ptr, (size_t)mapped_bytes);
  if (r != CUDA_SUCCESS) return fail("cuMemUnmap", r, __LINE__);
  r = afree((CUdeviceptr)(uintptr_t)ptr, (size_t)mapped_bytes);
  if (r != CUDA_SUCCESS) return fail("cuMemAddressFree", r, __LINE__);
  return ZOO_OK;
}

ZOO_API int zoo_close_fd(int fd) { return fd >= 0 ? ::close(fd) : 0; }

// ------------------------------------------------------------------ NVLS multicast objects (C1) --
// A multicast object gives every GPU of the team ONE extra virtual address range over the same offsets of everybody's
// buffer: a store to it is replicated by the NVSwitch into every GPU's copy (`multimem.st`), a `multimem.ld_reduce` returns the
// element-wise sum of all copies, reduced inside the switch.  The gradient exchange (csrc/comm/rs_opt_ag.cu) uses it so that the
// owner of a shard issues one load instead of `world` peer loads and one store instead of `world` peer stores.
//   rank 0     : zoo_mc_create  -> handle + POSIX fd (sent to the peers like the buffer fds)
//   other ranks: zoo_mc_import(fd) -> handle
//   every rank : zoo_mc_add_device(handle); barrier; zoo_mc_bind(handle, local buffer) -> multicast pointer
// The bound buffer must come from zoo_vmm_alloc_aligned with the multicast granularity (zoo_mc_granularity).
namespace {
typedef CUresult (*PFN_mcCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
typedef CUresult (*PFN_mcAddDevice)(CUmemGenericAllocationHandle, CUdevice);
typedef CUresult (*PFN_mcBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
typedef CUresult (*PFN_mcGran)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
typedef CUresult (*PFN_retain)(CUmemGenericAllocationHandle*, void*);
typedef CUresult (*PFN_devAttr)(int*, CUdevice_attribute, CUdevice);
typedef CUresult (*PFN_devGet)(CUdevice*, int);

CUmulticastObjectProp mc_prop(int world, size_t size) {
  CUmulticastObjectProp p{};
  p.numDevices = (unsigned)world;
  p.size = size;
  p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  p.flags = 0;
  return p;
}
}  // namespace

// 1 when the current device can take part in switch multicast / in-switch reductions.
ZOO_API int zoo_mc_supported(int* out) {
  if (!out) return ZOO_ERR_ARG;
  *out = 0;
  int dev = 0;
  ZOO_CUDA_CHECK(cudaGetDevice(&dev));
  ZOO_CUDA_CHECK(cudaFree(0));
  auto attr = drv<PFN_devAttr>("cuDeviceGetAttribute");
  auto dget = drv<PFN_devGet>("cuDeviceGet");
  if (!attr || !dget || !drv<PFN_mcCreate>("cuMulticastCreate")) return ZOO_OK;
  CUdevice d;
  if (dget(&d, dev) != CUDA_SUCCESS) return ZOO_OK;
  int v = 0;
  if (attr(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d) == CUDA_SUCCESS) *out = v;
  return ZOO_OK;
}

// Size granularity of multicast objects for a team of `world` GPUs (bind sizes and buffer sizes must be multiples of it).
ZOO_API int zoo_mc_granularity(int world, long long* gran) {
  if (!gran || world < 2) return ZOO_ERR_ARG;
  auto g = drv<PFN_mcGran>("cuMulticastGetGranularity");
  if (!g) return fail("resolving cuMulticastGetGranularity", CUDA_ERROR_NOT_SUPPORTED, __LINE__);
  CUmulticastObjectProp p = mc_prop(world, 0);
  size_t v = 0;
  CUresult r = g(&v, &p, CU_MULTICAST_GRANULARITY_RECOMMENDED);
  if (r != CUDA_SUCCESS || v == 0) return fail("cuMulticastGetGranularity", r, __LINE__);
  *gran = (long long)v;
  return ZOO_OK;
}

// zoo_vmm_alloc whose size is rounded up to a multiple of `align` bytes as well (multicast binding needs that).
ZOO_API int zoo_vmm_alloc_aligned(long long bytes, long long align, void** ptr, int* fd, long long* mapped_bytes) {
  if (align <= 0) return ZOO_ERR_ARG;
  const long long rounded = (bytes + align - 1) / align * align;
  return zoo_vmm_alloc(rounded, ptr, fd, mapped_bytes);
}

ZOO_API int zoo_mc_create(long long bytes, int world, unsigned long long* handle, int* fd) {
  if (!handle || !fd || bytes <= 0 || world < 2) return ZOO_ERR_ARG;
  ZOO_CUDA_CHECK(cudaFree(0));
  auto create = drv<PFN_mcCreate>("cuMulticastCreate");
  auto exp = drv<PFN_export>("cuMemExportToShareableHandle");
  if (!create || !exp) return fail("resolving cuMulticast* entry points", CUDA_ERROR_NOT_SUPPORTED, __LINE__);
  CUmulticastObjectProp p = mc_prop(world, (size_t)bytes);
  CUmemGenericAllocationHandle h;
  CUresult r = create(&h, &p);
  if (r != CUDA_SUCCESS) return fail("cuMulticastCreate", r, __LINE__);
  int out_fd = -1;
  r = exp(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) return fail("cuMemExportToShareableHandle(multicast)", r, __LINE__);
  *handle = (unsigned long long)h;
  *fd = out_fd;
  return ZOO_OK;
}

ZOO_API int zoo_mc_import(int fd, unsigned long long* handle) {
  if (!handle || fd < 0) return ZOO_ERR_ARG;
  ZOO_CUDA_CHECK(cudaFree(0));
  auto imp = drv<PFN_import>("cuMemImportFromShareableHandle");
  if (!imp) return fail("resolving cuMemImportFromShareableHandle", CUDA_ERROR_NOT_SUPPORTED, __LINE__);
  CUmemGenericAllocationHandle h;
  CUresult r = imp(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) return fail("cuMemImportFromShareableHandle(multicast)", r, __LINE__);
  *handle = (unsigned long long)h;
  return ZOO_OK;
}

ZOO_API int zoo_mc_add_device(unsigned long long handle) {
  int dev = 0;
  ZOO_CUDA_CHECK(cudaGetDevice(&dev));
  auto add = drv<PFN_mcAddDevice>("cuMulticastAddDevice");
  auto dget = drv<PFN_devGet>("cuDeviceGet");
  if (!add || !dget) return fail("resolving cuMulticastAddDevice", CUDA_ERROR_NOT_SUPPORTED, __LINE__);
  CUdevice d;
  CUresult r = dget(&d, dev);
  if (r != CUDA_SUCCESS) return fail("cuDeviceGet", r, __LINE__);
  r = add((CUmemGenericAllocationHandle)handle, d);
  if (r != CUDA_SUCCESS) return fail("cuMulticastAddDevice", r, __LINE__);
  return ZOO_OK;
}

// Bind the local buffer (from zoo_vmm_alloc_aligned; `bytes` = its mapped size) at offset 0 of the multicast object and map the
// object into this process: *mc_ptr addresses all GPUs' copies at once.  Call after EVERY rank has added its device.
ZOO_API int zoo_mc_bind(unsigned long long handle, void* local_ptr, long long bytes, void** mc_ptr) {
  if (!local_ptr || !mc_ptr || bytes <= 0) return ZOO_ERR_ARG;
  int dev = 0;
  ZOO_CUDA_CHECK(cudaGetDevice(&dev));
  auto bind = drv<PFN_mcBindMem>("cuMulticastBindMem");
  auto retain = drv<PFN_retain>("cuMemRetainAllocationHandle");
  auto release = drv<PFN_release>("cuMemRelease");
  if (!bind || !retain || !release) return fail("resolving cuMulticastBindMem", CUDA_ERROR_NOT_SUPPORTED, __LINE__);
  CUmemGenericAllocationHandle mem;
  CUresult r = retain(&mem, local_ptr);
  if (r != CUDA_SUCCESS) return fail("cuMemRetainAllocationHandle", r, __LINE__);
  r = bind((CUmemGenericAllocationHandle)handle, 0, mem, 0, (size_t)bytes, 0);
  release(mem);
  if (r != CUDA_SUCCESS) return fail("cuMulticastBindMem", r, __LINE__);
  CUmemAllocationProp prop = prop_for(dev);
  auto gran_fn = drv<PFN_gran>("cuMemGetAllocationGranularity");
  size_t gran = 0;
  if (!gran_fn || gran_fn(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || gran == 0) gran = 2u << 20;
  return map_handle((CUmemGenericAllocationHandle)handle, (size_t)bytes, gran, dev, mc_ptr);
}
