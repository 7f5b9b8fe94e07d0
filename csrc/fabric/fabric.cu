// fabric.cu — host side of the symmetric heap: device memory that every rank of the box
// maps into its own address space so kernels can issue NVLink peer loads / stores.
//
// One process per GPU.  Each rank allocates its segment with cudaMalloc, exports a CUDA IPC
// handle, the 64-byte handles travel over the control plane (torch.distributed store or the
// coordinator's ExchangeHandles extension RPC) and every peer opens them.  This replaces
// the reference's single-process NCCL clique (src/nccl_manager.cpp:28-85) and its
// per-iteration cudaMalloc / cudaFree (src/worker.cpp:420-443): memory is allocated once.
#include <cstring>

#include "../kernels/common.cuh"

using namespace psb;

extern "C" {

int psb_device_count(int* count) {
  PSB_REQUIRE(count, PSB_EINVAL, "psb_device_count: null argument");
  cudaError_t e = cudaGetDeviceCount(count);
  if (e != cudaSuccess) {
    *count = 0;
    cudaGetLastError();
  }
  return 0;
}

int psb_set_device(int dev) {
  PSB_CUDA_OK(cudaSetDevice(dev));
  return 0;
}

int psb_heap_alloc(size_t bytes, void** ptr) {
  PSB_REQUIRE(ptr && bytes > 0, PSB_EINVAL, "psb_heap_alloc: bad argument");
  PSB_CUDA_OK(cudaMalloc(ptr, bytes));
  PSB_CUDA_OK(cudaMemset(*ptr, 0, bytes));
  PSB_CUDA_OK(cudaDeviceSynchronize());
  return 0;
}

int psb_heap_free(void* ptr) {
  if (ptr) PSB_CUDA_OK(cudaFree(ptr));
  return 0;
}

int psb_ipc_export(void* ptr, unsigned char handle[PSB_IPC_HANDLE_BYTES]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == PSB_IPC_HANDLE_BYTES, "IPC handle size");
  cudaIpcMemHandle_t h;
  PSB_CUDA_OK(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle, &h, sizeof(h));
  return 0;
}

int psb_ipc_open(const unsigned char handle[PSB_IPC_HANDLE_BYTES], void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  PSB_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int psb_ipc_close(void* ptr) {
  if (ptr) PSB_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

int psb_can_access_peer(int dev, int peer, int* ok) {
  PSB_REQUIRE(ok, PSB_EINVAL, "psb_can_access_peer: null argument");
  PSB_CUDA_OK(cudaDeviceCanAccessPeer(ok, dev, peer));
  return 0;
}

int psb_enable_peer_access(int peer) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  PSB_CUDA_OK(e);
  return 0;
}

int psb_memcpy_d2h(void* dst_host, const void* src_dev, size_t bytes, psb_stream_t stream) {
  PSB_CUDA_OK(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost,
                              (cudaStream_t)stream));
  return 0;
}

int psb_memcpy_h2d(void* dst_dev, const void* src_host, size_t bytes, psb_stream_t stream) {
  PSB_CUDA_OK(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice,
                              (cudaStream_t)stream));
  return 0;
}

int psb_memset(void* dst_dev, int value, size_t bytes, psb_stream_t stream) {
  PSB_CUDA_OK(cudaMemsetAsync(dst_dev, value, bytes, (cudaStream_t)stream));
  return 0;
}

int psb_stream_sync(psb_stream_t stream) {
  PSB_CUDA_OK(cudaStreamSynchronize((cudaStream_t)stream));
  return 0;
}

int psb_device_sync(void) {
  PSB_CUDA_OK(cudaDeviceSynchronize());
  return 0;
}

}  // extern "C"

// =====================================================================================
// VMM heap + NVLS multicast.
//
// The cudaMalloc + cudaIpc* heap above is enough for peer loads / stores, but such memory
// cannot be bound to a multicast object.  The VMM path allocates the segment with cuMemCreate
// (shareable as a POSIX file descriptor — the descriptors travel between the ranks over a Unix
// socket with SCM_RIGHTS, parallel/fabric.py), maps every peer's segment, and — where the
// device reports CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED — binds all segments to ONE multicast
// object whose mapping `mc` addresses the same offset of every rank's segment at once:
// multimem.ld_reduce on mc + off returns the sum over all ranks computed inside the NVSwitch,
// multimem.st writes all replicas with one store (push_nvls.cu).
// The driver API is reached through cudaGetDriverEntryPoint so that the library keeps loading
// on a box without libcuda.
// =====================================================================================
#include <cuda.h>

#include <mutex>

namespace {

struct Drv {
  CUresult (*GetErrorString)(CUresult, const char**) = nullptr;
  CUresult (*DeviceGet)(CUdevice*, int) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*,
                        unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle,
                     unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*,
                                          CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle,
                                         CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*,
                                           CUmemAllocationHandleType) = nullptr;
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle,
                               size_t, size_t, unsigned long long) = nullptr;
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*,
                                      CUmulticastGranularity_flags) = nullptr;
  bool ok = false;
};

template <class F>
bool load_sym(F& fn, const char* name) {
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &sym, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || !sym) {
    cudaGetLastError();
    return false;
  }
  fn = reinterpret_cast<F>(sym);
  return true;
}

Drv& drv() {
  static Drv d;
  static std::once_flag once;
  std::call_once(once, [] {
    bool ok = true;
    ok &= load_sym(d.GetErrorString, "cuGetErrorString");
    ok &= load_sym(d.DeviceGet, "cuDeviceGet");
    ok &= load_sym(d.DeviceGetAttribute, "cuDeviceGetAttribute");
    ok &= load_sym(d.MemCreate, "cuMemCreate");
    ok &= load_sym(d.MemRelease, "cuMemRelease");
    ok &= load_sym(d.MemAddressReserve, "cuMemAddressReserve");
    ok &= load_sym(d.MemAddressFree, "cuMemAddressFree");
    ok &= load_sym(d.MemMap, "cuMemMap");
    ok &= load_sym(d.MemUnmap, "cuMemUnmap");
    ok &= load_sym(d.MemSetAccess, "cuMemSetAccess");
    ok &= load_sym(d.MemGetAllocationGranularity, "cuMemGetAllocationGranularity");
    ok &= load_sym(d.MemExportToShareableHandle, "cuMemExportToShareableHandle");
    ok &= load_sym(d.MemImportFromShareableHandle, "cuMemImportFromShareableHandle");
    d.ok = ok;
    // multicast entry points are optional (older drivers): checked where they are used
    load_sym(d.MulticastCreate, "cuMulticastCreate");
    load_sym(d.MulticastAddDevice, "cuMulticastAddDevice");
    load_sym(d.MulticastBindMem, "cuMulticastBindMem");
    load_sym(d.MulticastUnbind, "cuMulticastUnbind");
    load_sym(d.MulticastGetGranularity, "cuMulticastGetGranularity");
  });
  return d;
}

#define PSB_CU_OK(expr)                                                               \
  do {                                                                                \
    CUresult _r = (expr);                                                             \
    if (_r != CUDA_SUCCESS) {                                                         \
      const char* _s = nullptr;                                                       \
      if (drv().GetErrorString) drv().GetErrorString(_r, &_s);                        \
      psb::set_error("%s failed: %s (CUresult %d, %s:%d)", #expr, _s ? _s : "?", (int)_r, \
                     __FILE__, __LINE__);                                             \
      return (int)_r;                                                                 \
    }                                                                                 \
  } while (0)

int current_device(CUdevice* dev, int* ordinal) {
  int d = 0;
  PSB_CUDA_OK(cudaGetDevice(&d));
  PSB_CUDA_OK(cudaFree(0));  // make sure the primary context exists
  PSB_CU_OK(drv().DeviceGet(dev, d));
  if (ordinal) *ordinal = d;
  return 0;
}

CUmemAllocationProp device_prop(int ordinal) {
  CUmemAllocationProp prop = {};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = ordinal;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

int map_rw(CUmemGenericAllocationHandle h, size_t bytes, size_t align, int ordinal, void** ptr) {
  CUdeviceptr va = 0;
  PSB_CU_OK(drv().MemAddressReserve(&va, bytes, align, 0, 0));
  CUresult r = drv().MemMap(va, bytes, 0, h, 0);
  if (r != CUDA_SUCCESS) {
    drv().MemAddressFree(va, bytes);
    PSB_CU_OK(r);
  }
  CUmemAccessDesc acc = {};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = ordinal;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = drv().MemSetAccess(va, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) {
    drv().MemUnmap(va, bytes);
    drv().MemAddressFree(va, bytes);
    PSB_CU_OK(r);
  }
  *ptr = reinterpret_cast<void*>(va);
  return 0;
}

}  // namespace

extern "C" {

// What the current device offers: *vmm = cuMemCreate with POSIX-fd export, *multicast = NVLS
// multicast objects.  Both 0 when the driver entry points are missing.
int psb_vmm_support(int* vmm, int* multicast) {
  PSB_REQUIRE(vmm && multicast, PSB_EINVAL, "psb_vmm_support: null argument");
  *vmm = *multicast = 0;
  Drv& d = drv();
  if (!d.ok) return 0;
  CUdevice dev;
  if (int rc = current_device(&dev, nullptr)) return rc;
  int a = 0, b = 0, c = 0;
  d.DeviceGetAttribute(&a, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev);
  d.DeviceGetAttribute(&b, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
  d.DeviceGetAttribute(&c, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  *vmm = a && b;
  *multicast = *vmm && c && d.MulticastCreate && d.MulticastAddDevice && d.MulticastBindMem &&
               d.MulticastGetGranularity;
  return 0;
}

// Granularity to round a heap of `world` ranks up to: the allocation granularity, or — when a
// multicast object will be bound — the multicast one (`recommended` != 0: the size the driver
// recommends for performance, 512 MB on B200; else the minimum, 2 MB).
int psb_vmm_granularity(int world, int multicast, int recommended, size_t* gran) {
  PSB_REQUIRE(gran && world >= 1, PSB_EINVAL, "psb_vmm_granularity: bad argument");
  PSB_REQUIRE(drv().ok, PSB_EUNSUPPORTED, "VMM driver entry points are not available");
  CUdevice dev;
  int ord = 0;
  if (int rc = current_device(&dev, &ord)) return rc;
  CUmemAllocationProp prop = device_prop(ord);
  size_t g = 0;
  PSB_CU_OK(drv().MemGetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (multicast) {
    CUmulticastObjectProp mp = {};
    mp.numDevices = world;
    mp.size = g;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    PSB_CU_OK(drv().MulticastGetGranularity(
        &mg, &mp, recommended ? CU_MULTICAST_GRANULARITY_RECOMMENDED : CU_MULTICAST_GRANULARITY_MINIMUM));
    if (mg > g) g = mg;
  }
  *gran = g;
  return 0;
}

// Segment of the current device: physical allocation, mapped read-write locally and zeroed.
// *handle is the generic allocation handle (kept for multicast binding and the release),
// *fd a POSIX file descriptor that another process imports with psb_vmm_import.
int psb_vmm_alloc(size_t bytes, void** ptr, uint64_t* handle, int* fd) {
  PSB_REQUIRE(ptr && handle && fd && bytes > 0, PSB_EINVAL, "psb_vmm_alloc: bad argument");
  PSB_REQUIRE(drv().ok, PSB_EUNSUPPORTED, "VMM driver entry points are not available");
  CUdevice dev;
  int ord = 0;
  if (int rc = current_device(&dev, &ord)) return rc;
  CUmemAllocationProp prop = device_prop(ord);
  size_t g = 0;
  PSB_CU_OK(drv().MemGetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  PSB_REQUIRE(bytes % g == 0, PSB_EINVAL, "psb_vmm_alloc: %zu bytes is not a multiple of the granularity %zu",
              bytes, g);
  CUmemGenericAllocationHandle h;
  PSB_CU_OK(drv().MemCreate(&h, bytes, &prop, 0));
  if (int rc = map_rw(h, bytes, g, ord, ptr)) {
    drv().MemRelease(h);
    return rc;
  }
  PSB_CUDA_OK(cudaMemset(*ptr, 0, bytes));
  PSB_CUDA_OK(cudaDeviceSynchronize());
  int out_fd = -1;
  PSB_CU_OK(drv().MemExportToShareableHandle(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle = (uint64_t)h;
  *fd = out_fd;
  return 0;
}

// Map a peer's segment (its descriptor received over the Unix socket) into this process.
int psb_vmm_import(int fd, size_t bytes, void** ptr, uint64_t* handle) {
  PSB_REQUIRE(ptr && handle && fd >= 0 && bytes > 0, PSB_EINVAL, "psb_vmm_import: bad argument");
  PSB_REQUIRE(drv().ok, PSB_EUNSUPPORTED, "VMM driver entry points are not available");
  CUdevice dev;
  int ord = 0;
  if (int rc = current_device(&dev, &ord)) return rc;
  CUmemGenericAllocationHandle h;
  PSB_CU_OK(drv().MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd,
                                               CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  CUmemAllocationProp prop = device_prop(ord);
  size_t g = 0;
  PSB_CU_OK(drv().MemGetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (int rc = map_rw(h, bytes, g, ord, ptr)) {
    drv().MemRelease(h);
    return rc;
  }
  *handle = (uint64_t)h;
  return 0;
}

// Unmap a range mapped by psb_vmm_alloc / psb_vmm_import / psb_mc_bind_map and drop the handle.
int psb_vmm_release(void* ptr, size_t bytes, uint64_t handle) {
  if (ptr) {
    PSB_CU_OK(drv().MemUnmap((CUdeviceptr)ptr, bytes));
    PSB_CU_OK(drv().MemAddressFree((CUdeviceptr)ptr, bytes));
  }
  if (handle) PSB_CU_OK(drv().MemRelease((CUmemGenericAllocationHandle)handle));
  return 0;
}

// Multicast object over `world` devices (one rank creates it and passes the descriptor on).
int psb_mc_create(size_t bytes, int world, uint64_t* mc_handle, int* fd) {
  PSB_REQUIRE(mc_handle && fd && world >= 2, PSB_EINVAL, "psb_mc_create: bad argument");
  PSB_REQUIRE(drv().ok && drv().MulticastCreate, PSB_EUNSUPPORTED, "multicast is not available");
  CUmulticastObjectProp mp = {};
  mp.numDevices = world;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  PSB_CU_OK(drv().MulticastCreate(&h, &mp));
  int out_fd = -1;
  PSB_CU_OK(drv().MemExportToShareableHandle(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *mc_handle = (uint64_t)h;
  *fd = out_fd;
  return 0;
}

int psb_mc_import(int fd, uint64_t* mc_handle) {
  PSB_REQUIRE(mc_handle && fd >= 0, PSB_EINVAL, "psb_mc_import: bad argument");
  PSB_REQUIRE(drv().ok && drv().MulticastCreate, PSB_EUNSUPPORTED, "multicast is not available");
  CUmemGenericAllocationHandle h;
  PSB_CU_OK(drv().MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd,
                                               CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *mc_handle = (uint64_t)h;
  return 0;
}

// Step 1 (every rank, then a barrier): add the current device to the team.
int psb_mc_add_device(uint64_t mc_handle) {
  CUdevice dev;
  if (int rc = current_device(&dev, nullptr)) return rc;
  PSB_CU_OK(drv().MulticastAddDevice((CUmemGenericAllocationHandle)mc_handle, dev));
  return 0;
}

// Step 2 (every rank, after ALL devices were added): bind the local segment at offset 0 and map
// the multicast object read-write.  *mc_ptr + off then addresses off of every rank's segment.
int psb_mc_bind_map(uint64_t mc_handle, uint64_t local_handle, size_t bytes, void** mc_ptr) {
  PSB_REQUIRE(mc_ptr, PSB_EINVAL, "psb_mc_bind_map: null argument");
  CUdevice dev;
  int ord = 0;
  if (int rc = current_device(&dev, &ord)) return rc;
  PSB_CU_OK(drv().MulticastBindMem((CUmemGenericAllocationHandle)mc_handle, 0,
                                   (CUmemGenericAllocationHandle)local_handle, 0, bytes, 0));
  CUmulticastObjectProp mp = {};
  mp.numDevices = 2;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t mg = 0;
  PSB_CU_OK(drv().MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_MINIMUM));
  return map_rw((CUmemGenericAllocationHandle)mc_handle, bytes, mg, ord, mc_ptr);
}

int psb_mc_unbind(uint64_t mc_handle, size_t bytes) {
  CUdevice dev;
  if (int rc = current_device(&dev, nullptr)) return rc;
  if (drv().MulticastUnbind)
    PSB_CU_OK(drv().MulticastUnbind((CUmemGenericAllocationHandle)mc_handle, dev, 0, bytes));
  return 0;
}

}  // extern "C"
