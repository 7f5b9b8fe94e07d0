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
