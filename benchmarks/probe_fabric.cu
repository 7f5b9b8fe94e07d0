// probe_fabric.cu — what the fabric of this box can do, measured before the kernels rely on it.
//
// One process drives two (or more) GPUs:
//   A. device attributes: VMM, POSIX-FD / fabric handle export, multicast (NVLS)
//   B. cuMemCreate + map on every GPU, export -> import of the POSIX file descriptor
//   C. multicast object over all GPUs, bound to the VMM allocations; multimem.ld_reduce /
//      multimem.st correctness and bandwidth against the number of CTAs
//   D. TMA tensor-map loads (the GEMM's operand path) from LOCAL and from PEER memory: a ring of
//      16 KB boxes per CTA, bandwidth against #CTAs and ring depth — decides whether the fused
//      all-gather GEMM can take its B tiles straight from the owner's replica
// Output: JSON lines on stdout.  Build: see benchmarks/probe_fabric.sh.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CU(x)                                                                    \
  do {                                                                           \
    CUresult r_ = (x);                                                           \
    if (r_ != CUDA_SUCCESS) {                                                    \
      const char* s_ = nullptr;                                                  \
      cuGetErrorString(r_, &s_);                                                 \
      printf("{\"error\": \"%s -> %d %s\", \"line\": %d}\n", #x, (int)r_, s_ ? s_ : "?", __LINE__); \
      fflush(stdout);                                                            \
      return 1;                                                                  \
    }                                                                            \
  } while (0)
#define RT(x)                                                                    \
  do {                                                                           \
    cudaError_t r_ = (x);                                                        \
    if (r_ != cudaSuccess) {                                                     \
      printf("{\"error\": \"%s -> %s\", \"line\": %d}\n", #x, cudaGetErrorString(r_), __LINE__); \
      fflush(stdout);                                                            \
      return 1;                                                                  \
    }                                                                            \
  } while (0)

// ------------------------------------------------------------------------------------ kernels
__global__ void fill_bf16(__nv_bfloat16* p, size_t n, float base) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    p[i] = __float2bfloat16_rn(base + (float)(i % 7));
}

// out[i] (16 bytes = 8 bf16) = sum over all GPUs of mc[i]
__global__ void __launch_bounds__(512) mc_ld_reduce(const uint4* mc, uint4* out, size_t nvec) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride * 4) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t j = i + u * stride;
      if (j < nvec)
        asm volatile(
            "multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
            : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w)
            : "l"(mc + j)
            : "memory");
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t j = i + u * stride;
      if (j < nvec) out[j] = v[u];
    }
  }
}

__global__ void __launch_bounds__(512) mc_st(uint4* mc, const uint4* src, size_t nvec) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const uint4 v = src[i];
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc + i),
                 "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)),
                 "f"(__uint_as_float(v.w))
                 : "memory");
  }
}

__global__ void __launch_bounds__(512) peer_ldg(const uint4* src, uint4* dst, size_t nvec) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride * 4) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < nvec) v[u] = __ldcs(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < nvec) dst[i + u * stride] = v[u];
  }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tW_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::
          "r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// Ring of `stages` 16 KB boxes per CTA fed by TMA tensor loads ({64 x 128} bf16, 128B swizzle, as
// the GEMM's operand tiles); thread 0 produces, thread 32 consumes (waits and frees the slot).
// With store_back the consumer writes the box to `dst` with a bulk store (persisting a gathered
// tile), so the slot is freed only when that store has read the shared memory.
constexpr int kBox = 64 * 128 * 2;
__global__ void __launch_bounds__(64) tma_ring(const __grid_constant__ CUtensorMap map, int rows,
                                               int cols, int stages, char* dst, int store_back) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* ring = reinterpret_cast<uint8_t*>(((uintptr_t)smem + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(ring + (size_t)stages * kBox);
  uint64_t* empty = full + stages;
  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; ++i) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[i])));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&empty[i])));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async;" ::: "memory");
  }
  __syncthreads();
  const int boxes_r = rows / 128, boxes_c = cols / 64;
  const long total = (long)boxes_r * boxes_c;
  if (threadIdx.x == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (long b = blockIdx.x; b < total; b += gridDim.x) {
      mbar_wait(&empty[stage], phase ^ 1);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(
                       smem_u32(&full[stage])),
                   "r"(kBox)
                   : "memory");
      const int c0 = (int)(b % boxes_c) * 64, r0 = (int)(b / boxes_c) * 128;
      asm volatile(
          "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, "
          "{%3, %4}], [%2];" ::"r"(smem_u32(ring + (size_t)stage * kBox)),
          "l"(&map), "r"(smem_u32(&full[stage])), "r"(c0), "r"(r0)
          : "memory");
      if (++stage == stages) stage = 0, phase ^= 1;
    }
  } else if (threadIdx.x == 32) {
    int stage = 0;
    uint32_t phase = 0;
    for (long b = blockIdx.x; b < total; b += gridDim.x) {
      mbar_wait(&full[stage], phase);
      if (store_back) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(
                         dst + b * (long)kBox),
                     "r"(smem_u32(ring + (size_t)stage * kBox)), "r"(kBox)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      }
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&empty[stage]))
                   : "memory");
      if (++stage == stages) stage = 0, phase ^= 1;
    }
    if (store_back) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

// ------------------------------------------------------------------------------------ host
struct Seg {
  CUmemGenericAllocationHandle h;
  CUdeviceptr va;
};

static float time_ms(cudaStream_t s, cudaEvent_t a, cudaEvent_t b) {
  float ms = 0;
  cudaEventSynchronize(b);
  cudaEventElapsedTime(&ms, a, b);
  return ms;
}

int main(int argc, char** argv) {
  CU(cuInit(0));
  int ndev = 0;
  RT(cudaGetDeviceCount(&ndev));
  int use = ndev < 2 ? ndev : (argc > 1 ? atoi(argv[1]) : 2);
  if (use > ndev) use = ndev;
  printf("{\"probe\": \"devices\", \"visible\": %d, \"used\": %d}\n", ndev, use);
  // ---- A. attributes
  bool mc_ok = true, fd_ok = true;
  for (int d = 0; d < use; ++d) {
    CUdevice dev;
    CU(cuDeviceGet(&dev, d));
    int vmm = 0, fd = 0, fab = 0, mc = 0;
    cuDeviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev);
    cuDeviceGetAttribute(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
    cuDeviceGetAttribute(&fab, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, dev);
    cuDeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
    printf("{\"probe\": \"attr\", \"dev\": %d, \"vmm\": %d, \"posix_fd\": %d, \"fabric_handle\": %d, "
           "\"multicast\": %d}\n", d, vmm, fd, fab, mc);
    mc_ok = mc_ok && mc;
    fd_ok = fd_ok && fd && vmm;
    RT(cudaSetDevice(d));
    RT(cudaFree(0));
    for (int p = 0; p < use; ++p)
      if (p != d) {
        int can = 0;
        cudaDeviceCanAccessPeer(&can, d, p);
        if (can) {
          cudaError_t e = cudaDeviceEnablePeerAccess(p, 0);
          if (e != cudaSuccess) cudaGetLastError();
        }
      }
  }
  fflush(stdout);
  const size_t bytes = 256ull << 20;
  // ---- B. VMM segments, mapped on every device
  std::vector<Seg> seg(use);
  size_t gran = 0;
  if (fd_ok) {
    for (int d = 0; d < use; ++d) {
      CUmemAllocationProp prop = {};
      prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
      prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
      prop.location.id = d;
      prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      CU(cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
      RT(cudaSetDevice(d));
      CU(cuMemCreate(&seg[d].h, bytes, &prop, 0));
      CU(cuMemAddressReserve(&seg[d].va, bytes, gran, 0, 0));
      CU(cuMemMap(seg[d].va, bytes, 0, seg[d].h, 0));
      std::vector<CUmemAccessDesc> acc(use);
      for (int p = 0; p < use; ++p) {
        acc[p].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc[p].location.id = p;
        acc[p].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
      }
      CU(cuMemSetAccess(seg[d].va, bytes, acc.data(), use));
    }
    int fdv = -1;
    CUresult r = cuMemExportToShareableHandle(&fdv, seg[0].h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    CUmemGenericAllocationHandle imp = 0;
    CUresult r2 = CUDA_ERROR_UNKNOWN;
    if (r == CUDA_SUCCESS)
      r2 = cuMemImportFromShareableHandle(&imp, (void*)(uintptr_t)fdv,
                                          CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    printf("{\"probe\": \"vmm\", \"granularity\": %zu, \"export_fd\": %d, \"import\": %d}\n", gran,
           (int)r, (int)r2);
    if (r2 == CUDA_SUCCESS) cuMemRelease(imp);
    if (fdv >= 0) close(fdv);
  } else {
    printf("{\"probe\": \"vmm\", \"skipped\": \"no VMM / POSIX fd support\"}\n");
    return 0;
  }
  fflush(stdout);
  for (int d = 0; d < use; ++d) {
    RT(cudaSetDevice(d));
    fill_bf16<<<592, 512>>>((__nv_bfloat16*)seg[d].va, bytes / 2, (float)(d + 1));
    RT(cudaDeviceSynchronize());
  }
  RT(cudaSetDevice(0));
  cudaStream_t s0;
  RT(cudaStreamCreate(&s0));
  cudaEvent_t e0, e1;
  RT(cudaEventCreate(&e0));
  RT(cudaEventCreate(&e1));
  uint4* scratch = nullptr;
  RT(cudaMalloc(&scratch, bytes));

  // ---- reference: 16-byte peer loads (the K1 / K2 path)
  if (use > 1) {
    for (int grid : {16, 32, 74, 148}) {
      peer_ldg<<<grid, 512, 0, s0>>>((const uint4*)seg[1].va, scratch, bytes / 16);
      RT(cudaEventRecord(e0, s0));
      for (int i = 0; i < 5; ++i)
        peer_ldg<<<grid, 512, 0, s0>>>((const uint4*)seg[1].va, scratch, bytes / 16);
      RT(cudaEventRecord(e1, s0));
      RT(cudaStreamSynchronize(s0));
      const float ms = time_ms(s0, e0, e1) / 5;
      printf("{\"probe\": \"peer_ldg\", \"grid\": %d, \"ms\": %.4f, \"GBps\": %.1f}\n", grid, ms,
             bytes / ms / 1e6);
    }
    fflush(stdout);
  }

  // ---- C. multicast
  if (mc_ok && use > 1) {
    CUmulticastObjectProp mp = {};
    mp.numDevices = use;
    mp.size = bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mgran = 0, mgran_rec = 0;
    CU(cuMulticastGetGranularity(&mgran, &mp, CU_MULTICAST_GRANULARITY_MINIMUM));
    CU(cuMulticastGetGranularity(&mgran_rec, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    CUmemGenericAllocationHandle mch;
    CUresult r = cuMulticastCreate(&mch, &mp);
    printf("{\"probe\": \"multicast_create\", \"rc\": %d, \"gran_min\": %zu, \"gran_rec\": %zu}\n",
           (int)r, mgran, mgran_rec);
    fflush(stdout);
    if (r == CUDA_SUCCESS) {
      for (int d = 0; d < use; ++d) {
        CUdevice dev;
        CU(cuDeviceGet(&dev, d));
        CU(cuMulticastAddDevice(mch, dev));
      }
      for (int d = 0; d < use; ++d) CU(cuMulticastBindMem(mch, 0, seg[d].h, 0, bytes, 0));
      CUdeviceptr mcva;
      CU(cuMemAddressReserve(&mcva, bytes, mgran_rec, 0, 0));
      CU(cuMemMap(mcva, bytes, 0, mch, 0));
      std::vector<CUmemAccessDesc> acc(use);
      for (int p = 0; p < use; ++p) {
        acc[p].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc[p].location.id = p;
        acc[p].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
      }
      CU(cuMemSetAccess(mcva, bytes, acc.data(), use));
      int mfd = -1;
      CUresult re = cuMemExportToShareableHandle(&mfd, mch, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
      printf("{\"probe\": \"multicast_export_fd\", \"rc\": %d}\n", (int)re);
      if (mfd >= 0) close(mfd);
      // correctness of ld_reduce: element i holds (d+1) + i%7 on GPU d
      mc_ld_reduce<<<148, 512, 0, s0>>>((const uint4*)mcva, scratch, bytes / 16);
      RT(cudaStreamSynchronize(s0));
      std::vector<__nv_bfloat16> host(4096);
      RT(cudaMemcpy(host.data(), scratch, host.size() * 2, cudaMemcpyDeviceToHost));
      int bad = 0;
      for (size_t i = 0; i < host.size(); ++i) {
        float exp = 0;
        for (int d = 0; d < use; ++d) exp += (float)(d + 1) + (float)(i % 7);
        if (fabsf(__bfloat162float(host[i]) - exp) > 0.51f) ++bad;
      }
      printf("{\"probe\": \"multimem_ld_reduce_check\", \"bad\": %d, \"of\": %zu}\n", bad, host.size());
      for (int grid : {8, 16, 32, 74, 148}) {
        RT(cudaEventRecord(e0, s0));
        for (int i = 0; i < 5; ++i)
          mc_ld_reduce<<<grid, 512, 0, s0>>>((const uint4*)mcva, scratch, bytes / 16);
        RT(cudaEventRecord(e1, s0));
        RT(cudaStreamSynchronize(s0));
        const float ms = time_ms(s0, e0, e1) / 5;
        printf("{\"probe\": \"multimem_ld_reduce\", \"grid\": %d, \"ms\": %.4f, \"out_GBps\": %.1f}\n",
               grid, ms, bytes / ms / 1e6);
      }
      // multimem.st: scratch -> every GPU's segment
      RT(cudaMemset(scratch, 0x3c, bytes));
      for (int grid : {8, 16, 32, 74, 148}) {
        RT(cudaEventRecord(e0, s0));
        for (int i = 0; i < 5; ++i) mc_st<<<grid, 512, 0, s0>>>((uint4*)mcva, scratch, bytes / 16);
        RT(cudaEventRecord(e1, s0));
        RT(cudaStreamSynchronize(s0));
        const float ms = time_ms(s0, e0, e1) / 5;
        printf("{\"probe\": \"multimem_st\", \"grid\": %d, \"ms\": %.4f, \"GBps\": %.1f}\n", grid, ms,
               bytes / ms / 1e6);
      }
      int bad_st = 0;
      for (int d = 0; d < use; ++d) {
        std::vector<uint8_t> hb(4096);
        RT(cudaMemcpy(hb.data(), (void*)(seg[d].va + bytes - 4096), 4096, cudaMemcpyDeviceToHost));
        for (auto b : hb) bad_st += b != 0x3c;
      }
      printf("{\"probe\": \"multimem_st_check\", \"bad\": %d}\n", bad_st);
      fflush(stdout);
    }
  } else {
    printf("{\"probe\": \"multicast\", \"skipped\": \"attribute says unsupported or one GPU\"}\n");
  }

  // ---- D. TMA tensor loads from local / peer memory
  {
    const int cols = 4096, rows = (int)(bytes / 2 / cols);
    RT(cudaFuncSetAttribute(tma_ring, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    for (int src = 0; src < (use > 1 ? 2 : 1); ++src) {
      CUtensorMap map;
      cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
      cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
      cuuint32_t box[2] = {64, 128};
      cuuint32_t elem[2] = {1, 1};
      CU(cuTensorMapEncodeTiled(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)seg[src].va, dims,
                                strides, box, elem, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE));
      for (int store_back = 0; store_back < 2; ++store_back)
        for (int stages : {4, 8, 12})
          for (int grid : {8, 16, 32, 74, 148}) {
            const size_t smem = (size_t)stages * kBox + 1024 + 2 * stages * 8 + 64;
            tma_ring<<<grid, 64, smem, s0>>>(map, rows, cols, stages, (char*)scratch, store_back);
            RT(cudaEventRecord(e0, s0));
            for (int i = 0; i < 3; ++i)
              tma_ring<<<grid, 64, smem, s0>>>(map, rows, cols, stages, (char*)scratch, store_back);
            RT(cudaEventRecord(e1, s0));
            RT(cudaStreamSynchronize(s0));
            const float ms = time_ms(s0, e0, e1) / 3;
            printf("{\"probe\": \"tma_ring\", \"src\": \"%s\", \"store_back\": %d, \"stages\": %d, "
                   "\"grid\": %d, \"ms\": %.4f, \"GBps\": %.1f}\n",
                   src ? "peer" : "local", store_back, stages, grid, ms, bytes / ms / 1e6);
          }
      fflush(stdout);
    }
  }
  printf("{\"probe\": \"done\"}\n");
  return 0;
}
