// push_tma.cu — K1, bulk-copy variant: the fused gradient push on a handful of SMs.
//
// The register-path kernel in push_pull.cu needs ~148 CTAs of 512 threads to keep enough
// 16-byte peer loads in flight to fill NVLink, and while it is resident no compute CTA fits on
// those SMs — "overlap" with backward is then time-slicing.  Here ONE producer thread per CTA
// keeps a shared-memory ring full with cp.async.bulk copies (UBLKCP) of the N peer gradient
// chunks plus the local master / m / v chunks, and eight consumer warps reduce, update and
// store.  ~170 KB are in flight per SM, so 16-32 SMs saturate the link and the other ~120 SMs
// keep running cuBLAS/cuDNN backward kernels.  Same protocol, flags and results as
// push_fused_kernel (selected with PSB_PUSH_TMA); REDUCE_ONLY / FROM_REDUCED stay on the
// register path.
//
// Reference paths replaced: see push_pull.cu.
#include "common.cuh"

namespace psb {
namespace tma_push {

constexpr int kConsumers = 256;            // 8 warps, 8 elements per thread per pass
constexpr int kThreads = kConsumers + 32;  // + producer warp
constexpr int kChunk = kConsumers * 8;     // elements per stage
constexpr int kMaxStages = 8;
constexpr int kSmemBudget = 200 * 1024;

__device__ __forceinline__ uint4 lds16(const void* p) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(smem_u32(p)));
  return r;
}
__device__ __forceinline__ void lds8f(const float* p, float f[8]) {
  uint4 a = lds16(p), b = lds16(p + 4);
  f[0] = __uint_as_float(a.x), f[1] = __uint_as_float(a.y), f[2] = __uint_as_float(a.z);
  f[3] = __uint_as_float(a.w), f[4] = __uint_as_float(b.x), f[5] = __uint_as_float(b.y);
  f[6] = __uint_as_float(b.z), f[7] = __uint_as_float(b.w);
}
__device__ __forceinline__ void stg8f(float* p, const float f[8]) {
  st_stream_f8(p, f);
}

// Stage layout: [grad of peer 0 .. world-1][master][m][v], each kChunk elements.
template <int KIND, int GDT>
__global__ void __launch_bounds__(kThreads, 1) push_fused_tma_kernel(const PsbPush a, int stages) {
  constexpr int GSZ = GDT == PSB_BF16 ? 2 : 4;
  constexpr bool HAS_M = KIND != PSB_OPT_SGD;
  constexpr bool HAS_V = KIND == PSB_OPT_ADAM || KIND == PSB_OPT_ADAMW;
  extern __shared__ __align__(128) uint8_t smem[];
  const int rank = a.fabric.rank, world = a.fabric.world;
  const int flags = a.flags;
  const uint32_t e = *reinterpret_cast<volatile const uint32_t*>(a.epoch) + 1u;
  const uint32_t mask = (flags & PSB_PUSH_MASKED) ? *a.peer_mask : 0xffffffffu;
  if ((flags & PSB_PUSH_MASKED) && (mask & ((1u << world) - 1u)) == 0u) return;

  const int grad_bytes = kChunk * GSZ;
  const int stage_bytes = world * grad_bytes + kChunk * 4 * (1 + (HAS_M ? 1 : 0) + (HAS_V ? 1 : 0));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty = full + kMaxStages;
  uint8_t* ring = smem + 128;
  __shared__ int is_last;

  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], kConsumers / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // ---- protocol steps 1-2 (same as the register kernel)
  if (!(flags & PSB_PUSH_NO_WAIT)) {
    if (blockIdx.x == 0 && threadIdx.x < world) {
      fence_sys();
      st_release_sys(a.fabric.pad[threadIdx.x] + PSB_PAD_READY_OFF + a.bucket * PSB_MAX_PEERS + rank,
                     e);
    }
    if (threadIdx.x < world && ((mask >> threadIdx.x) & 1u))
      wait_ge_sys(a.fabric.pad[rank] + PSB_PAD_READY_OFF + a.bucket * PSB_MAX_PEERS + threadIdx.x, e);
  }
  __syncthreads();

  const size_t slice = a.slice_elems;
  const size_t base = (size_t)rank * slice;
  const size_t nchunks = (slice + kChunk - 1) / kChunk;
  const int warp = threadIdx.x >> 5;

  if (warp == kConsumers / 32) {
    // =============== producer: one thread feeds the ring ===============
    if ((threadIdx.x & 31) == 0) {
      asm volatile("fence.proxy.async;" ::: "memory");  // flag acquire (generic) -> bulk reads
      int stage = 0;
      uint32_t phase = 0;
      for (size_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const size_t e0 = c * kChunk;
        const size_t n = slice - e0 < (size_t)kChunk ? slice - e0 : (size_t)kChunk;
        mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* st = ring + (size_t)stage * stage_bytes;
        uint32_t tx = 0;
        for (int w = 0; w < world; ++w)
          if ((mask >> w) & 1u) tx += (uint32_t)(n * GSZ);
        tx += (uint32_t)(n * 4) * (1 + (HAS_M ? 1 : 0) + (HAS_V ? 1 : 0));
        mbar_expect_tx(&full[stage], tx);
        for (int w = 0; w < world; ++w) {
          if (!((mask >> w) & 1u)) continue;
          const char* g = reinterpret_cast<const char*>(a.grad[w]) +
                          (a.grad_off ? a.grad_off[w] : 0ull) + (base + e0) * GSZ;
          bulk_load(st + (size_t)w * grad_bytes, g, (uint32_t)(n * GSZ), &full[stage]);
        }
        uint8_t* sp = st + (size_t)world * grad_bytes;
        bulk_load(sp, a.master + e0, (uint32_t)(n * 4), &full[stage]);
        if (HAS_M) bulk_load(sp + kChunk * 4, a.m + e0, (uint32_t)(n * 4), &full[stage]);
        if (HAS_V) bulk_load(sp + kChunk * 8, a.v + e0, (uint32_t)(n * 4), &full[stage]);
        if (++stage == stages) stage = 0, phase ^= 1;
      }
    }
  } else {
    // =============== consumers: reduce + update + store ===============
    const OptimCoef cf = make_coef(a.opt, a.step ? *a.step : 0u);
    const float gscale = a.opt.grad_scale * (a.clip_coef ? *a.clip_coef : 1.f);
    int stage = 0;
    uint32_t phase = 0;
    for (size_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
      const size_t e0 = c * kChunk;
      const size_t n = slice - e0 < (size_t)kChunk ? slice - e0 : (size_t)kChunk;
      mbar_wait(&full[stage], phase);
      const uint8_t* st = ring + (size_t)stage * stage_bytes;
      const size_t le = (size_t)threadIdx.x * 8;  // element inside the chunk
      if (le < n) {                               // slices are multiples of 512: n % 8 == 0
        float g[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] = 0.f;
        for (int w = 0; w < world; ++w) {  // fixed order => deterministic sum
          if (!((mask >> w) & 1u)) continue;
          const uint8_t* gp = st + (size_t)w * grad_bytes + le * GSZ;
          if (GDT == PSB_BF16) {
            float f[8];
            unpack8(lds16(gp), f);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] += f[k];
          } else {
            float f[8];
            lds8f(reinterpret_cast<const float*>(gp), f);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] += f[k];
          }
        }
        const float* sp = reinterpret_cast<const float*>(st + (size_t)world * grad_bytes);
        float p[8], mm[8] = {}, vv[8] = {};
        lds8f(sp + le, p);
        if (HAS_M) lds8f(sp + kChunk + le, mm);
        if (HAS_V) lds8f(sp + 2 * kChunk + le, vv);
#pragma unroll
        for (int k = 0; k < 8; ++k) update1<KIND>(cf, p[k], mm[k], vv[k], g[k] * gscale);
        stg8f(a.master + e0 + le, p);
        if (HAS_M) stg8f(a.m + e0 + le, mm);
        if (HAS_V) stg8f(a.v + e0 + le, vv);
        const uint4 packed = pack8(p);
        if (flags & PSB_PUSH_BROADCAST) {
          for (int w = 0; w < world; ++w)
            st_stream_16(reinterpret_cast<__nv_bfloat16*>(a.param[w]) + base + e0 + le, packed);
        } else {
          st_stream_16(reinterpret_cast<__nv_bfloat16*>(a.param[rank]) + base + e0 + le, packed);
        }
      }
      __syncwarp();
      if ((threadIdx.x & 31) == 0) mbar_arrive(&empty[stage]);  // this warp is done with the stage
      if (++stage == stages) stage = 0, phase ^= 1;
    }
  }

  // ---- protocol step 4: last CTA publishes the new version
  __syncthreads();
  if (threadIdx.x == 0) {
    fence_sys();
    is_last = (atomicAdd(a.done_counter, 1u) == gridDim.x - 1);
    if (is_last) fence_sys();
  }
  __syncthreads();
  if (is_last) {
    if (threadIdx.x < world)
      st_release_sys(a.fabric.pad[threadIdx.x] + PSB_PAD_VERSION_OFF + a.bucket * PSB_MAX_PEERS + rank,
                     (flags & PSB_PUSH_SEQLOCK) ? 2u * e : e);
    if (threadIdx.x == 0) {
      *a.epoch = e;
      *a.done_counter = 0u;
    }
  }
}

template <int KIND, int GDT>
static int launch(const PsbPush& a, cudaStream_t s) {
  const int gsz = GDT == PSB_BF16 ? 2 : 4;
  const int nstate = 1 + (KIND != PSB_OPT_SGD ? 1 : 0) +
                     ((KIND == PSB_OPT_ADAM || KIND == PSB_OPT_ADAMW) ? 1 : 0);
  const int stage_bytes = a.fabric.world * kChunk * gsz + kChunk * 4 * nstate;
  int stages = kSmemBudget / stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  PSB_REQUIRE(stages >= 2, PSB_EUNSUPPORTED, "push_fused_tma: stage of %d bytes leaves no ring",
              stage_bytes);
  const int smem = 128 + stages * stage_bytes;
  auto kern = push_fused_tma_kernel<KIND, GDT>;
  PSB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const size_t nchunks = (a.slice_elems + kChunk - 1) / kChunk;
  int grid = a.grid > 0 ? a.grid : 24;
  if ((size_t)grid > nchunks) grid = (int)(nchunks ? nchunks : 1);
  if (grid > kNumSMs) grid = kNumSMs;
  kern<<<grid, kThreads, smem, s>>>(a, stages);
  PSB_LAUNCH_OK("push_fused_tma");
  return 0;
}

}  // namespace tma_push

// called by psb_push_fused when PSB_PUSH_TMA is set (arguments already validated there)
int push_fused_tma_dispatch(const PsbPush& a, cudaStream_t s) {
#define PSB_DISPATCH(K)                                                          \
  case K:                                                                        \
    return a.grad_dtype == PSB_BF16 ? tma_push::launch<K, PSB_BF16>(a, s)         \
                                    : tma_push::launch<K, PSB_F32>(a, s);
  switch (a.opt.kind) {
    PSB_DISPATCH(PSB_OPT_SGD)
    PSB_DISPATCH(PSB_OPT_MOMENTUM)
    PSB_DISPATCH(PSB_OPT_ADAM)
    PSB_DISPATCH(PSB_OPT_ADAMW)
  }
#undef PSB_DISPATCH
  set_error("push_fused_tma: unknown optimizer kind %d", a.opt.kind);
  return PSB_EINVAL;
}

}  // namespace psb

PSB_EXPORT_SPIN_LIMIT(push_tma)
