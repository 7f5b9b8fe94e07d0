// push_nvls.cu — K1 / K2 over NVLS: the NVSwitch does the reduction and the replication.
//
// Same protocol as push_fused_kernel (push_pull.cu): READY flags -> reduce the owned slice over
// all workers -> server-side optimizer on the fp32 master shard -> refreshed bf16 slice into
// every replica -> VERSION flags.  The data moves differently.  The gradient space and the
// parameter replicas live in a VMM heap whose segments are all bound to one multicast object
// (fabric.cu): an address `mc + off` names offset `off` of EVERY rank's segment.
//   * reduce:    multimem.ld_reduce.add.acc::f32.v4.bf16x2 [grad_mc + i]  — one load returns
//                the sum of the N ranks' 8 gradient values, accumulated in fp32 inside the
//                switch and rounded to bf16 once.  NVLink ingress of the owner: slice bytes,
//                instead of (N-1) x slice bytes with peer loads.
//   * broadcast: multimem.st [param_mc + i] — one store, the switch writes all N replicas.
//                Egress of the owner: slice bytes instead of (N-1) x slice.
// At N = 8 a bucket of S bytes costs each GPU S per direction instead of 1.75 S, and the kernel
// issues one load + one store per 16 bytes instead of 8 + 8, so a handful of CTAs is enough.
// The sum is rounded to bf16 before the optimizer sees it (the peer-load kernel keeps it in
// fp32): results agree to bf16 tolerance, not bit for bit — the peer-load kernel stays the
// deterministic default, this one is selected with PSB_PUSH_NVLS.
//
// Reference path replaced: the worker's ncclAllReduce (/root/reference/src/nccl_manager.cpp:
// 102-121), which on an NVSwitch box runs NCCL's own NVLS kernels; here the reduction is fused
// with the server update (src/parameter_server.cpp:49-60, 83-90) in one kernel.
#include "common.cuh"

namespace psb {
namespace nvls {

constexpr int kThreads = 512;

__device__ __forceinline__ uint4 mc_ld_reduce_bf16x8(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ void mc_st_16(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc),
               "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)),
               "f"(__uint_as_float(v.w))
               : "memory");
}

template <int KIND>
__global__ void __launch_bounds__(kThreads, 1) push_nvls_kernel(const PsbPush a) {
  constexpr int U = KIND == PSB_OPT_SGD ? 4 : 2;
  const int rank = a.fabric.rank, world = a.fabric.world;
  const int flags = a.flags;
  uint32_t* const my_pad = a.fabric.pad[rank];
  const uint32_t e = *reinterpret_cast<volatile const uint32_t*>(a.epoch) + 1u;

  if (!(flags & PSB_PUSH_NO_WAIT)) {
    if (blockIdx.x == 0 && threadIdx.x < world) {
      fence_sys();
      st_release_sys(a.fabric.pad[threadIdx.x] + PSB_PAD_READY_OFF + a.bucket * PSB_MAX_PEERS + rank,
                     e);
    }
    if (threadIdx.x < world)
      wait_ge_sys(my_pad + PSB_PAD_READY_OFF + a.bucket * PSB_MAX_PEERS + threadIdx.x, e);
    __syncthreads();
  }

  const OptimCoef c = make_coef(a.opt, a.step ? *a.step : 0u);
  const float gscale = a.opt.grad_scale * (a.clip_coef ? *a.clip_coef : 1.f);
  const size_t nvec = a.slice_elems / 8;
  const size_t base = (size_t)rank * a.slice_elems;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const __nv_bfloat16* gmc = reinterpret_cast<const __nv_bfloat16*>(a.grad_mc) + base;
  __nv_bfloat16* pmc = reinterpret_cast<__nv_bfloat16*>(a.param_mc) + base;
  __nv_bfloat16* plocal = reinterpret_cast<__nv_bfloat16*>(a.param[rank]) + base;

  for (size_t i0 = tid; i0 < nvec; i0 += stride * U) {
    uint4 raw[U];
    float ps[U][8], ms[U][8] = {}, vs[U][8] = {};
    bool valid[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      valid[u] = (i0 + u * stride) < nvec;
      if (valid[u]) raw[u] = mc_ld_reduce_bf16x8(gmc + (i0 + u * stride) * 8);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!valid[u]) continue;
      const size_t le = (i0 + u * stride) * 8;
      ld_stream_f8(a.master + le, ps[u]);
      if constexpr (KIND != PSB_OPT_SGD) ld_stream_f8(a.m + le, ms[u]);
      if constexpr (KIND == PSB_OPT_ADAM || KIND == PSB_OPT_ADAMW) ld_stream_f8(a.v + le, vs[u]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!valid[u]) continue;
      const size_t le = (i0 + u * stride) * 8;
      float g[8];
      unpack8(raw[u], g);
#pragma unroll
      for (int k = 0; k < 8; ++k) update1<KIND>(c, ps[u][k], ms[u][k], vs[u][k], g[k] * gscale);
      st_stream_f8(a.master + le, ps[u]);
      if constexpr (KIND != PSB_OPT_SGD) st_stream_f8(a.m + le, ms[u]);
      if constexpr (KIND == PSB_OPT_ADAM || KIND == PSB_OPT_ADAMW) st_stream_f8(a.v + le, vs[u]);
      const uint4 packed = pack8(ps[u]);
      if (flags & PSB_PUSH_BROADCAST) mc_st_16(pmc + le, packed);
      else st_stream_16(plocal + le, packed);
    }
  }

  __shared__ int is_last;
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
                     e);
    if (threadIdx.x == 0) {
      *a.epoch = e;
      *a.done_counter = 0u;
    }
  }
}

// K2 over NVLS: every owner stores its slice of the bucket into all replicas with multimem.st
// (push-mode all-gather; the consumers only wait for the version flags).
__global__ void __launch_bounds__(kThreads, 1)
    bcast_slice_nvls_kernel(PsbFabric f, const void* param_local, void* param_mc,
                            uint64_t slice_bytes, int bucket, const uint32_t* epoch,
                            uint32_t* done_counter, int publish) {
  const size_t nvec = slice_bytes / 16;
  const size_t base = (size_t)f.rank * slice_bytes;
  const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(param_local) + base);
  uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<char*>(param_mc) + base);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride * 4) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < nvec) v[u] = ld_stream_16(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < nvec) mc_st_16(dst + i + u * stride, v[u]);
  }
  if (!publish) return;
  __shared__ int is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    fence_sys();
    is_last = (atomicAdd(done_counter, 1u) == gridDim.x - 1);
    if (is_last) fence_sys();
  }
  __syncthreads();
  if (is_last) {
    if (threadIdx.x < f.world)
      st_release_sys(f.pad[threadIdx.x] + PSB_PAD_VERSION_OFF + bucket * PSB_MAX_PEERS + f.rank,
                     *epoch);
    if (threadIdx.x == 0) *done_counter = 0u;
  }
}

}  // namespace nvls

int push_nvls_dispatch(const PsbPush& a, cudaStream_t s) {
  using namespace nvls;
  PSB_REQUIRE(a.grad_mc && (a.param_mc || !(a.flags & PSB_PUSH_BROADCAST)), PSB_EINVAL,
              "psb_push_fused(NVLS): multicast addresses of the bucket are missing");
  PSB_REQUIRE(a.grad_dtype == PSB_BF16, PSB_EUNSUPPORTED,
              "psb_push_fused(NVLS): gradients must be bf16 (multimem.ld_reduce .bf16x2)");
  PSB_REQUIRE(!(a.flags & (PSB_PUSH_MASKED | PSB_PUSH_REDUCE_ONLY | PSB_PUSH_FROM_REDUCED |
                           PSB_PUSH_SEQLOCK)),
              PSB_EUNSUPPORTED, "psb_push_fused(NVLS): masked / clip / seqlock modes use the peer-load kernel");
  PSB_REQUIRE(a.grad_off == nullptr, PSB_EUNSUPPORTED, "psb_push_fused(NVLS): no gradient ring");
  PSB_REQUIRE((((uintptr_t)a.master | (uintptr_t)a.m | (uintptr_t)a.v) & 31) == 0, PSB_EINVAL,
              "psb_push_fused(NVLS): master / m / v must be 32-byte aligned");
  PSB_REQUIRE((((uintptr_t)a.grad_mc | (uintptr_t)a.param_mc) & 15) == 0 &&
                  (a.slice_elems * 2) % 16 == 0,
              PSB_EINVAL, "psb_push_fused(NVLS): 16-byte alignment of the multicast ranges");
  const size_t nvec = a.slice_elems / 8;
  int grid = a.grid;
  if (grid <= 0) {
    size_t want = (nvec + kThreads * 2 - 1) / (kThreads * 2);
    grid = (int)(want < 1 ? 1 : (want > (size_t)kNumSMs ? kNumSMs : want));
  }
  switch (a.opt.kind) {
    case PSB_OPT_SGD: push_nvls_kernel<PSB_OPT_SGD><<<grid, kThreads, 0, s>>>(a); break;
    case PSB_OPT_MOMENTUM: push_nvls_kernel<PSB_OPT_MOMENTUM><<<grid, kThreads, 0, s>>>(a); break;
    case PSB_OPT_ADAM: push_nvls_kernel<PSB_OPT_ADAM><<<grid, kThreads, 0, s>>>(a); break;
    case PSB_OPT_ADAMW: push_nvls_kernel<PSB_OPT_ADAMW><<<grid, kThreads, 0, s>>>(a); break;
    default: PSB_REQUIRE(false, PSB_EINVAL, "psb_push_fused(NVLS): unknown optimizer %d", a.opt.kind);
  }
  PSB_LAUNCH_OK("push_nvls");
  return 0;
}

}  // namespace psb

extern "C" int psb_bcast_slice_nvls(const PsbFabric* fabric, const void* param_local, void* param_mc,
                                    uint64_t slice_bytes, int bucket, const uint32_t* epoch,
                                    uint32_t* done_counter, int grid, psb_stream_t stream) {
  using namespace psb;
  PSB_REQUIRE(fabric && param_local && param_mc, PSB_EINVAL, "psb_bcast_slice_nvls: null argument");
  PSB_REQUIRE(slice_bytes % 16 == 0, PSB_EINVAL, "psb_bcast_slice_nvls: slice_bytes %% 16 != 0");
  PSB_REQUIRE(!epoch || done_counter, PSB_EINVAL, "psb_bcast_slice_nvls: done_counter missing");
  if (grid <= 0) {
    size_t want = (slice_bytes / 16 + 2047) / 2048;
    grid = (int)(want < 1 ? 1 : (want > 64 ? 64 : want));
  }
  nvls::bcast_slice_nvls_kernel<<<grid, nvls::kThreads, 0, (cudaStream_t)stream>>>(
      *fabric, param_local, param_mc, slice_bytes, bucket, epoch, done_counter, epoch != nullptr);
  PSB_LAUNCH_OK("bcast_slice_nvls");
  return 0;
}

PSB_EXPORT_SPIN_LIMIT(push_nvls)
