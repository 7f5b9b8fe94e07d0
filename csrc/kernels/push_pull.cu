// push_pull.cu — K1 (fused gradient push), K2 (parameter pull), K9 (device waits / barriers).
//
// The data plane of the parameter server on one NVSwitch box: every GPU hosts one worker
// and owns one shard (slice) of every bucket.  Gradients live in peer-mapped symmetric
// buffers; the owner of a slice pulls the N worker contributions straight over NVLink with
// 16-byte peer loads, reduces them in fp32 in a fixed order, applies the server-side
// optimizer to its fp32 master shard and writes the refreshed bf16 slice back into every
// replica with peer stores — one kernel, no NCCL, no host staging.
//
// Reference paths replaced (below /root/reference):
//   src/nccl_manager.cpp:102-121   ncclAllReduce per tensor on the null stream      (N1)
//   src/worker.cpp:416-444         cudaMalloc + H2D + all-reduce + D2H + host divide (C9)
//   src/parameter_server.cpp:49-60 cross-worker sum and mean                        (H1, H2)
//   src/parameter_server.cpp:83-90 p -= g                                           (H3)
//   src/parameter_server.cpp:93-97 + src/worker.cpp:240-252  full-model pull        (H7)
//   src/worker.cpp:359-375         500 ms sleeps + 50 ms CheckSyncStatus polling    (K9)
#include "common.cuh"

namespace psb {

constexpr int kPushThreads = 512;

template <int GDT>
struct GradVec;  // 8 gradient elements of one peer, raw
template <>
struct GradVec<PSB_BF16> {
  uint4 a;
  __device__ __forceinline__ void load(const void* base, size_t elem) {
    a = ld_stream_16(reinterpret_cast<const __nv_bfloat16*>(base) + elem);
  }
  __device__ __forceinline__ void add_to(float g[8]) const {
    float f[8];
    unpack8(a, f);
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] += f[k];
  }
};
template <>
struct GradVec<PSB_F32> {
  uint4 a, b;
  __device__ __forceinline__ void load(const void* base, size_t elem) {
    a = ld_stream_16(reinterpret_cast<const float*>(base) + elem);
    b = ld_stream_16(reinterpret_cast<const float*>(base) + elem + 4);
  }
  __device__ __forceinline__ void add_to(float g[8]) const {
    g[0] += __uint_as_float(a.x), g[1] += __uint_as_float(a.y);
    g[2] += __uint_as_float(a.z), g[3] += __uint_as_float(a.w);
    g[4] += __uint_as_float(b.x), g[5] += __uint_as_float(b.y);
    g[6] += __uint_as_float(b.z), g[7] += __uint_as_float(b.w);
  }
};

__device__ __forceinline__ void load8(const float* p, float f[8]) { ld_stream_f8(p, f); }
__device__ __forceinline__ void store8(float* p, const float f[8]) { st_stream_f8(p, f); }

// ------------------------------------------------------------------------------------
// K1.  WORLD > 0: compile-time peer count (fully unrolled peer loads); WORLD == 0: dynamic.
// U vectors of 8 elements are in flight per thread (2 for bf16 gradients from <= 4 peers, else 1:
// 8 peers x 16 B already keeps 128 B per thread in flight and more would spill).
// ------------------------------------------------------------------------------------
template <int KIND, int GDT, int WORLD>
__global__ void __launch_bounds__(kPushThreads, 1) push_fused_kernel(const PsbPush a) {
  constexpr int MAXW = WORLD > 0 ? WORLD : PSB_MAX_PEERS;
  constexpr int U = GDT != PSB_BF16 ? 1
                    : (KIND == PSB_OPT_SGD && MAXW <= 2) ? 4
                    : MAXW <= 4 ? 2 : 1;
  const int rank = a.fabric.rank;
  const int world = WORLD > 0 ? WORLD : a.fabric.world;
  const int flags = a.flags;
  uint32_t* const my_pad = a.fabric.pad[rank];
  const uint32_t e = *reinterpret_cast<volatile const uint32_t*>(a.epoch) + 1u;
  const uint32_t mask = (flags & PSB_PUSH_MASKED) ? *a.peer_mask : 0xffffffffu;
  if ((flags & PSB_PUSH_MASKED) && (mask & ((1u << world) - 1u)) == 0u) return;  // nothing pending

  // ---- 1. announce "my gradients of this bucket are complete", 2. wait for everyone's
  if (!(flags & (PSB_PUSH_NO_WAIT | PSB_PUSH_FROM_REDUCED))) {
    if (blockIdx.x == 0 && threadIdx.x < world) {
      fence_sys();
      st_release_sys(a.fabric.pad[threadIdx.x] + PSB_PAD_READY_OFF + a.bucket * PSB_MAX_PEERS + rank,
                     e);
    }
    if (threadIdx.x < world && ((mask >> threadIdx.x) & 1u)) {
      wait_ge_sys(my_pad + PSB_PAD_READY_OFF + a.bucket * PSB_MAX_PEERS + threadIdx.x, e);
    }
    __syncthreads();
  }

  const OptimCoef c = make_coef(a.opt, a.step ? *a.step : 0u);
  const float gscale = a.opt.grad_scale;
  const float clip = a.clip_coef ? *a.clip_coef : 1.f;
  const size_t nvec = a.slice_elems / 8;
  const size_t base = (size_t)rank * a.slice_elems;  // my slice inside the bucket
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  float sumsq = 0.f;

  for (size_t i0 = tid; i0 < nvec; i0 += stride * U) {
    float g[U][8];
    bool valid[U];
#pragma unroll
    for (int u = 0; u < U; ++u) valid[u] = (i0 + u * stride) < nvec;

    if (flags & PSB_PUSH_FROM_REDUCED) {
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (valid[u]) load8(a.reduced + (i0 + u * stride) * 8, g[u]);
    } else {
      // issue every peer load of every in-flight vector before the first use
      GradVec<GDT> raw[U][MAXW];
#pragma unroll
      for (int u = 0; u < U; ++u) {
#pragma unroll
        for (int w = 0; w < MAXW; ++w) {
          if (w < world && valid[u] && ((mask >> w) & 1u))
            raw[u][w].load(reinterpret_cast<const char*>(a.grad[w]) +
                               (a.grad_off ? __ldg(a.grad_off + w) : 0ull),
                           base + (i0 + u * stride) * 8);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
#pragma unroll
        for (int k = 0; k < 8; ++k) g[u][k] = 0.f;
#pragma unroll
        for (int w = 0; w < MAXW; ++w) {  // fixed order => deterministic sum
          if (w < world && valid[u] && ((mask >> w) & 1u)) raw[u][w].add_to(g[u]);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) g[u][k] *= gscale;
      }
    }

    if (flags & PSB_PUSH_REDUCE_ONLY) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!valid[u]) continue;
        store8(a.reduced + (i0 + u * stride) * 8, g[u]);
#pragma unroll
        for (int k = 0; k < 8; ++k) sumsq = fmaf(g[u][k], g[u][k], sumsq);
      }
      continue;
    }

    // state of every in-flight vector is requested before the first update
    float ps[U][8], ms[U][8] = {}, vs[U][8] = {};
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!valid[u]) continue;
      const size_t le = (i0 + u * stride) * 8;
      load8(a.master + le, ps[u]);
      if constexpr (KIND != PSB_OPT_SGD) load8(a.m + le, ms[u]);
      if constexpr (KIND == PSB_OPT_ADAM || KIND == PSB_OPT_ADAMW) load8(a.v + le, vs[u]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!valid[u]) continue;
      const size_t le = (i0 + u * stride) * 8;  // element inside my slice
      float(&p)[8] = ps[u];
      float(&mm)[8] = ms[u];
      float(&vv)[8] = vs[u];
#pragma unroll
      for (int k = 0; k < 8; ++k) update1<KIND>(c, p[k], mm[k], vv[k], g[u][k] * clip);
      store8(a.master + le, p);
      if constexpr (KIND != PSB_OPT_SGD) store8(a.m + le, mm);
      if constexpr (KIND == PSB_OPT_ADAM || KIND == PSB_OPT_ADAMW) store8(a.v + le, vv);
      const uint4 packed = pack8(p);
      if (flags & PSB_PUSH_BROADCAST) {
#pragma unroll
        for (int w = 0; w < MAXW; ++w) {
          if (w < world)
            st_stream_16(reinterpret_cast<__nv_bfloat16*>(a.param[w]) + base + le, packed);
        }
      } else {
        st_stream_16(reinterpret_cast<__nv_bfloat16*>(a.param[rank]) + base + le, packed);
      }
    }
  }

  // ---- Σg² of the reduced gradient (clip path)
  __shared__ float warp_part[kPushThreads / 32];
  __shared__ int is_last;
  if (flags & PSB_PUSH_REDUCE_ONLY) {
    sumsq = warp_sum(sumsq);
    if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = sumsq;
    __syncthreads();
    if (threadIdx.x < 32) {
      float w = threadIdx.x < (blockDim.x >> 5) ? warp_part[threadIdx.x] : 0.f;
      w = warp_sum(w);
      if (threadIdx.x == 0 && a.sumsq_acc) atomicAdd(a.sumsq_acc, w);
    }
  }

  // ---- 4. last CTA publishes the new version to every replica holder
  __syncthreads();
  if (threadIdx.x == 0) {
    fence_sys();
    is_last = (atomicAdd(a.done_counter, 1u) == gridDim.x - 1);
    if (is_last) fence_sys();
  }
  __syncthreads();
  if (is_last) {
    if (!(flags & PSB_PUSH_REDUCE_ONLY)) {
      if (threadIdx.x < world) {
        st_release_sys(
            a.fabric.pad[threadIdx.x] + PSB_PAD_VERSION_OFF + a.bucket * PSB_MAX_PEERS + rank,
            (flags & PSB_PUSH_SEQLOCK) ? 2u * e : e);
      }
      if (threadIdx.x == 0) *a.epoch = e;
    }
    if (threadIdx.x == 0) *a.done_counter = 0u;
  }
}

// ------------------------------------------------------------------------------------
// Steps 1-2 of K1 as a kernel of their own: one warp announces "my gradients of this bucket
// are complete" on every owner's pad and waits for everybody's announcement.  Launched in front
// of a K1 with PSB_PUSH_NO_WAIT, the data kernel starts only when its inputs exist — its CTAs no
// longer sit on SMs (next to a GEMM that wants them) while the slowest rank finishes its dW.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) push_ready_kernel(PsbFabric f, int bucket,
                                                        const uint32_t* epoch) {
  const uint32_t e = *reinterpret_cast<volatile const uint32_t*>(epoch) + 1u;
  if (threadIdx.x < f.world) {
    fence_sys();
    st_release_sys(f.pad[threadIdx.x] + PSB_PAD_READY_OFF + bucket * PSB_MAX_PEERS + f.rank, e);
    wait_ge_sys(f.pad[f.rank] + PSB_PAD_READY_OFF + bucket * PSB_MAX_PEERS + threadIdx.x, e);
  }
}

// ------------------------------------------------------------------------------------
// K2.  Pull-mode all-gather: each warp copies 512-byte runs from one owner at a time.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512, 1) pull_allgather_kernel(const PsbPull a) {
  const int rank = a.fabric.rank, world = a.fabric.world;
  const uint32_t* vpad = a.fabric.pad[rank] + PSB_PAD_VERSION_OFF + a.bucket * PSB_MAX_PEERS;
  if (!a.skip_wait && !a.seqlock) {
    const uint32_t e = *a.epoch;
    if (threadIdx.x < world) wait_ge_sys(vpad + threadIdx.x, e);
    __syncthreads();
  }
  if (world == 1) return;
  const size_t slice_vec = a.slice_bytes / 16;
  const int lane = threadIdx.x & 31;
  const size_t warps = (size_t)gridDim.x * (blockDim.x >> 5);
  const size_t warp_id = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  constexpr int U = 4;
  if (a.seqlock) {
    // Async mode: owners rewrite their slices at any time.  Each CTA copies its share of one
    // owner's slice between two reads of that owner's version word and repeats the share if
    // the word was odd (update in flight) or changed meanwhile.
    __shared__ uint32_t v0s, v1s;
    const size_t chunks = (slice_vec + 31) / 32;
    for (int ord = 0; ord < world - 1; ++ord) {
      const int owner = (rank + 1 + ord) % world;
      while (true) {
        if (threadIdx.x == 0) {
          uint32_t v;
          SpinWatch watch;
          while ((v = ld_acquire_sys(vpad + owner)) & 1u) {
            __nanosleep(64);
            if (watch.tick()) break;
          }
          v0s = v;
        }
        __syncthreads();
        for (size_t c = warp_id; c < chunks; c += warps) {
          const size_t vec = c * 32 + lane;
          if (vec < slice_vec) {
            const size_t off = (size_t)owner * a.slice_bytes + vec * 16;
            st_stream_16(reinterpret_cast<char*>(a.param[rank]) + off,
                         ld_stream_16(reinterpret_cast<const char*>(a.param[owner]) + off));
          }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
          fence_sys();
          v1s = ld_acquire_sys(vpad + owner);
        }
        __syncthreads();
        const bool stable = v1s == v0s || abort_requested();
        __syncthreads();  // nobody may still be comparing when thread 0 rewrites v0s
        if (stable) break;
      }
    }
    return;
  }
  const size_t chunks_per_slice = (slice_vec + 31) / 32;  // one warp-wide run = 32 x 16 B
  const size_t total_chunks = chunks_per_slice * (size_t)(world - 1);
  for (size_t c0 = warp_id; c0 < total_chunks; c0 += warps * U) {
    uint4 val[U];
    size_t off[U];
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t c = c0 + (size_t)u * warps;
      ok[u] = false;
      if (c < total_chunks) {
        const int ord = (int)(c % (size_t)(world - 1));
        const int owner = (rank + 1 + ord) % world;
        const size_t vec = (c / (size_t)(world - 1)) * 32 + lane;
        if (vec < slice_vec) {
          off[u] = (size_t)owner * a.slice_bytes + vec * 16;
          val[u] = ld_stream_16(reinterpret_cast<const char*>(a.param[owner]) + off[u]);
          ok[u] = true;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (ok[u]) st_stream_16(reinterpret_cast<char*>(a.param[rank]) + off[u], val[u]);
  }
}

// ------------------------------------------------------------------------------------
// Async / SSP control kernels (tiny: one thread per (bucket, peer) or per bucket).
// ------------------------------------------------------------------------------------
__global__ void async_signal_kernel(PsbAsync a, uint32_t clock) {
  const int rank = a.fabric.rank, world = a.fabric.world;
  const int total = a.num_buckets * world;
  __threadfence_system();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int b = i / world, r = i % world;
    st_release_sys(a.fabric.pad[r] + PSB_PAD_READY_OFF + b * PSB_MAX_PEERS + rank, clock);
  }
  if (blockIdx.x == 0 && threadIdx.x < world)
    st_release_sys(a.fabric.pad[threadIdx.x] + PSB_PAD_CLOCK_OFF + rank, clock);
}

__global__ void pad_post_kernel(PsbFabric f, uint32_t word_base, uint32_t value) {
  if (threadIdx.x < f.world) {
    fence_sys();
    st_release_sys(f.pad[threadIdx.x] + word_base + f.rank, value);
  }
}

__global__ void async_plan_kernel(PsbAsync a) {
  const int rank = a.fabric.rank, world = a.fabric.world;
  const uint32_t* ready = a.fabric.pad[rank] + PSB_PAD_READY_OFF;
  for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < a.num_buckets;
       b += gridDim.x * blockDim.x) {
    uint32_t mask = 0;
    for (int w = 0; w < world; ++w) {
      const uint32_t rdy = ld_acquire_sys(ready + b * PSB_MAX_PEERS + w);
      const uint32_t ap = a.applied[b * PSB_MAX_PEERS + w];
      if ((int32_t)(rdy - ap) > 0) {  // oldest unapplied clock of this worker
        const uint32_t clock = ap + 1u;
        mask |= 1u << w;
        a.grad_off[b * PSB_MAX_PEERS + w] = (uint64_t)(clock % (uint32_t)a.ring) * a.ring_stride_bytes;
        a.applied[b * PSB_MAX_PEERS + w] = clock;
      } else {
        a.grad_off[b * PSB_MAX_PEERS + w] = 0;
      }
    }
    a.masks[b] = mask;
    if (mask) {  // seqlock: odd = "slice r of bucket b is being rewritten"
      const uint32_t odd = 2u * a.epoch[b] + 1u;
      for (int t = 0; t < world; ++t)
        st_release_sys(a.fabric.pad[t] + PSB_PAD_VERSION_OFF + b * PSB_MAX_PEERS + rank, odd);
    }
  }
}

__global__ void async_ack_kernel(PsbAsync a) {
  const int rank = a.fabric.rank, world = a.fabric.world;
  const int total = a.num_buckets * world;
  __threadfence_system();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int b = i / world, w = i % world;
    st_release_sys(a.fabric.pad[w] + PSB_PAD_ACK_OFF + b * PSB_MAX_PEERS + rank,
                   a.applied[b * PSB_MAX_PEERS + w]);
  }
}

__global__ void async_gate_kernel(PsbAsync a, uint32_t clock, int staleness) {
  const int rank = a.fabric.rank, world = a.fabric.world;
  const uint32_t* ack = a.fabric.pad[rank] + PSB_PAD_ACK_OFF;
  const uint32_t* clk = a.fabric.pad[rank] + PSB_PAD_CLOCK_OFF;
  const int total = a.num_buckets * world;
  if (clock > (uint32_t)a.ring) {  // the slot of `clock` last held clock - ring
    const uint32_t need = clock - (uint32_t)a.ring;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x)
      wait_ge_sys(ack + (i / world) * PSB_MAX_PEERS + (i % world), need);
  }
  if (staleness >= 0 && blockIdx.x == 0 && threadIdx.x < world &&
      clock > (uint32_t)staleness + 1u) {
    wait_ge_sys(clk + threadIdx.x, clock - 1u - (uint32_t)staleness);
  }
}

// ------------------------------------------------------------------------------------
// K9.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    wait_versions_kernel(PsbFabric f, const uint32_t* __restrict__ epoch_words, int lo, int hi) {
  const uint32_t* pad = f.pad[f.rank] + PSB_PAD_VERSION_OFF;
  const int total = (hi - lo) * f.world;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int b = lo + i / f.world, r = i % f.world;
    wait_ge_sys(pad + b * PSB_MAX_PEERS + r, epoch_words[b]);
  }
}

__global__ void __launch_bounds__(32) device_barrier_kernel(PsbFabric f, int slot, uint32_t* seq) {
  const uint32_t s = *seq + 1u;
  if (threadIdx.x < f.world) {
    fence_sys();
    st_release_sys(f.pad[threadIdx.x] + PSB_PAD_BARRIER_OFF + slot * PSB_MAX_PEERS + f.rank, s);
    wait_ge_sys(f.pad[f.rank] + PSB_PAD_BARRIER_OFF + slot * PSB_MAX_PEERS + threadIdx.x, s);
  }
  __syncwarp();
  if (threadIdx.x == 0) *seq = s;
}

__device__ __forceinline__ void st_release_sys_u64(uint64_t* p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// Cross-shard all-reduce of one float per rank through the signal pads: each rank stores
// {seq, value} as ONE 64-bit word into slot [seq & 1][rank] of every pad, then reads the N
// words of its own pad and sums them in rank order (identical result on every rank).
__global__ void __launch_bounds__(32)
    norm_allreduce_kernel(PsbFabric f, float* local_sumsq, uint32_t* seq, float max_norm,
                          float eps, float* coef_out, float* norm_out) {
  const uint32_t s = *seq + 1u;
  const int par = s & 1u;
  const float mine = *local_sumsq;
  float got = 0.f;
  if (threadIdx.x < f.world) {
    uint64_t word = ((uint64_t)s << 32) | (uint64_t)__float_as_uint(mine);
    uint64_t* dst = reinterpret_cast<uint64_t*>(f.pad[threadIdx.x] + PSB_PAD_NORM_OFF) +
                    par * PSB_MAX_PEERS + f.rank;
    st_release_sys_u64(dst, word);
    const uint64_t* src = reinterpret_cast<const uint64_t*>(f.pad[f.rank] + PSB_PAD_NORM_OFF) +
                          par * PSB_MAX_PEERS + threadIdx.x;
    uint64_t w;
    SpinWatch watch;
    do {
      w = ld_acquire_sys_u64(src);
      if ((int32_t)((uint32_t)(w >> 32) - s) >= 0) break;
      __nanosleep(32);
      if (watch.tick()) break;
    } while (true);
    got = __uint_as_float((uint32_t)w);
  }
  float total = 0.f;
  for (int r = 0; r < f.world; ++r) total += __shfl_sync(0xffffffffu, got, r);
  if (threadIdx.x == 0) {
    const float norm = sqrtf(total);
    if (norm_out) *norm_out = norm;
    float c = max_norm / (norm + eps);
    *coef_out = (max_norm > 0.f && c < 1.f) ? c : 1.f;
    *local_sumsq = 0.f;
    *seq = s;
  }
}

static int check_fabric(const PsbFabric& f, const char* who) {
  PSB_REQUIRE(f.world >= 1 && f.world <= PSB_MAX_PEERS, PSB_ERANGE,
              "%s: world %d outside [1, %d]", who, f.world, PSB_MAX_PEERS);
  PSB_REQUIRE(f.rank >= 0 && f.rank < f.world, PSB_EINVAL, "%s: rank %d outside world %d", who,
              f.rank, f.world);
  for (int i = 0; i < f.world; ++i)
    PSB_REQUIRE(f.pad[i] != nullptr, PSB_EINVAL, "%s: signal pad of peer %d is null", who, i);
  return 0;
}

template <int KIND, int GDT>
static void launch_push(const PsbPush& a, int grid, cudaStream_t s) {
  const int kPushThreads = a.block > 0 ? a.block : psb::kPushThreads;
  switch (a.fabric.world) {
    case 1: push_fused_kernel<KIND, GDT, 1><<<grid, kPushThreads, 0, s>>>(a); break;
    case 2: push_fused_kernel<KIND, GDT, 2><<<grid, kPushThreads, 0, s>>>(a); break;
    case 4: push_fused_kernel<KIND, GDT, 4><<<grid, kPushThreads, 0, s>>>(a); break;
    case 8: push_fused_kernel<KIND, GDT, 8><<<grid, kPushThreads, 0, s>>>(a); break;
    default: push_fused_kernel<KIND, GDT, 0><<<grid, kPushThreads, 0, s>>>(a); break;
  }
}

}  // namespace psb

using namespace psb;

extern "C" {

size_t psb_pad_bytes(void) { return (size_t)PSB_PAD_WORDS * sizeof(uint32_t); }

int psb_push_fused(const PsbPush* args, psb_stream_t stream) {
  PSB_REQUIRE(args, PSB_EINVAL, "psb_push_fused: null args");
  const PsbPush& a = *args;
  if (int rc = check_fabric(a.fabric, "psb_push_fused")) return rc;
  PSB_REQUIRE(a.bucket >= 0 && a.bucket < PSB_MAX_BUCKETS, PSB_ERANGE,
              "psb_push_fused: bucket %d outside [0, %d)", a.bucket, PSB_MAX_BUCKETS);
  PSB_REQUIRE(a.slice_elems % 8 == 0, PSB_EINVAL,
              "psb_push_fused: slice_elems %llu must be a multiple of 8",
              (unsigned long long)a.slice_elems);
  PSB_REQUIRE(a.grad_dtype == PSB_F32 || a.grad_dtype == PSB_BF16, PSB_EINVAL,
              "psb_push_fused: bad grad dtype %d", a.grad_dtype);
  PSB_REQUIRE(a.epoch && a.done_counter, PSB_EINVAL, "psb_push_fused: epoch/done_counter missing");
  const bool reduce_only = a.flags & PSB_PUSH_REDUCE_ONLY;
  const bool from_reduced = a.flags & PSB_PUSH_FROM_REDUCED;
  PSB_REQUIRE(!(reduce_only && from_reduced), PSB_EINVAL,
              "psb_push_fused: REDUCE_ONLY and FROM_REDUCED are exclusive");
  PSB_REQUIRE(!(reduce_only || from_reduced) || a.reduced, PSB_EINVAL,
              "psb_push_fused: `reduced` scratch missing");
  PSB_REQUIRE(!(a.flags & PSB_PUSH_MASKED) || a.peer_mask, PSB_EINVAL,
              "psb_push_fused: peer_mask missing");
  if (!reduce_only) {
    PSB_REQUIRE(a.master, PSB_EINVAL, "psb_push_fused: master shard missing");
    PSB_REQUIRE(a.opt.kind == PSB_OPT_SGD || a.m, PSB_EINVAL, "psb_push_fused: m missing");
    PSB_REQUIRE((a.opt.kind != PSB_OPT_ADAM && a.opt.kind != PSB_OPT_ADAMW) || a.v, PSB_EINVAL,
                "psb_push_fused: v missing");
    for (int i = 0; i < a.fabric.world; ++i)
      if ((a.flags & PSB_PUSH_BROADCAST) || i == a.fabric.rank)
        PSB_REQUIRE(a.param[i], PSB_EINVAL, "psb_push_fused: replica of peer %d is null", i);
  }
  if (!from_reduced)
    for (int i = 0; i < a.fabric.world; ++i)
      PSB_REQUIRE(a.grad[i], PSB_EINVAL, "psb_push_fused: gradient buffer of peer %d is null", i);

  if (a.flags & PSB_PUSH_NVLS) return push_nvls_dispatch(a, (cudaStream_t)stream);
  if ((a.flags & PSB_PUSH_TMA) && !reduce_only && !from_reduced)
    return push_fused_tma_dispatch(a, (cudaStream_t)stream);
  PSB_REQUIRE((((uintptr_t)a.master | (uintptr_t)a.m | (uintptr_t)a.v | (uintptr_t)a.reduced) & 31) == 0,
              PSB_EINVAL, "psb_push_fused: master / m / v / reduced must be 32-byte aligned");
  PSB_REQUIRE(a.block == 0 || (a.block % 32 == 0 && a.block >= 64 && a.block <= kPushThreads),
              PSB_EINVAL, "psb_push_fused: block=%d must be a multiple of 32 in [64, %d]", a.block,
              kPushThreads);
  const size_t nvec = a.slice_elems / 8;
  int grid = a.grid;
  if (grid <= 0) {
    // enough CTAs to keep ~2 MB in flight over NVLink, never more than one wave
    const size_t per_cta = (size_t)(a.block > 0 ? a.block : kPushThreads) * 2;
    size_t want = (nvec + per_cta - 1) / per_cta;
    grid = (int)(want < 1 ? 1 : (want > (size_t)kNumSMs ? kNumSMs : want));
  }
  cudaStream_t s = (cudaStream_t)stream;
#define PSB_DISPATCH(K)                                              \
  case K:                                                            \
    if (a.grad_dtype == PSB_BF16)                                    \
      launch_push<K, PSB_BF16>(a, grid, s);                          \
    else                                                             \
      launch_push<K, PSB_F32>(a, grid, s);                           \
    break;
  switch (a.opt.kind) {
    PSB_DISPATCH(PSB_OPT_SGD)
    PSB_DISPATCH(PSB_OPT_MOMENTUM)
    PSB_DISPATCH(PSB_OPT_ADAM)
    PSB_DISPATCH(PSB_OPT_ADAMW)
    default:
      PSB_REQUIRE(false, PSB_EINVAL, "psb_push_fused: unknown optimizer kind %d", a.opt.kind);
  }
#undef PSB_DISPATCH
  PSB_LAUNCH_OK("push_fused");
  return 0;
}

int psb_push_ready(const PsbPush* args, psb_stream_t stream) {
  PSB_REQUIRE(args && args->epoch, PSB_EINVAL, "psb_push_ready: null argument");
  if (int rc = check_fabric(args->fabric, "psb_push_ready")) return rc;
  PSB_REQUIRE(args->bucket >= 0 && args->bucket < PSB_MAX_BUCKETS, PSB_ERANGE,
              "psb_push_ready: bucket %d outside [0, %d)", args->bucket, PSB_MAX_BUCKETS);
  push_ready_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(args->fabric, args->bucket, args->epoch);
  PSB_LAUNCH_OK("push_ready");
  return 0;
}

int psb_pull_allgather(const PsbPull* args, psb_stream_t stream) {
  PSB_REQUIRE(args, PSB_EINVAL, "psb_pull_allgather: null args");
  const PsbPull& a = *args;
  if (int rc = check_fabric(a.fabric, "psb_pull_allgather")) return rc;
  PSB_REQUIRE(a.bucket >= 0 && a.bucket < PSB_MAX_BUCKETS, PSB_ERANGE,
              "psb_pull_allgather: bucket %d outside [0, %d)", a.bucket, PSB_MAX_BUCKETS);
  PSB_REQUIRE(a.slice_bytes % 16 == 0, PSB_EINVAL, "psb_pull_allgather: slice_bytes %% 16 != 0");
  PSB_REQUIRE(a.skip_wait || a.epoch, PSB_EINVAL, "psb_pull_allgather: epoch missing");
  for (int i = 0; i < a.fabric.world; ++i)
    PSB_REQUIRE(a.param[i], PSB_EINVAL, "psb_pull_allgather: replica of peer %d is null", i);
  int grid = a.grid;
  if (grid <= 0) {
    size_t chunks = (a.slice_bytes / 16 + 31) / 32 * (size_t)(a.fabric.world - 1);
    size_t want = (chunks + 16 * 4 - 1) / (16 * 4);
    grid = (int)(want < 1 ? 1 : (want > 64 ? 64 : want));
  }
  pull_allgather_kernel<<<grid, 512, 0, (cudaStream_t)stream>>>(a);
  PSB_LAUNCH_OK("pull_allgather");
  return 0;
}

int psb_wait_versions(const PsbFabric* fabric, const uint32_t* epoch_words, int bucket_lo,
                      int bucket_hi, psb_stream_t stream) {
  PSB_REQUIRE(fabric && epoch_words, PSB_EINVAL, "psb_wait_versions: null argument");
  if (int rc = check_fabric(*fabric, "psb_wait_versions")) return rc;
  PSB_REQUIRE(0 <= bucket_lo && bucket_lo <= bucket_hi && bucket_hi <= PSB_MAX_BUCKETS, PSB_ERANGE,
              "psb_wait_versions: bucket range [%d, %d) invalid", bucket_lo, bucket_hi);
  if (bucket_lo == bucket_hi) return 0;
  wait_versions_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(*fabric, epoch_words, bucket_lo,
                                                             bucket_hi);
  PSB_LAUNCH_OK("wait_versions");
  return 0;
}

int psb_device_barrier(const PsbFabric* fabric, int slot, uint32_t* seq, psb_stream_t stream) {
  PSB_REQUIRE(fabric && seq, PSB_EINVAL, "psb_device_barrier: null argument");
  if (int rc = check_fabric(*fabric, "psb_device_barrier")) return rc;
  PSB_REQUIRE(slot >= 0 && slot < 64, PSB_ERANGE, "psb_device_barrier: slot %d outside [0, 64)",
              slot);
  device_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*fabric, slot, seq);
  PSB_LAUNCH_OK("device_barrier");
  return 0;
}

int psb_norm_allreduce(const PsbFabric* fabric, float* local_sumsq, uint32_t* seq,
                       float max_norm, float eps, float* coef_out, float* norm_out,
                       psb_stream_t stream) {
  PSB_REQUIRE(fabric && local_sumsq && seq && coef_out, PSB_EINVAL,
              "psb_norm_allreduce: null argument");
  if (int rc = check_fabric(*fabric, "psb_norm_allreduce")) return rc;
  norm_allreduce_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(
      *fabric, local_sumsq, seq, max_norm, eps, coef_out, norm_out);
  PSB_LAUNCH_OK("norm_allreduce");
  return 0;
}

static int check_async(const PsbAsync* a, const char* who) {
  PSB_REQUIRE(a, PSB_EINVAL, "%s: null args", who);
  if (int rc = check_fabric(a->fabric, who)) return rc;
  PSB_REQUIRE(a->num_buckets > 0 && a->num_buckets <= PSB_MAX_BUCKETS, PSB_ERANGE,
              "%s: num_buckets %d outside (0, %d]", who, a->num_buckets, PSB_MAX_BUCKETS);
  PSB_REQUIRE(a->ring >= 1 && a->ring <= 8, PSB_ERANGE, "%s: ring %d outside [1, 8]", who, a->ring);
  PSB_REQUIRE(a->applied && a->masks && a->grad_off && a->epoch, PSB_EINVAL,
              "%s: state arrays missing", who);
  return 0;
}

int psb_async_signal(const PsbAsync* a, uint32_t clock, psb_stream_t stream) {
  if (int rc = check_async(a, "psb_async_signal")) return rc;
  async_signal_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(*a, clock);
  PSB_LAUNCH_OK("async_signal");
  return 0;
}

int psb_async_plan(const PsbAsync* a, psb_stream_t stream) {
  if (int rc = check_async(a, "psb_async_plan")) return rc;
  async_plan_kernel<<<1, 128, 0, (cudaStream_t)stream>>>(*a);
  PSB_LAUNCH_OK("async_plan");
  return 0;
}

int psb_async_ack(const PsbAsync* a, psb_stream_t stream) {
  if (int rc = check_async(a, "psb_async_ack")) return rc;
  async_ack_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(*a);
  PSB_LAUNCH_OK("async_ack");
  return 0;
}

int psb_async_gate(const PsbAsync* a, uint32_t clock, int32_t staleness, psb_stream_t stream) {
  if (int rc = check_async(a, "psb_async_gate")) return rc;
  async_gate_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(*a, clock, staleness);
  PSB_LAUNCH_OK("async_gate");
  return 0;
}

int psb_pad_post(const PsbFabric* fabric, uint32_t word_base, uint32_t value, psb_stream_t stream) {
  PSB_REQUIRE(fabric, PSB_EINVAL, "psb_pad_post: null fabric");
  if (int rc = check_fabric(*fabric, "psb_pad_post")) return rc;
  PSB_REQUIRE(word_base + PSB_MAX_PEERS <= PSB_PAD_WORDS, PSB_ERANGE,
              "psb_pad_post: word %u outside the pad", word_base);
  pad_post_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*fabric, word_base, value);
  PSB_LAUNCH_OK("pad_post");
  return 0;
}

}  // extern "C"

PSB_EXPORT_SPIN_LIMIT(push_pull)
extern "C" int psb_spin_limit_push_tma(unsigned long long ns);
extern "C" int psb_spin_limit_gemm1(unsigned long long ns);
extern "C" int psb_spin_limit_gemm2(unsigned long long ns);
extern "C" int psb_spin_limit_push_nvls(unsigned long long ns);

extern "C" int psb_abort_word_push_tma(const uint32_t* p);
extern "C" int psb_abort_word_gemm1(const uint32_t* p);
extern "C" int psb_abort_word_gemm2(const uint32_t* p);
extern "C" int psb_abort_word_push_nvls(const uint32_t* p);

// Host-mapped abort word (see common.cuh).  create: allocates the word (zero), points the
// kernels of the CURRENT device at it and returns its host address; the caller (or the control
// plane's heartbeat thread) stores 1 there to make every peer spin-wait on this device give up.
extern "C" int psb_abort_word_create(uint32_t** host_word) {
  PSB_REQUIRE(host_word, PSB_EINVAL, "psb_abort_word_create: null argument");
  uint32_t* h = nullptr;
  PSB_CUDA_OK(cudaHostAlloc(reinterpret_cast<void**>(&h), sizeof(uint32_t),
                            cudaHostAllocMapped | cudaHostAllocPortable));
  *h = 0u;
  uint32_t* d = nullptr;
  PSB_CUDA_OK(cudaHostGetDevicePointer(reinterpret_cast<void**>(&d), h, 0));
  int (*setters[])(const uint32_t*) = {psb_abort_word_push_pull, psb_abort_word_push_tma,
                                       psb_abort_word_gemm1, psb_abort_word_gemm2,
                                       psb_abort_word_push_nvls};
  for (auto set : setters) PSB_CUDA_OK((cudaError_t)set(d));
  *host_word = h;
  return 0;
}

// Detaches the kernels of the current device from the word and frees it.
extern "C" int psb_abort_word_destroy(uint32_t* host_word) {
  int (*setters[])(const uint32_t*) = {psb_abort_word_push_pull, psb_abort_word_push_tma,
                                       psb_abort_word_gemm1, psb_abort_word_gemm2,
                                       psb_abort_word_push_nvls};
  for (auto set : setters) PSB_CUDA_OK((cudaError_t)set(nullptr));
  PSB_CUDA_OK(cudaDeviceSynchronize());
  if (host_word) PSB_CUDA_OK(cudaFreeHost(host_word));
  return 0;
}

extern "C" int psb_set_spin_timeout_ms(uint64_t ms) {
  PSB_REQUIRE(ms >= 1, PSB_EINVAL, "psb_set_spin_timeout_ms: the limit must be at least 1 ms");
  const unsigned long long ns = (unsigned long long)ms * 1000000ull;
  // one copy of the limit per translation unit that spins on peer words (current device)
  int (*setters[])(unsigned long long) = {psb_spin_limit_push_pull, psb_spin_limit_push_tma,
                                          psb_spin_limit_gemm1, psb_spin_limit_gemm2,
                                          psb_spin_limit_push_nvls};
  for (auto set : setters) {
    PSB_CUDA_OK((cudaError_t)set(ns));
  }
  return 0;
}
