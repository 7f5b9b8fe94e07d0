// common.cuh — shared device helpers for libpsb_kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "psb_kernels.h"

namespace psb {

// ---- host-side error plumbing -------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;
inline void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

#define PSB_CUDA_OK(expr)                                                                   \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      psb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,      \
                     __LINE__);                                                             \
      return (int)_e;                                                                       \
    }                                                                                       \
  } while (0)

#define PSB_LAUNCH_OK(name)                                                                 \
  do {                                                                                      \
    cudaError_t _e = cudaGetLastError();                                                    \
    if (_e != cudaSuccess) {                                                                \
      psb::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));              \
      return (int)_e;                                                                       \
    }                                                                                       \
    psb::count_launch();                                                                    \
  } while (0)

#define PSB_REQUIRE(cond, code, ...)                                                        \
  do {                                                                                      \
    if (!(cond)) {                                                                          \
      psb::set_error(__VA_ARGS__);                                                          \
      return (code);                                                                        \
    }                                                                                       \
  } while (0)

constexpr int kNumSMs = 148;  // B200

// push_tma.cu
int push_fused_tma_dispatch(const PsbPush& a, cudaStream_t s);
// push_nvls.cu
int push_nvls_dispatch(const PsbPush& a, cudaStream_t s);

// ---- memory-model primitives (system scope: visible across NVLink peers) ------
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }

// ---- watchdog of the peer spin-waits ---------------------------------------------------
// A peer that died (or whose kernel never launched) would leave a waiting kernel spinning for
// ever and the GPU unusable.  Every spin loop on a word written by a peer ticks a SpinWatch;
// when the loop has run longer than the limit the kernel traps: the launch ends with a CUDA
// error that the host reports (the process has to restart from a checkpoint, as after any
// sticky CUDA error) and the GPU is free again.  The limit is a per-device word in every
// translation unit that waits (static __device__), set by psb_set_spin_timeout_ms().
static __device__ unsigned long long g_spin_limit_ns = 300ull * 1000000000ull;
// Abort word: a 32-bit word in host-mapped memory (psb_abort_word_create).  The control plane
// sets it when a peer was evicted (control/client.cpp: psbc_watch); spin-waits then give up and
// return instead of running into the watchdog's trap, so the CUDA context of the surviving
// ranks stays usable (they can still write a checkpoint or leave cleanly).  Data produced by a
// kernel that gave up is undefined — the caller must not use it.
static __device__ const uint32_t* g_abort_word = nullptr;

__device__ __forceinline__ bool abort_requested() {
  const uint32_t* w = g_abort_word;
  return w != nullptr && *reinterpret_cast<const volatile uint32_t*>(w) != 0u;
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

struct SpinWatch {
  unsigned long long t0 = 0;
  uint32_t n = 0;
  // true: stop waiting (the job is being torn down, see g_abort_word)
  __device__ __forceinline__ bool tick() {
    if ((++n & 1023u) != 0) return false;  // look at the clock / abort word once per 1024 polls
    if (abort_requested()) return true;
    const unsigned long long now = globaltimer_ns();
    if (t0 == 0) t0 = now;
    else if (now - t0 > g_spin_limit_ns) __trap();
    return false;
  }
};

// host side of the limit for the translation unit that includes this header
static inline cudaError_t set_spin_limit_in_this_unit(unsigned long long ns) {
  return cudaMemcpyToSymbol(g_spin_limit_ns, &ns, sizeof(ns));
}
static inline cudaError_t set_abort_word_in_this_unit(const uint32_t* dev_ptr) {
  return cudaMemcpyToSymbol(g_abort_word, &dev_ptr, sizeof(dev_ptr));
}
#define PSB_EXPORT_SPIN_LIMIT(tag)                                         \
  extern "C" int psb_spin_limit_##tag(unsigned long long ns) {             \
    return (int)psb::set_spin_limit_in_this_unit(ns);                      \
  }                                                                        \
  extern "C" int psb_abort_word_##tag(const uint32_t* dev_ptr) {           \
    return (int)psb::set_abort_word_in_this_unit(dev_ptr);                 \
  }

// spin until *p >= target (wrap-safe signed distance)
__device__ __forceinline__ void wait_ge_sys(const uint32_t* p, uint32_t target) {
  SpinWatch watch;
  while ((int32_t)(ld_acquire_sys(p) - target) < 0) {
    __nanosleep(32);
    if (watch.tick()) return;
  }
}

// ---- 16-byte streaming loads / stores ------------------------------------------
// Peer (NVLink) addresses are cached only in the local L1, which is invalidated at
// every launch; no_allocate keeps streaming data from thrashing it.
__device__ __forceinline__ uint4 ld_stream_16(const void* p) {
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream_16(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
  uint4 r = ld_stream_16(p);
  return make_float4(__uint_as_float(r.x), __uint_as_float(r.y), __uint_as_float(r.z),
                     __uint_as_float(r.w));
}
// 256-bit accesses (sm_100: LDG.E.256 / STG.E.256), p 32-byte aligned.  One instruction per
// 8 floats: a warp covers whole 32-byte sectors instead of two half-sector passes.
__device__ __forceinline__ void ld_stream_f8(const float* p, float f[8]) {
  asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(f[0]), "=f"(f[1]), "=f"(f[2]), "=f"(f[3]), "=f"(f[4]), "=f"(f[5]), "=f"(f[6]),
                 "=f"(f[7])
               : "l"(p));
}
__device__ __forceinline__ void st_stream_f8(float* p, const float f[8]) {
  asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p),
               "f"(f[0]), "f"(f[1]), "f"(f[2]), "f"(f[3]), "f"(f[4]), "f"(f[5]), "f"(f[6]), "f"(f[7])
               : "memory");
}
__device__ __forceinline__ void st_stream_f4(float* p, const float4& v) {
  st_stream_16(p, make_uint4(__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z),
                             __float_as_uint(v.w)));
}


// ---- Blackwell PTX helpers shared by the TMA / tcgen05 kernels (gemm_tcgen05*.cu, push_tma.cu) --
// mbarriers, bulk and tensor copies of the async proxy, tcgen05 fences and TMEM loads.  One copy:
// three kernels grew their own in round 1.
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t r[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
        "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu_add(uint32_t* p, uint32_t v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void bulk_load(void* smem, const void* gmem, uint32_t bytes,
                                          uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(smem)),
      "l"(gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void bulk_store(void* gmem, const void* smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem),
               "r"(smem_u32(smem)), "r"(bytes)
               : "memory");
}

// ---- bf16 packing ----------------------------------------------------------------
__device__ __forceinline__ void unpack_bf16x2(uint32_t u, float& lo, float& hi) {
  lo = __uint_as_float(u << 16);
  hi = __uint_as_float(u & 0xffff0000u);
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void unpack8(const uint4& u, float f[8]) {
  unpack_bf16x2(u.x, f[0], f[1]);
  unpack_bf16x2(u.y, f[2], f[3]);
  unpack_bf16x2(u.z, f[4], f[5]);
  unpack_bf16x2(u.w, f[6], f[7]);
}
__device__ __forceinline__ uint4 pack8(const float f[8]) {
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]),
                    pack_bf16x2(f[6], f[7]));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- optimizer math on 8 contiguous elements --------------------------------------
// Everything the server-side update needs, pre-digested once per thread.
struct OptimCoef {
  int kind;
  float lr, momentum, beta1, beta2, eps, wd;
  float inv_bias1;       // 1 / (1 - beta1^t)
  float inv_sqrt_bias2;  // 1 / sqrt(1 - beta2^t)
};

__device__ __forceinline__ OptimCoef make_coef(const PsbOptim& o, uint32_t completed_steps) {
  OptimCoef c;
  c.kind = o.kind;
  c.lr = o.lr;
  c.momentum = o.momentum;
  c.beta1 = o.beta1;
  c.beta2 = o.beta2;
  c.eps = o.eps;
  c.wd = o.weight_decay;
  c.inv_bias1 = 1.f;
  c.inv_sqrt_bias2 = 1.f;
  if (o.kind == PSB_OPT_ADAM || o.kind == PSB_OPT_ADAMW) {
    float t = (float)(completed_steps + 1u);
    c.inv_bias1 = 1.f / (1.f - powf(o.beta1, t));
    c.inv_sqrt_bias2 = rsqrtf(1.f - powf(o.beta2, t));
  }
  return c;
}

// p, m, v are updated in place; g is the already scaled / clipped mean gradient.
template <int KIND>
__device__ __forceinline__ void update1(const OptimCoef& c, float& p, float& m, float& v, float g) {
  if constexpr (KIND == PSB_OPT_SGD) {
    g = fmaf(c.wd, p, g);
    p = fmaf(-c.lr, g, p);
  } else if constexpr (KIND == PSB_OPT_MOMENTUM) {
    g = fmaf(c.wd, p, g);
    m = fmaf(c.momentum, m, g);
    p = fmaf(-c.lr, m, p);
  } else {
    if constexpr (KIND == PSB_OPT_ADAM) {
      g = fmaf(c.wd, p, g);
    } else {
      p = p * (1.f - c.lr * c.wd);
    }
    m = fmaf(c.beta1, m, (1.f - c.beta1) * g);
    v = fmaf(c.beta2, v, (1.f - c.beta2) * g * g);
    float denom = fmaf(sqrtf(v), c.inv_sqrt_bias2, c.eps);
    p = fmaf(-c.lr * c.inv_bias1, m / denom, p);
  }
}

}  // namespace psb
