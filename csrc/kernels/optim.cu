// optim.cu — K3 (flat-shard optimizer), K4 (norm / clip), K6 (cast / scale / multi-tensor
// gather), K7 (fill).  All are HBM-bound streaming kernels: 16-byte vector accesses,
// grid sized to a multiple of the 148 SMs, no shared-memory staging (no reuse).
//
// Reference hot loops these replace (paths below /root/reference):
//   H3  src/parameter_server.cpp:83-90   p[j] -= g[j]
//   H2  src/parameter_server.cpp:57-60   acc[j] /= W
//   H4  src/worker.cpp:436-438           g[j] /= num_gpus
//   H5  src/worker.cpp:317-320           g[j] = 0.01f
#include "common.cuh"

namespace psb {

thread_local char g_err[512] = {0};
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static inline int grid_for(size_t nvec, int threads, int waves_cap = 8) {
  size_t blocks = (nvec + threads - 1) / threads;
  size_t cap = (size_t)kNumSMs * waves_cap;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

// ---------------------------------------------------------------- K7 fill
template <typename T>
__global__ void __launch_bounds__(256) fill_kernel(T* __restrict__ dst, size_t n, T value) {
  constexpr int V = 16 / sizeof(T);
  size_t nvec = n / V;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  union {
    uint4 u;
    T t[V];
  } pk;
#pragma unroll
  for (int i = 0; i < V; ++i) pk.t[i] = value;
  // head: elements before the first 16-byte boundary are handled with the tail
  uintptr_t addr = reinterpret_cast<uintptr_t>(dst);
  size_t head = ((16 - (addr & 15)) & 15) / sizeof(T);
  if (head > n) head = n;
  T* body = dst + head;
  nvec = (n - head) / V;
  for (size_t i = tid; i < nvec; i += stride) reinterpret_cast<uint4*>(body)[i] = pk.u;
  size_t done = head + nvec * V;
  for (size_t i = tid; i < head; i += stride) dst[i] = value;
  for (size_t i = done + tid; i < n; i += stride) dst[i] = value;
}

// ---------------------------------------------------------------- K6 cast / scale
__global__ void __launch_bounds__(256)
    cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, size_t n,
                         float scale) {
  size_t nvec = n / 8;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    float4 a = ld_stream_f4(src + i * 8);
    float4 b = ld_stream_f4(src + i * 8 + 4);
    float f[8] = {a.x * scale, a.y * scale, a.z * scale, a.w * scale,
                  b.x * scale, b.y * scale, b.z * scale, b.w * scale};
    st_stream_16(dst + i * 8, pack8(f));
  }
  for (size_t i = nvec * 8 + tid; i < n; i += stride) dst[i] = __float2bfloat16_rn(src[i] * scale);
}

__global__ void __launch_bounds__(256)
    cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, size_t n,
                         float scale) {
  size_t nvec = n / 8;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    float f[8];
    unpack8(ld_stream_16(src + i * 8), f);
    st_stream_f4(dst + i * 8, make_float4(f[0] * scale, f[1] * scale, f[2] * scale, f[3] * scale));
    st_stream_f4(dst + i * 8 + 4,
                 make_float4(f[4] * scale, f[5] * scale, f[6] * scale, f[7] * scale));
  }
  for (size_t i = nvec * 8 + tid; i < n; i += stride) dst[i] = __bfloat162float(src[i]) * scale;
}

// Multi-tensor gather: blockIdx.y selects the tensor, blockIdx.x strides over it.
template <typename SRC, typename DST>
__global__ void __launch_bounds__(256) gather_kernel(PsbGatherList l, DST* __restrict__ bucket) {
  const int t = blockIdx.y;
  const SRC* __restrict__ src = reinterpret_cast<const SRC*>(l.src[t]);
  DST* __restrict__ dst = bucket + l.dst_offset[t];
  const size_t n = l.numel[t];
  const float scale = l.scale;
  const size_t nvec = n / 8;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    float f[8];
    if constexpr (sizeof(SRC) == 2) {
      unpack8(ld_stream_16(src + i * 8), f);
    } else {
      float4 a = ld_stream_f4(reinterpret_cast<const float*>(src) + i * 8);
      float4 b = ld_stream_f4(reinterpret_cast<const float*>(src) + i * 8 + 4);
      f[0] = a.x, f[1] = a.y, f[2] = a.z, f[3] = a.w, f[4] = b.x, f[5] = b.y, f[6] = b.z, f[7] = b.w;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] *= scale;
    if constexpr (sizeof(DST) == 2) {
      st_stream_16(dst + i * 8, pack8(f));
    } else {
      st_stream_f4(reinterpret_cast<float*>(dst) + i * 8, make_float4(f[0], f[1], f[2], f[3]));
      st_stream_f4(reinterpret_cast<float*>(dst) + i * 8 + 4, make_float4(f[4], f[5], f[6], f[7]));
    }
  }
  for (size_t i = nvec * 8 + tid; i < n; i += stride) {
    float x;
    if constexpr (sizeof(SRC) == 2) {
      x = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(src)[i]);
    } else {
      x = reinterpret_cast<const float*>(src)[i];
    }
    x *= scale;
    if constexpr (sizeof(DST) == 2) {
      reinterpret_cast<__nv_bfloat16*>(dst)[i] = __float2bfloat16_rn(x);
    } else {
      reinterpret_cast<float*>(dst)[i] = x;
    }
  }
}

// ---------------------------------------------------------------- K3 optimizer
// One thread owns 8 contiguous elements per iteration: 16 B of bf16 gradient (or 32 B fp32),
// 32 B each of master / m / v, 16 B of bf16 output.  Two vectors are kept in flight.
template <int KIND, int GDT, bool OUT>
__global__ void __launch_bounds__(256)
    optim_kernel(PsbOptim opt, float* __restrict__ master, float* __restrict__ m,
                 float* __restrict__ v, const void* __restrict__ grad,
                 __nv_bfloat16* __restrict__ out, size_t n, const float* __restrict__ clip_coef,
                 const uint32_t* __restrict__ step) {
  const OptimCoef c = make_coef(opt, step ? *step : 0u);
  const float gs = opt.grad_scale * (clip_coef ? *clip_coef : 1.f);
  const size_t nvec = n / 8;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    const size_t e = i * 8;
    float g[8], p[8], mm[8] = {}, vv[8] = {};
    if constexpr (GDT == PSB_BF16) {
      unpack8(ld_stream_16(reinterpret_cast<const __nv_bfloat16*>(grad) + e), g);
    } else {
      float4 a = ld_stream_f4(reinterpret_cast<const float*>(grad) + e);
      float4 b = ld_stream_f4(reinterpret_cast<const float*>(grad) + e + 4);
      g[0] = a.x, g[1] = a.y, g[2] = a.z, g[3] = a.w, g[4] = b.x, g[5] = b.y, g[6] = b.z, g[7] = b.w;
    }
    {
      float4 a = ld_stream_f4(master + e), b = ld_stream_f4(master + e + 4);
      p[0] = a.x, p[1] = a.y, p[2] = a.z, p[3] = a.w, p[4] = b.x, p[5] = b.y, p[6] = b.z, p[7] = b.w;
    }
    if constexpr (KIND != PSB_OPT_SGD) {
      float4 a = ld_stream_f4(m + e), b = ld_stream_f4(m + e + 4);
      mm[0] = a.x, mm[1] = a.y, mm[2] = a.z, mm[3] = a.w, mm[4] = b.x, mm[5] = b.y, mm[6] = b.z,
      mm[7] = b.w;
    }
    if constexpr (KIND == PSB_OPT_ADAM || KIND == PSB_OPT_ADAMW) {
      float4 a = ld_stream_f4(v + e), b = ld_stream_f4(v + e + 4);
      vv[0] = a.x, vv[1] = a.y, vv[2] = a.z, vv[3] = a.w, vv[4] = b.x, vv[5] = b.y, vv[6] = b.z,
      vv[7] = b.w;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) update1<KIND>(c, p[k], mm[k], vv[k], g[k] * gs);
    st_stream_f4(master + e, make_float4(p[0], p[1], p[2], p[3]));
    st_stream_f4(master + e + 4, make_float4(p[4], p[5], p[6], p[7]));
    if constexpr (KIND != PSB_OPT_SGD) {
      st_stream_f4(m + e, make_float4(mm[0], mm[1], mm[2], mm[3]));
      st_stream_f4(m + e + 4, make_float4(mm[4], mm[5], mm[6], mm[7]));
    }
    if constexpr (KIND == PSB_OPT_ADAM || KIND == PSB_OPT_ADAMW) {
      st_stream_f4(v + e, make_float4(vv[0], vv[1], vv[2], vv[3]));
      st_stream_f4(v + e + 4, make_float4(vv[4], vv[5], vv[6], vv[7]));
    }
    if constexpr (OUT) st_stream_16(out + e, pack8(p));
  }
  // scalar tail (n not a multiple of 8)
  for (size_t e = nvec * 8 + tid; e < n; e += stride) {
    float g = GDT == PSB_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(grad)[e])
                              : reinterpret_cast<const float*>(grad)[e];
    float p = master[e], mm = 0.f, vv = 0.f;
    if constexpr (KIND != PSB_OPT_SGD) mm = m[e];
    if constexpr (KIND == PSB_OPT_ADAM || KIND == PSB_OPT_ADAMW) vv = v[e];
    update1<KIND>(c, p, mm, vv, g * gs);
    master[e] = p;
    if constexpr (KIND != PSB_OPT_SGD) m[e] = mm;
    if constexpr (KIND == PSB_OPT_ADAM || KIND == PSB_OPT_ADAMW) v[e] = vv;
    if constexpr (OUT) out[e] = __float2bfloat16_rn(p);
  }
}

__global__ void counter_add_kernel(uint32_t* c, uint32_t d) { *c += d; }

// ---------------------------------------------------------------- K4 norm / clip
constexpr int kSumsqBlocks = kNumSMs * 4;
constexpr int kSumsqThreads = 256;

template <int DT>
__global__ void __launch_bounds__(kSumsqThreads)
    sumsq_partial_kernel(const void* __restrict__ x, size_t n, float* __restrict__ partials) {
  const size_t nvec = n / 8;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  float acc = 0.f;
  for (size_t i = tid; i < nvec; i += stride) {
    float f[8];
    if constexpr (DT == PSB_BF16) {
      unpack8(ld_stream_16(reinterpret_cast<const __nv_bfloat16*>(x) + i * 8), f);
    } else {
      float4 a = ld_stream_f4(reinterpret_cast<const float*>(x) + i * 8);
      float4 b = ld_stream_f4(reinterpret_cast<const float*>(x) + i * 8 + 4);
      f[0] = a.x, f[1] = a.y, f[2] = a.z, f[3] = a.w, f[4] = b.x, f[5] = b.y, f[6] = b.z, f[7] = b.w;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) acc = fmaf(f[k], f[k], acc);
  }
  for (size_t e = nvec * 8 + tid; e < n; e += stride) {
    float f = DT == PSB_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(x)[e])
                             : reinterpret_cast<const float*>(x)[e];
    acc = fmaf(f, f, acc);
  }
  __shared__ float warp_part[kSumsqThreads / 32];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float w = threadIdx.x < kSumsqThreads / 32 ? warp_part[threadIdx.x] : 0.f;
    w = warp_sum(w);
    if (threadIdx.x == 0) partials[blockIdx.x] = w;  // every slot is overwritten: no zeroing needed
  }
}

// fixed-order final reduction => deterministic result
__global__ void __launch_bounds__(256)
    sumsq_final_kernel(const float* __restrict__ partials, int n_parts, float* __restrict__ out) {
  __shared__ float sh[256];
  float acc = 0.f;
  for (int i = threadIdx.x; i < n_parts; i += 256) acc += partials[i];
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = sh[0];
}

__global__ void clip_coef_kernel(const float* __restrict__ parts, int n_parts, float max_norm,
                                 float eps, float* __restrict__ coef_out,
                                 float* __restrict__ norm_out) {
  float s = 0.f;
  for (int i = 0; i < n_parts; ++i) s += parts[i];
  float norm = sqrtf(s);
  if (norm_out) *norm_out = norm;
  float c = max_norm / (norm + eps);
  *coef_out = c < 1.f ? c : 1.f;
}

}  // namespace psb

using namespace psb;

extern "C" {

const char* psb_last_error(void) { return psb::g_err; }
int psb_version(void) { return 100; }
uint64_t psb_launch_count(void) { return psb::g_launches.load(); }

int psb_fill_f32(float* dst, size_t n, float value, psb_stream_t stream) {
  if (n == 0) return 0;
  fill_kernel<float><<<grid_for(n / 4 + 1, 256), 256, 0, (cudaStream_t)stream>>>(dst, n, value);
  PSB_LAUNCH_OK("fill_f32");
  return 0;
}

int psb_fill_bf16(void* dst, size_t n, float value, psb_stream_t stream) {
  if (n == 0) return 0;
  fill_kernel<__nv_bfloat16><<<grid_for(n / 8 + 1, 256), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<__nv_bfloat16*>(dst), n, __float2bfloat16_rn(value));
  PSB_LAUNCH_OK("fill_bf16");
  return 0;
}

int psb_cast_f32_to_bf16(const float* src, void* dst, size_t n, float scale, psb_stream_t stream) {
  if (n == 0) return 0;
  PSB_REQUIRE(((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 15) == 0, PSB_EINVAL,
              "psb_cast_f32_to_bf16: pointers must be 16-byte aligned");
  cast_f32_bf16_kernel<<<grid_for(n / 8 + 1, 256), 256, 0, (cudaStream_t)stream>>>(
      src, reinterpret_cast<__nv_bfloat16*>(dst), n, scale);
  PSB_LAUNCH_OK("cast_f32_to_bf16");
  return 0;
}

int psb_cast_bf16_to_f32(const void* src, float* dst, size_t n, float scale, psb_stream_t stream) {
  if (n == 0) return 0;
  PSB_REQUIRE(((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 15) == 0, PSB_EINVAL,
              "psb_cast_bf16_to_f32: pointers must be 16-byte aligned");
  cast_bf16_f32_kernel<<<grid_for(n / 8 + 1, 256), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(src), dst, n, scale);
  PSB_LAUNCH_OK("cast_bf16_to_f32");
  return 0;
}

int psb_gather_to_bucket(const PsbGatherList* l, void* bucket, psb_stream_t stream) {
  PSB_REQUIRE(l && bucket, PSB_EINVAL, "psb_gather_to_bucket: null argument");
  PSB_REQUIRE(l->count >= 0 && l->count <= PSB_MAX_GATHER, PSB_ERANGE,
              "psb_gather_to_bucket: count %d exceeds PSB_MAX_GATHER=%d", l->count, PSB_MAX_GATHER);
  if (l->count == 0) return 0;
  size_t max_n = 0;
  for (int i = 0; i < l->count; ++i) {
    PSB_REQUIRE(((uintptr_t)l->src[i] & 15) == 0 && (l->dst_offset[i] & 7) == 0, PSB_EINVAL,
                "psb_gather_to_bucket: tensor %d is not 16-byte aligned", i);
    if (l->numel[i] > max_n) max_n = l->numel[i];
  }
  int gx = grid_for(max_n / 8 + 1, 256, 2);
  // keep the total grid near a few waves: many small tensors need few blocks each
  if (l->count > 8 && gx > kNumSMs) gx = kNumSMs;
  dim3 grid(gx, l->count);
  cudaStream_t s = (cudaStream_t)stream;
  if (l->src_dtype == PSB_BF16 && l->dst_dtype == PSB_BF16) {
    gather_kernel<__nv_bfloat16, __nv_bfloat16>
        <<<grid, 256, 0, s>>>(*l, reinterpret_cast<__nv_bfloat16*>(bucket));
  } else if (l->src_dtype == PSB_F32 && l->dst_dtype == PSB_BF16) {
    gather_kernel<float, __nv_bfloat16>
        <<<grid, 256, 0, s>>>(*l, reinterpret_cast<__nv_bfloat16*>(bucket));
  } else if (l->src_dtype == PSB_F32 && l->dst_dtype == PSB_F32) {
    gather_kernel<float, float><<<grid, 256, 0, s>>>(*l, reinterpret_cast<float*>(bucket));
  } else if (l->src_dtype == PSB_BF16 && l->dst_dtype == PSB_F32) {
    gather_kernel<__nv_bfloat16, float><<<grid, 256, 0, s>>>(*l, reinterpret_cast<float*>(bucket));
  } else {
    PSB_REQUIRE(false, PSB_EINVAL, "psb_gather_to_bucket: bad dtypes %d -> %d", l->src_dtype,
                l->dst_dtype);
  }
  PSB_LAUNCH_OK("gather_to_bucket");
  return 0;
}

}  // extern "C"

template <int KIND, int GDT>
static void launch_optim(const PsbOptim* opt, float* master, float* m, float* v, const void* grad,
                         void* out, size_t n, const float* clip, const uint32_t* step,
                         cudaStream_t s) {
  int grid = grid_for(n / 8 + 1, 256, 4);
  if (out) {
    optim_kernel<KIND, GDT, true><<<grid, 256, 0, s>>>(
        *opt, master, m, v, grad, reinterpret_cast<__nv_bfloat16*>(out), n, clip, step);
  } else {
    optim_kernel<KIND, GDT, false><<<grid, 256, 0, s>>>(*opt, master, m, v, grad, nullptr, n, clip,
                                                        step);
  }
}

extern "C" {

int psb_optim_step(const PsbOptim* opt, float* master, float* m, float* v, const void* grad,
                   int grad_dtype, void* param_bf16_out, size_t n, const float* clip_coef,
                   const uint32_t* step, psb_stream_t stream) {
  PSB_REQUIRE(opt && master && grad, PSB_EINVAL, "psb_optim_step: null argument");
  if (n == 0) return 0;
  PSB_REQUIRE(grad_dtype == PSB_F32 || grad_dtype == PSB_BF16, PSB_EINVAL,
              "psb_optim_step: bad grad dtype %d", grad_dtype);
  PSB_REQUIRE(opt->kind == PSB_OPT_SGD || m, PSB_EINVAL, "psb_optim_step: momentum buffer missing");
  PSB_REQUIRE((opt->kind != PSB_OPT_ADAM && opt->kind != PSB_OPT_ADAMW) || v, PSB_EINVAL,
              "psb_optim_step: second-moment buffer missing");
  uintptr_t al = (uintptr_t)master | (uintptr_t)grad | (uintptr_t)param_bf16_out | (uintptr_t)m |
                 (uintptr_t)v;
  PSB_REQUIRE((al & 15) == 0, PSB_EINVAL, "psb_optim_step: pointers must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
#define PSB_DISPATCH(K)                                                                     \
  case K:                                                                                   \
    if (grad_dtype == PSB_BF16)                                                             \
      launch_optim<K, PSB_BF16>(opt, master, m, v, grad, param_bf16_out, n, clip_coef, step, s); \
    else                                                                                    \
      launch_optim<K, PSB_F32>(opt, master, m, v, grad, param_bf16_out, n, clip_coef, step, s);  \
    break;
  switch (opt->kind) {
    PSB_DISPATCH(PSB_OPT_SGD)
    PSB_DISPATCH(PSB_OPT_MOMENTUM)
    PSB_DISPATCH(PSB_OPT_ADAM)
    PSB_DISPATCH(PSB_OPT_ADAMW)
    default:
      PSB_REQUIRE(false, PSB_EINVAL, "psb_optim_step: unknown optimizer kind %d", opt->kind);
  }
#undef PSB_DISPATCH
  PSB_LAUNCH_OK("optim_step");
  return 0;
}

int psb_counter_add(uint32_t* counter, uint32_t delta, psb_stream_t stream) {
  counter_add_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(counter, delta);
  PSB_LAUNCH_OK("counter_add");
  return 0;
}

size_t psb_sumsq_scratch_floats(void) { return kSumsqBlocks; }

int psb_sumsq(const void* x, int dtype, size_t n, float* partials, float* out, psb_stream_t stream) {
  PSB_REQUIRE(x && partials && out, PSB_EINVAL, "psb_sumsq: null argument");
  PSB_REQUIRE(((uintptr_t)x & 15) == 0, PSB_EINVAL, "psb_sumsq: input must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  int blocks = grid_for(n / 8 + 1, kSumsqThreads, 4);
  if (blocks > kSumsqBlocks) blocks = kSumsqBlocks;
  if (dtype == PSB_BF16) {
    sumsq_partial_kernel<PSB_BF16><<<blocks, kSumsqThreads, 0, s>>>(x, n, partials);
  } else if (dtype == PSB_F32) {
    sumsq_partial_kernel<PSB_F32><<<blocks, kSumsqThreads, 0, s>>>(x, n, partials);
  } else {
    PSB_REQUIRE(false, PSB_EINVAL, "psb_sumsq: bad dtype %d", dtype);
  }
  PSB_LAUNCH_OK("sumsq_partial");
  sumsq_final_kernel<<<1, 256, 0, s>>>(partials, blocks, out);
  PSB_LAUNCH_OK("sumsq_final");
  return 0;
}

int psb_clip_coef(const float* sumsq_parts, int n_parts, float max_norm, float eps, float* coef_out,
                  float* norm_out, psb_stream_t stream) {
  PSB_REQUIRE(sumsq_parts && coef_out && n_parts > 0, PSB_EINVAL, "psb_clip_coef: bad argument");
  clip_coef_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(sumsq_parts, n_parts, max_norm, eps, coef_out,
                                                      norm_out);
  PSB_LAUNCH_OK("clip_coef");
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------
// psb_sumsq_noalloc — Σx² for callers that cannot allocate device memory.
//
// Same two-phase reduction as psb_sumsq, but the per-block partial sums go to the file-scope
// array below instead of a caller-provided buffer.  Every GPU that loads this library keeps
// the array (PSB_SUMSQ_NOALLOC_SLOTS floats = 4 KB) until the process exits.
//
// In host code `g_sumsq_noalloc_scratch` is only a host-side stand-in for the symbol, not a
// device address; the address the kernels receive is looked up once per device with
// cudaGetSymbolAddress while that device is current (noalloc_scratch_for_current_device).
// ---------------------------------------------------------------------------------------
namespace psb {

static __device__ __align__(256) float g_sumsq_noalloc_scratch[PSB_SUMSQ_NOALLOC_SLOTS];

constexpr int kMaxDevices = 64;
static float* g_noalloc_ptr[kMaxDevices] = {nullptr};
static std::atomic<bool> g_noalloc_ready[kMaxDevices];

// Returns the device address of the scratch array on the current device (set-up on first use).
static int noalloc_scratch_for_current_device(float** out) {
  int dev = -1;
  PSB_CUDA_OK(cudaGetDevice(&dev));
  PSB_REQUIRE(dev >= 0 && dev < kMaxDevices, PSB_ERANGE,
              "psb_sumsq_noalloc: device ordinal %d outside [0, %d)", dev, kMaxDevices);
  if (!g_noalloc_ready[dev].load(std::memory_order_acquire)) {
    float* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_sumsq_noalloc_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      set_error("psb_sumsq_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : PSB_EINVAL;
    }
    g_noalloc_ptr[dev] = ptr;  // racing first calls store the same value
    g_noalloc_ready[dev].store(true, std::memory_order_release);
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}

}  // namespace psb

extern "C" int psb_sumsq_noalloc(const void* x, int dtype, size_t n, float* out,
                                 psb_stream_t stream) {
  PSB_REQUIRE(x && out, PSB_EINVAL, "psb_sumsq_noalloc: null argument");
  PSB_REQUIRE(((uintptr_t)x & 15) == 0, PSB_EINVAL,
              "psb_sumsq_noalloc: input must be 16-byte aligned");
  PSB_REQUIRE(dtype == PSB_BF16 || dtype == PSB_F32, PSB_EINVAL,
              "psb_sumsq_noalloc: bad dtype %d", dtype);
  PSB_REQUIRE(n <= (size_t)PSB_SUMSQ_NOALLOC_MAX_ELEMS, PSB_ERANGE,
              "psb_sumsq_noalloc: n = %llu exceeds PSB_SUMSQ_NOALLOC_MAX_ELEMS = %llu (the static "
              "scratch holds %d partial sums of at most %d elements each); use psb_sumsq with "
              "caller-provided scratch for larger inputs",
              (unsigned long long)n, (unsigned long long)PSB_SUMSQ_NOALLOC_MAX_ELEMS,
              PSB_SUMSQ_NOALLOC_SLOTS, PSB_SUMSQ_NOALLOC_ELEMS_PER_SLOT);
  float* scratch = nullptr;
  if (int rc = psb::noalloc_scratch_for_current_device(&scratch)) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  // one block (= one partial slot) per PSB_SUMSQ_NOALLOC_ELEMS_PER_SLOT elements at most, and
  // never more blocks than the regular entry point would launch
  size_t by_span = (n + PSB_SUMSQ_NOALLOC_ELEMS_PER_SLOT - 1) / PSB_SUMSQ_NOALLOC_ELEMS_PER_SLOT;
  int blocks = psb::grid_for(n / 8 + 1, psb::kSumsqThreads, 4);
  if (blocks > PSB_SUMSQ_NOALLOC_SLOTS) blocks = PSB_SUMSQ_NOALLOC_SLOTS;
  if ((size_t)blocks < by_span) blocks = (int)by_span;  // by_span <= SLOTS by the check above
  // Every slot in [0, blocks) is overwritten by the first kernel and only those are read by the
  // second, so nothing in the shared array has to start at zero.
  if (dtype == PSB_BF16) {
    psb::sumsq_partial_kernel<PSB_BF16><<<blocks, psb::kSumsqThreads, 0, s>>>(x, n, scratch);
  } else {
    psb::sumsq_partial_kernel<PSB_F32><<<blocks, psb::kSumsqThreads, 0, s>>>(x, n, scratch);
  }
  PSB_LAUNCH_OK("sumsq_noalloc_partial");
  psb::sumsq_final_kernel<<<1, 256, 0, s>>>(scratch, blocks, out);
  PSB_LAUNCH_OK("sumsq_noalloc_final");
  return 0;
}
