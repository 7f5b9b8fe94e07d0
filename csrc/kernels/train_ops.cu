// train_ops.cu — small memory-bound kernels of the training step that sit between the GEMMs of
// the flagship MLP: bias gradient (column sum of dZ) and the MSE loss with its gradient.
// The reference has no model code at all (src/worker.cpp:316-320 fabricates gradients); these
// replace chains of generic PyTorch elementwise / reduce launches, which took ~100 us
// (reduce_kernel) and ~60 us (loss: two casts, mse, mean, fills, backward) of a 2.1 ms step
// before; now 16-19 us per 64 MB column sum and 9 + 7 us for the loss
// (profiles/timeline_mlp_1gpu.tsv).
#include "common.cuh"

namespace psb {

// ------------------------------------------------------------------------------------
// Column sum: out[n] = sum_m x[m, n], x bf16 [M, N] row-major, fp32 accumulation.
// Block (strip, slab): 256 threads = 8 warps; a warp covers 256 columns with 16-byte loads,
// the warps of a block take rows slab_begin + warp, + 8, ...  Partials [slabs][N] in fp32; the
// last block of a strip (arrival counter) adds them in slab order, so the result does not
// depend on scheduling.
// ------------------------------------------------------------------------------------
constexpr int kColThreads = 256;
constexpr int kColStrip = 256;
constexpr int kColMaxStrips = 1024;  // arrival counters sit at the FRONT of the scratch, so their
                                     // place does not depend on the problem size (a scratch is
                                     // reused across shapes and must find them at zero)

__global__ void __launch_bounds__(kColThreads) colsum_bf16_kernel(
    const __nv_bfloat16* __restrict__ x, void* out, int out_dtype, int M, int N, int slabs,
    float* partial, uint32_t* counters) {
  const int strip = blockIdx.x, slab = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int col = strip * kColStrip + lane * 8;
  const int rows_per_slab = (M + slabs - 1) / slabs;
  const int r0 = slab * rows_per_slab;
  const int r1 = min(M, r0 + rows_per_slab);
  float acc[8] = {};
  if (col < N) {
    const __nv_bfloat16* src = x + col;
    int r = r0 + warp;
    for (; r + 24 < r1; r += 32) {  // 4 rows in flight per thread
      uint4 v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] = ld_stream_16(src + (size_t)(r + 8 * k) * N);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float f[8];
        unpack8(v[k], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
    for (; r < r1; r += 8) {
      float f[8];
      unpack8(ld_stream_16(src + (size_t)r * N), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
  }
  __shared__ float red[kColThreads / 32][kColStrip];
  __shared__ int is_last;
#pragma unroll
  for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = acc[j];
  __syncthreads();
  const int c = strip * kColStrip + threadIdx.x;  // one column per thread from here on
  float s = 0.f;
#pragma unroll
  for (int w = 0; w < kColThreads / 32; ++w) s += red[w][threadIdx.x];
  if (slabs == 1) {
    if (c < N) {
      if (out_dtype == PSB_BF16) reinterpret_cast<__nv_bfloat16*>(out)[c] = __float2bfloat16_rn(s);
      else reinterpret_cast<float*>(out)[c] = s;
    }
    return;
  }
  if (c < N) partial[(size_t)slab * N + c] = s;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    is_last = atomicAdd(counters + strip, 1u) == (uint32_t)slabs - 1;
    if (is_last) counters[strip] = 0u;  // ready for the next launch on this stream
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (c < N) {
    float t = 0.f;
    for (int k = 0; k < slabs; ++k) t += __ldcg(partial + (size_t)k * N + c);
    if (out_dtype == PSB_BF16) reinterpret_cast<__nv_bfloat16*>(out)[c] = __float2bfloat16_rn(t);
    else reinterpret_cast<float*>(out)[c] = t;
  }
}

// ------------------------------------------------------------------------------------
// MSE loss, mean over all n elements, bf16 inputs.  Forward: loss = sum((a-b)^2) / n with a
// fixed-order two-level reduction (block partials, last block adds them).  Backward:
// da = 2 (a-b) / n * (*gout), bf16.
// ------------------------------------------------------------------------------------
constexpr int kMseThreads = 256;

__global__ void __launch_bounds__(kMseThreads) mse_fwd_kernel(const __nv_bfloat16* __restrict__ a,
                                                              const __nv_bfloat16* __restrict__ b,
                                                              size_t n, float* partial,
                                                              uint32_t* counter, float* loss) {
  const size_t nvec = n / 8;
  float s = 0.f;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (size_t)gridDim.x * blockDim.x) {
    float fa[8], fb[8];
    unpack8(ld_stream_16(a + i * 8), fa);
    unpack8(ld_stream_16(b + i * 8), fb);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float d = fa[k] - fb[k];
      s = fmaf(d, d, s);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (size_t i = nvec * 8; i < n; ++i) {  // tail
      const float d = __bfloat162float(a[i]) - __bfloat162float(b[i]);
      s = fmaf(d, d, s);
    }
  __shared__ float wsum[kMseThreads / 32];
  __shared__ int is_last;
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < kMseThreads / 32; ++w) t += wsum[w];
    partial[blockIdx.x] = t;
    __threadfence();
    is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
    if (is_last) *counter = 0u;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  float t = 0.f;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) t += __ldcg(partial + i);
  t = warp_sum(t);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    float r = 0.f;
    for (int w = 0; w < kMseThreads / 32; ++w) r += wsum[w];
    *loss = r / (float)n;
  }
}

__global__ void __launch_bounds__(kMseThreads) mse_bwd_kernel(const __nv_bfloat16* __restrict__ a,
                                                              const __nv_bfloat16* __restrict__ b,
                                                              const float* __restrict__ gout,
                                                              size_t n, __nv_bfloat16* da) {
  const float scale = 2.f / (float)n * (gout ? __ldg(gout) : 1.f);
  const size_t nvec = n / 8;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (size_t)gridDim.x * blockDim.x) {
    float fa[8], fb[8];
    unpack8(ld_stream_16(a + i * 8), fa);
    unpack8(ld_stream_16(b + i * 8), fb);
#pragma unroll
    for (int k = 0; k < 8; ++k) fa[k] = (fa[k] - fb[k]) * scale;
    st_stream_16(da + i * 8, pack8(fa));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (size_t i = nvec * 8; i < n; ++i)
      da[i] = __float2bfloat16_rn((__bfloat162float(a[i]) - __bfloat162float(b[i])) * scale);
}

}  // namespace psb

using namespace psb;

extern "C" {

int psb_colsum_slabs(int M, int N) {
  // enough blocks for ~4 per SM, at least 64 rows per slab, at most 64 slabs
  const int strips = (N + kColStrip - 1) / kColStrip;
  int slabs = (4 * kNumSMs + strips - 1) / strips;
  const int max_by_rows = M / 64 > 0 ? M / 64 : 1;
  if (slabs > max_by_rows) slabs = max_by_rows;
  if (slabs > 64) slabs = 64;
  return slabs < 1 ? 1 : slabs;
}

size_t psb_colsum_scratch_bytes(int M, int N) {
  const int strips = (N + kColStrip - 1) / kColStrip;
  (void)strips;
  return (size_t)kColMaxStrips * sizeof(uint32_t) + (size_t)psb_colsum_slabs(M, N) * N * sizeof(float);
}

int psb_colsum_bf16(const void* x, void* out, int out_dtype, int M, int N, void* scratch,
                    psb_stream_t stream) {
  PSB_REQUIRE(x && out && scratch, PSB_EINVAL, "psb_colsum_bf16: null pointer");
  PSB_REQUIRE(M > 0 && N > 0 && N % 8 == 0, PSB_EINVAL,
              "psb_colsum_bf16: need M > 0 and N a positive multiple of 8 (got %d x %d)", M, N);
  PSB_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)scratch & 15) == 0, PSB_EINVAL,
              "psb_colsum_bf16: x and scratch must be 16-byte aligned");
  PSB_REQUIRE(out_dtype == PSB_BF16 || out_dtype == PSB_F32, PSB_EINVAL,
              "psb_colsum_bf16: bad output dtype %d", out_dtype);
  const int strips = (N + kColStrip - 1) / kColStrip;
  const int slabs = psb_colsum_slabs(M, N);
  PSB_REQUIRE(strips <= kColMaxStrips, PSB_ERANGE, "psb_colsum_bf16: N=%d exceeds %d columns", N,
              kColMaxStrips * kColStrip);
  uint32_t* counters = reinterpret_cast<uint32_t*>(scratch);
  float* partial = reinterpret_cast<float*>(counters + kColMaxStrips);
  colsum_bf16_kernel<<<dim3(strips, slabs), kColThreads, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), out, out_dtype, M, N, slabs, partial, counters);
  PSB_LAUNCH_OK("colsum_bf16");
  return 0;
}

int psb_mse_blocks(size_t n) {
  size_t want = (n / 8 + kMseThreads * 4 - 1) / (kMseThreads * 4);
  return (int)(want < 1 ? 1 : want > (size_t)(2 * kNumSMs) ? 2 * kNumSMs : want);
}

int psb_mse_fwd_bf16(const void* a, const void* b, size_t n, float* scratch, float* loss,
                     psb_stream_t stream) {
  PSB_REQUIRE(a && b && scratch && loss && n > 0, PSB_EINVAL, "psb_mse_fwd_bf16: bad argument");
  PSB_REQUIRE((((uintptr_t)a | (uintptr_t)b) & 15) == 0, PSB_EINVAL,
              "psb_mse_fwd_bf16: inputs must be 16-byte aligned");
  const int grid = psb_mse_blocks(n);
  // scratch: [grid] partials, then the arrival counter (zero before the first use)
  mse_fwd_kernel<<<grid, kMseThreads, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(a), reinterpret_cast<const __nv_bfloat16*>(b), n,
      scratch, reinterpret_cast<uint32_t*>(scratch + 2 * kNumSMs), loss);
  PSB_LAUNCH_OK("mse_fwd_bf16");
  return 0;
}

int psb_mse_bwd_bf16(const void* a, const void* b, const float* gout, size_t n, void* da,
                     psb_stream_t stream) {
  PSB_REQUIRE(a && b && da && n > 0, PSB_EINVAL, "psb_mse_bwd_bf16: bad argument");
  PSB_REQUIRE((((uintptr_t)a | (uintptr_t)b | (uintptr_t)da) & 15) == 0, PSB_EINVAL,
              "psb_mse_bwd_bf16: tensors must be 16-byte aligned");
  mse_bwd_kernel<<<psb_mse_blocks(n), kMseThreads, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(a), reinterpret_cast<const __nv_bfloat16*>(b), gout, n,
      reinterpret_cast<__nv_bfloat16*>(da));
  PSB_LAUNCH_OK("mse_bwd_bf16");
  return 0;
}

}  // extern "C"
