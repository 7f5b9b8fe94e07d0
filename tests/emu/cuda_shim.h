// cuda_shim.h — just enough of the CUDA device language to run the bodies of simple kernels on
// the host, one "thread" after the other (tests/test_sparse_kernel_emulation.py).  Valid only
// for kernels without barriers, shuffles or spin-waits: every thread runs to completion before
// the next one starts, so atomics degenerate to plain read-modify-writes.
#pragma once
#include <stdint.h>

#include <cmath>
#include <cstring>

#include "psb_kernels.h"

#define __global__
#define __device__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(x)

struct uint3_ { unsigned x = 0, y = 0, z = 0; };
static uint3_ threadIdx, blockIdx, blockDim, gridDim;

struct uint4 { uint32_t x, y, z, w; };
struct alignas(16) float4 { float x, y, z, w; };
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return {x, y, z, w}; }
static inline float4 make_float4(float x, float y, float z, float w) { return {x, y, z, w}; }

struct __nv_bfloat16 { uint16_t bits; };
static inline float bf16_to_float(__nv_bfloat16 h) {
  uint32_t u = (uint32_t)h.bits << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static inline __nv_bfloat16 float_to_bf16_rn(float f) {  // round to nearest even (finite inputs)
  uint32_t u;
  memcpy(&u, &f, 4);
  u += 0x7fffu + ((u >> 16) & 1u);
  return {(uint16_t)(u >> 16)};
}

static inline int atomicCAS(int* p, int cmp, int val) {
  int old = *p;
  if (old == cmp) *p = val;
  return old;
}
static inline uint32_t atomicAdd(uint32_t* p, uint32_t v) {
  uint32_t old = *p;
  *p += v;
  return old;
}
static inline float4 atomicAdd(float4* p, float4 v) {
  float4 old = *p;
  p->x += v.x, p->y += v.y, p->z += v.z, p->w += v.w;
  return old;
}
static inline float rsqrtf(float x) { return 1.f / sqrtf(x); }

namespace psb {
constexpr int kNumSMs = 148;
static inline uint4 ld_stream_16(const void* p) {
  uint4 r;
  memcpy(&r, p, 16);
  return r;
}
static inline void st_stream_16(void* p, const uint4& v) { memcpy(p, &v, 16); }
static inline float4 ld_stream_f4(const float* p) {
  float4 r;
  memcpy(&r, p, 16);
  return r;
}
static inline void unpack8(const uint4& u, float f[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = bf16_to_float({(uint16_t)(w[i] & 0xffffu)});
    f[2 * i + 1] = bf16_to_float({(uint16_t)(w[i] >> 16)});
  }
}
static inline uint4 pack8(const float f[8]) {
  uint32_t w[4];
  for (int i = 0; i < 4; ++i)
    w[i] = (uint32_t)float_to_bf16_rn(f[2 * i]).bits | ((uint32_t)float_to_bf16_rn(f[2 * i + 1]).bits << 16);
  return {w[0], w[1], w[2], w[3]};
}
}  // namespace psb

// run `kernel(args...)` for every thread of a 1-D grid, in order
template <class K, class... A>
static void emu_launch(K kernel, int grid, int block, A... args) {
  gridDim.x = grid, blockDim.x = block;
  for (int b = 0; b < grid; ++b)
    for (int t = 0; t < block; ++t) {
      blockIdx.x = b, threadIdx.x = t;
      kernel(args...);
    }
}
