// sparse.cu — K5: sparse embedding push / pull over the symmetric heap (sm_100a).
//
// The reference has no sparse path: its Tensor is dense (proto/parameter_server.proto:19-24)
// and a push always carries the whole model (src/worker.cpp:253-275).  An embedding table is the
// case where that is hopeless — a batch touches a few thousand of millions of rows — so the
// table is row-sharded over the GPUs and only the touched rows move:
//
//   pull    one sub-warp per requested row: 16-byte loads from the owner's bf16 shard (a peer
//           address when the owner is another GPU), 16-byte stores into the local output.
//   stage   worker side of a push, one launch: ids, gradient rows (cast to the staging dtype) and
//           the count into this rank's range of the symmetric heap.
//   apply   owner side of a push, three launches between two device barriers:
//             mark        a thread per staged (worker, entry): rows this rank owns get a slot in
//                         a compact accumulator (atomicCAS on slot[row] + an atomic counter)
//             accumulate  a sub-warp per staged entry: the gradient row is read from the
//                         worker's staging range (NVLink carries only rows owned here) and
//                         added into the slot's fp32 row with float4 atomics (RED.ADD.F32x4)
//             apply       a sub-warp per touched row: mean gradient through update1<> (the
//                         optimizer math of K1 / K3), fp32 master (+ m, v) and the bf16 row
//                         rewritten, accumulator row zeroed and slot released
//           Untouched rows cost nothing: the work is O(staged entries), not O(rows).
//
// All three are bandwidth / latency kernels; nothing here is GEMM-shaped.  Rows narrower than a
// warp's 512 bytes share a warp (8 or 16 lanes per row) so that small embedding widths keep the
// lanes busy.
#include "common.cuh"

namespace psb {
namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ int lanes_per_row(int vecs) { return vecs <= 8 ? 8 : (vecs <= 16 ? 16 : 32); }

// peer words written before the caller's barrier: strong (uncached) loads
__device__ __forceinline__ uint32_t ld_peer_u32(const void* p) {
  return *reinterpret_cast<const volatile uint32_t*>(p);
}
__device__ __forceinline__ int64_t ld_peer_i64(const void* p) {
  return *reinterpret_cast<const volatile int64_t*>(p);
}

// owner and local row of a global index; false for padding / out-of-range indices
__device__ __forceinline__ bool locate(const PsbSparse& s, int64_t id, int& owner, int64_t& lrow) {
  if (id < 0 || id >= s.vocab) return false;
  owner = (int)(id / s.rows_per_rank);
  lrow = id - (int64_t)owner * s.rows_per_rank;
  return owner < s.fabric.world;
}

__global__ void __launch_bounds__(kThreads)
    sparse_pull_kernel(PsbSparse s, const int64_t* __restrict__ idx, int64_t n,
                       __nv_bfloat16* __restrict__ out) {
  const int vecs = s.dim / 8;
  const int lpr = lanes_per_row(vecs);
  const int lane = threadIdx.x & 31, l = lane % lpr;
  const int64_t groups_per_warp = 32 / lpr;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = warp * groups_per_warp + lane / lpr; i < n; i += nwarps * groups_per_warp) {
    uint4* dst = reinterpret_cast<uint4*>(out + i * s.dim);
    int owner;
    int64_t lrow;
    if (!locate(s, idx[i], owner, lrow)) {
      for (int v = l; v < vecs; v += lpr) dst[v] = make_uint4(0u, 0u, 0u, 0u);
      continue;
    }
    const uint4* src = reinterpret_cast<const uint4*>(
        reinterpret_cast<const __nv_bfloat16*>(s.table[owner]) + lrow * s.dim);
    for (int v = l; v < vecs; v += lpr) st_stream_16(dst + v, ld_stream_16(src + v));
  }
}

// Worker half of a push in one launch: ids, gradient rows (cast to the staging dtype on the way)
// and the entry count into this rank's staging range.  A thread moves 8 elements.
template <int IN, int OUT>
__global__ void __launch_bounds__(kThreads)
    sparse_stage_kernel(const int64_t* __restrict__ idx, const void* __restrict__ grad, int64_t n,
                        int dim, int64_t* __restrict__ stage_idx, void* __restrict__ stage_grad,
                        uint32_t* __restrict__ stage_cnt) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  if (tid == 0) *stage_cnt = (uint32_t)n;
  for (int64_t i = tid; i < n; i += stride) stage_idx[i] = idx[i];
  const int64_t nvec = n * dim / 8;
  for (int64_t v = tid; v < nvec; v += stride) {
    float g[8];
    if constexpr (IN == PSB_BF16) {
      unpack8(ld_stream_16(reinterpret_cast<const __nv_bfloat16*>(grad) + v * 8), g);
    } else {
      const float* src = reinterpret_cast<const float*>(grad) + v * 8;
      const float4 x = ld_stream_f4(src), y = ld_stream_f4(src + 4);
      g[0] = x.x, g[1] = x.y, g[2] = x.z, g[3] = x.w, g[4] = y.x, g[5] = y.y, g[6] = y.z, g[7] = y.w;
    }
    if constexpr (OUT == PSB_BF16) {
      *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(stage_grad) + v * 8) = pack8(g);
    } else {
      float* dst = reinterpret_cast<float*>(stage_grad) + v * 8;
      *reinterpret_cast<float4*>(dst) = make_float4(g[0], g[1], g[2], g[3]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(g[4], g[5], g[6], g[7]);
    }
  }
}

// staged entry t of the flattened (worker, entry) space -> is it a row of this rank?
__device__ __forceinline__ bool staged_entry(const PsbSparse& s, int64_t t, int& w, int64_t& i,
                                             int64_t& lrow) {
  w = (int)(t / s.cap);
  i = t - (int64_t)w * s.cap;
  if (w >= s.fabric.world) return false;
  uint32_t cnt = ld_peer_u32(s.stage_cnt[w]);
  if (cnt > (uint32_t)s.cap) cnt = (uint32_t)s.cap;
  if (i >= (int64_t)cnt) return false;
  int owner;
  const int64_t id = ld_peer_i64(reinterpret_cast<const int64_t*>(s.stage_idx[w]) + i);
  return locate(s, id, owner, lrow) && owner == s.fabric.rank;
}

__global__ void __launch_bounds__(kThreads) sparse_mark_kernel(PsbSparse s) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int w;
  int64_t i, lrow;
  if (!staged_entry(s, t, w, i, lrow)) return;
  // the first arrival claims the row (-1 -> -2), takes the next accumulator slot and publishes
  // it; later arrivals of the same row see -2 or the slot and have nothing to do.  Readers of
  // slot[] run in the next launch.
  if (atomicCAS(&s.slot[lrow], -1, -2) == -1) {
    const uint32_t k = atomicAdd(s.n_touched, 1u);
    s.touched[k] = (int32_t)lrow;
    s.slot[lrow] = (int32_t)k;
  }
}

template <int GDT>
__global__ void __launch_bounds__(kThreads) sparse_accumulate_kernel(PsbSparse s) {
  const int vecs = s.dim / 8;
  const int lpr = lanes_per_row(vecs);
  const int lane = threadIdx.x & 31, l = lane % lpr;
  const int64_t groups_per_warp = 32 / lpr;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t total = (int64_t)s.fabric.world * s.cap;
  for (int64_t t = warp * groups_per_warp + lane / lpr; t < total; t += nwarps * groups_per_warp) {
    int w;
    int64_t i, lrow;
    if (!staged_entry(s, t, w, i, lrow)) continue;
    float* a = s.acc + (int64_t)s.slot[lrow] * s.dim;
    for (int v = l; v < vecs; v += lpr) {
      float g[8];
      if constexpr (GDT == PSB_BF16) {
        unpack8(ld_stream_16(reinterpret_cast<const __nv_bfloat16*>(s.stage_grad[w]) +
                             i * s.dim + v * 8), g);
      } else {
        const float* src = reinterpret_cast<const float*>(s.stage_grad[w]) + i * s.dim + v * 8;
        const float4 x = ld_stream_f4(src), y = ld_stream_f4(src + 4);
        g[0] = x.x, g[1] = x.y, g[2] = x.z, g[3] = x.w, g[4] = y.x, g[5] = y.y, g[6] = y.z, g[7] = y.w;
      }
      atomicAdd(reinterpret_cast<float4*>(a + v * 8), make_float4(g[0], g[1], g[2], g[3]));
      atomicAdd(reinterpret_cast<float4*>(a + v * 8 + 4), make_float4(g[4], g[5], g[6], g[7]));
    }
  }
}

template <int KIND>
__global__ void __launch_bounds__(kThreads) sparse_apply_kernel(PsbSparse s) {
  const OptimCoef c = make_coef(s.opt, s.step ? *s.step : 0u);
  const float gs = s.opt.grad_scale;
  const int vecs = s.dim / 8;
  const int lpr = lanes_per_row(vecs);
  const int lane = threadIdx.x & 31, l = lane % lpr;
  const int64_t groups_per_warp = 32 / lpr;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t n = *s.n_touched;
  __nv_bfloat16* table =
      const_cast<__nv_bfloat16*>(reinterpret_cast<const __nv_bfloat16*>(s.table[s.fabric.rank]));
  for (int64_t k = warp * groups_per_warp + lane / lpr; k < n; k += nwarps * groups_per_warp) {
    const int64_t row = s.touched[k];
    float* a = s.acc + k * s.dim;
    for (int v = l; v < vecs; v += lpr) {
      const int64_t e = row * s.dim + v * 8;
      float g[8], p[8], mm[8] = {}, vv[8] = {};
      {
        const float4 x = *reinterpret_cast<const float4*>(a + v * 8);
        const float4 y = *reinterpret_cast<const float4*>(a + v * 8 + 4);
        g[0] = x.x, g[1] = x.y, g[2] = x.z, g[3] = x.w, g[4] = y.x, g[5] = y.y, g[6] = y.z, g[7] = y.w;
        *reinterpret_cast<float4*>(a + v * 8) = make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4*>(a + v * 8 + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      {
        const float4 x = *reinterpret_cast<const float4*>(s.master + e);
        const float4 y = *reinterpret_cast<const float4*>(s.master + e + 4);
        p[0] = x.x, p[1] = x.y, p[2] = x.z, p[3] = x.w, p[4] = y.x, p[5] = y.y, p[6] = y.z, p[7] = y.w;
      }
      if constexpr (KIND != PSB_OPT_SGD) {
        const float4 x = *reinterpret_cast<const float4*>(s.m + e);
        const float4 y = *reinterpret_cast<const float4*>(s.m + e + 4);
        mm[0] = x.x, mm[1] = x.y, mm[2] = x.z, mm[3] = x.w, mm[4] = y.x, mm[5] = y.y, mm[6] = y.z,
        mm[7] = y.w;
      }
      if constexpr (KIND == PSB_OPT_ADAM || KIND == PSB_OPT_ADAMW) {
        const float4 x = *reinterpret_cast<const float4*>(s.v + e);
        const float4 y = *reinterpret_cast<const float4*>(s.v + e + 4);
        vv[0] = x.x, vv[1] = x.y, vv[2] = x.z, vv[3] = x.w, vv[4] = y.x, vv[5] = y.y, vv[6] = y.z,
        vv[7] = y.w;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) update1<KIND>(c, p[j], mm[j], vv[j], g[j] * gs);
      *reinterpret_cast<float4*>(s.master + e) = make_float4(p[0], p[1], p[2], p[3]);
      *reinterpret_cast<float4*>(s.master + e + 4) = make_float4(p[4], p[5], p[6], p[7]);
      if constexpr (KIND != PSB_OPT_SGD) {
        *reinterpret_cast<float4*>(s.m + e) = make_float4(mm[0], mm[1], mm[2], mm[3]);
        *reinterpret_cast<float4*>(s.m + e + 4) = make_float4(mm[4], mm[5], mm[6], mm[7]);
      }
      if constexpr (KIND == PSB_OPT_ADAM || KIND == PSB_OPT_ADAMW) {
        *reinterpret_cast<float4*>(s.v + e) = make_float4(vv[0], vv[1], vv[2], vv[3]);
        *reinterpret_cast<float4*>(s.v + e + 4) = make_float4(vv[4], vv[5], vv[6], vv[7]);
      }
      *reinterpret_cast<uint4*>(table + e) = pack8(p);
    }
    if (l == 0) s.slot[row] = -1;
  }
}

// CTAs for `groups` sub-warp groups of `lpr` lanes (grid-stride loops take the rest)
int grid_for(int64_t groups, int lpr) {
  const int64_t per_block = (int64_t)(kThreads / 32) * (32 / lpr);
  int64_t g = (groups + per_block - 1) / per_block;
  const int64_t cap = (int64_t)kNumSMs * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// ---- host side: argument checks and launches ----
int validate(const PsbSparse* s, const char* who, bool owner_state) {
  PSB_REQUIRE(s != nullptr, PSB_EINVAL, "%s: null argument", who);
  const int world = s->fabric.world, rank = s->fabric.rank;
  PSB_REQUIRE(world >= 1 && world <= PSB_MAX_PEERS && rank >= 0 && rank < world, PSB_EINVAL,
              "%s: rank %d of world %d", who, rank, world);
  PSB_REQUIRE(s->dim > 0 && s->dim % 8 == 0, PSB_EINVAL,
              "%s: dim %d is not a positive multiple of 8 (16-byte bf16 vectors)", who, s->dim);
  PSB_REQUIRE(s->vocab > 0 && s->rows_per_rank > 0 &&
                  s->rows_per_rank * (int64_t)world >= s->vocab,
              PSB_EINVAL, "%s: %lld rows per rank x %d ranks do not cover %lld rows", who,
              (long long)s->rows_per_rank, world, (long long)s->vocab);
  for (int r = 0; r < world; ++r)
    PSB_REQUIRE(s->table[r] != nullptr, PSB_EINVAL, "%s: table shard of rank %d is missing", who, r);
  if (!owner_state) return 0;
  PSB_REQUIRE(s->cap > 0, PSB_EINVAL, "%s: staging capacity %d", who, s->cap);
  PSB_REQUIRE(s->grad_dtype == PSB_F32 || s->grad_dtype == PSB_BF16, PSB_EINVAL,
              "%s: gradient dtype %d", who, s->grad_dtype);
  for (int r = 0; r < world; ++r)
    PSB_REQUIRE(s->stage_idx[r] && s->stage_grad[r] && s->stage_cnt[r], PSB_EINVAL,
                "%s: staging range of rank %d is missing", who, r);
  PSB_REQUIRE(s->master && s->slot && s->touched && s->n_touched && s->acc, PSB_EINVAL,
              "%s: owner state (master / slot / touched / n_touched / acc) is missing", who);
  const int kind = s->opt.kind;
  PSB_REQUIRE(kind >= PSB_OPT_SGD && kind <= PSB_OPT_ADAMW, PSB_EINVAL, "%s: optimizer kind %d",
              who, kind);
  PSB_REQUIRE(kind == PSB_OPT_SGD || s->m, PSB_EINVAL, "%s: optimizer needs m", who);
  PSB_REQUIRE((kind != PSB_OPT_ADAM && kind != PSB_OPT_ADAMW) || s->v, PSB_EINVAL,
              "%s: Adam needs v", who);
  return 0;
}

}  // namespace
}  // namespace psb

using namespace psb;

extern "C" int psb_sparse_init(int32_t* slot, int64_t rows, psb_stream_t stream) {
  PSB_REQUIRE(slot != nullptr && rows > 0, PSB_EINVAL, "psb_sparse_init: null slot array / no rows");
  PSB_CUDA_OK(cudaMemsetAsync(slot, 0xFF, (size_t)rows * sizeof(int32_t), (cudaStream_t)stream));
  return 0;
}

extern "C" int psb_sparse_pull(const PsbSparse* s, const int64_t* idx, int64_t n, void* out_bf16,
                               psb_stream_t stream) {
  if (int rc = validate(s, "psb_sparse_pull", false)) return rc;
  if (n == 0) return 0;
  PSB_REQUIRE(idx != nullptr && out_bf16 != nullptr && n > 0, PSB_EINVAL,
              "psb_sparse_pull: null indices / output");
  PSB_REQUIRE((reinterpret_cast<uintptr_t>(out_bf16) & 15u) == 0, PSB_EINVAL,
              "psb_sparse_pull: output is not 16-byte aligned");
  const int vecs = s->dim / 8;
  const int lpr = vecs <= 8 ? 8 : (vecs <= 16 ? 16 : 32);
  sparse_pull_kernel<<<grid_for(n, lpr), kThreads, 0, (cudaStream_t)stream>>>(
      *s, idx, n, reinterpret_cast<__nv_bfloat16*>(out_bf16));
  PSB_LAUNCH_OK("sparse_pull_kernel");
  return 0;
}

extern "C" int psb_sparse_stage(const PsbSparse* s, const int64_t* idx, const void* grad,
                                int grad_dtype, int64_t n, psb_stream_t stream) {
  if (int rc = validate(s, "psb_sparse_stage", true)) return rc;
  PSB_REQUIRE(n >= 0 && n <= s->cap, PSB_ERANGE, "psb_sparse_stage: %lld rows, staging capacity %d",
              (long long)n, s->cap);
  PSB_REQUIRE(n == 0 || (idx != nullptr && grad != nullptr), PSB_EINVAL,
              "psb_sparse_stage: null ids / gradients");
  PSB_REQUIRE(grad_dtype == PSB_F32 || grad_dtype == PSB_BF16, PSB_EINVAL,
              "psb_sparse_stage: gradient dtype %d", grad_dtype);
  PSB_REQUIRE((reinterpret_cast<uintptr_t>(grad) & 15u) == 0, PSB_EINVAL,
              "psb_sparse_stage: gradients are not 16-byte aligned");
  const int r = s->fabric.rank;
  int64_t* sidx = const_cast<int64_t*>(reinterpret_cast<const int64_t*>(s->stage_idx[r]));
  void* sgrad = const_cast<void*>(s->stage_grad[r]);
  uint32_t* scnt = const_cast<uint32_t*>(reinterpret_cast<const uint32_t*>(s->stage_cnt[r]));
  const int64_t work = n * s->dim / 8 > n ? n * s->dim / 8 : n;
  int64_t g = (work + kThreads - 1) / kThreads;
  const int grid = (int)(g < 1 ? 1 : (g > (int64_t)kNumSMs * 8 ? (int64_t)kNumSMs * 8 : g));
  cudaStream_t st = (cudaStream_t)stream;
  const bool in_bf = grad_dtype == PSB_BF16, out_bf = s->grad_dtype == PSB_BF16;
  if (in_bf && out_bf)
    sparse_stage_kernel<PSB_BF16, PSB_BF16><<<grid, kThreads, 0, st>>>(idx, grad, n, s->dim, sidx, sgrad, scnt);
  else if (in_bf)
    sparse_stage_kernel<PSB_BF16, PSB_F32><<<grid, kThreads, 0, st>>>(idx, grad, n, s->dim, sidx, sgrad, scnt);
  else if (out_bf)
    sparse_stage_kernel<PSB_F32, PSB_BF16><<<grid, kThreads, 0, st>>>(idx, grad, n, s->dim, sidx, sgrad, scnt);
  else
    sparse_stage_kernel<PSB_F32, PSB_F32><<<grid, kThreads, 0, st>>>(idx, grad, n, s->dim, sidx, sgrad, scnt);
  PSB_LAUNCH_OK("sparse_stage_kernel");
  return 0;
}

extern "C" int psb_sparse_apply(const PsbSparse* s, psb_stream_t stream) {
  if (int rc = validate(s, "psb_sparse_apply", true)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t total = (int64_t)s->fabric.world * s->cap;
  const int vecs = s->dim / 8;
  const int lpr = vecs <= 8 ? 8 : (vecs <= 16 ? 16 : 32);
  sparse_mark_kernel<<<(unsigned)((total + kThreads - 1) / kThreads), kThreads, 0, st>>>(*s);
  PSB_LAUNCH_OK("sparse_mark_kernel");
  if (s->grad_dtype == PSB_BF16)
    sparse_accumulate_kernel<PSB_BF16><<<grid_for(total, lpr), kThreads, 0, st>>>(*s);
  else
    sparse_accumulate_kernel<PSB_F32><<<grid_for(total, lpr), kThreads, 0, st>>>(*s);
  PSB_LAUNCH_OK("sparse_accumulate_kernel");
  const int g = grid_for(total, lpr);
  switch (s->opt.kind) {
    case PSB_OPT_SGD: sparse_apply_kernel<PSB_OPT_SGD><<<g, kThreads, 0, st>>>(*s); break;
    case PSB_OPT_MOMENTUM: sparse_apply_kernel<PSB_OPT_MOMENTUM><<<g, kThreads, 0, st>>>(*s); break;
    case PSB_OPT_ADAM: sparse_apply_kernel<PSB_OPT_ADAM><<<g, kThreads, 0, st>>>(*s); break;
    default: sparse_apply_kernel<PSB_OPT_ADAMW><<<g, kThreads, 0, st>>>(*s); break;
  }
  PSB_LAUNCH_OK("sparse_apply_kernel");
  PSB_CUDA_OK(cudaMemsetAsync(s->n_touched, 0, sizeof(uint32_t), st));
  return 0;
}
