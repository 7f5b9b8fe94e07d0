// gemm_tcgen05.cu — K2f: bf16 GEMM on the 5th-generation tensor cores, optionally fused with
// the all-gather of its row-sharded weight over NVLink peer memory.
//
//   Y[M,N] = X[M,K] · W[N,K]^T (+ bias),  fp32 accumulation in TMEM.
//
// Structure (one persistent CTA per SM, 192 threads, warp-specialised):
//   warp 0      TMA producer: cp.async.bulk.tensor 2-D tiles (128B swizzle) of X and W into a
//               4-stage shared-memory ring, completion on mbarriers
//   warp 1      MMA issuer: one elected lane issues tcgen05.mma (128x256x16, kind::f16) from
//               shared-memory descriptors into one of two 256-column TMEM accumulators and
//               releases ring slots / publishes accumulators with tcgen05.commit
//   warps 2..5  epilogue: tcgen05.ld 32x32b.x32 from TMEM, bias, bf16/fp32 conversion, global
//               stores; overlaps the next tile's main loop through the second accumulator
// Fused variant (psb_gemm_allgather_bf16): the first `gather_ctas` CTAs do no math.  Each
// runs a ring of 1-D bulk copies (global(peer) -> smem -> global(local)) that pulls the
// remote row-slices of W from their owners' replicas into the local replica and bumps a
// per-N-tile arrival counter; GEMM CTAs start with the locally owned N-tiles and acquire
// the counter (plus fence.proxy.async) before their TMA loads touch a gathered tile, so
// the transfer of slice r+1 overlaps the math on slice r.
//
// No counterpart in the reference: its only "pull" is ServeParameters' deep copy
// (src/parameter_server.cpp:93-97) and it has no GEMM at all.
#include <cuda.h>

#include <cstdlib>
#include <mutex>

#include "common.cuh"

namespace psb {
namespace gemm {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;  // 64 bf16 = 128 B = one swizzle atom row
constexpr int UMMA_K = 16;
constexpr int STAGES = 4;
constexpr int A_BYTES = BM * BK * 2;  // 16 KB
constexpr int B_BYTES = BN * BK * 2;  // 32 KB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int THREADS = 192;
constexpr int TMEM_COLS = 512;  // two fp32 accumulators of BN columns
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;

constexpr int G_CHUNK = 32768;  // gather ring: bytes per bulk copy
constexpr int G_STAGES = 6;

// ---- PTX wrappers ------------------------------------------------------------------
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* smem,
                                            int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void bulk_commit() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(dst_smem)),
               "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols)
               : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once every previously issued tcgen05.mma of this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}

// Shared-memory matrix descriptor of a K-major tile stored as rows of 64 bf16 (128 B) with
// the 128-byte swizzle TMA applied: 8-row groups are 1024 B apart (SBO), LBO is unused for
// swizzled K-major layouts, version = 1 (Blackwell), layout type 2 = SWIZZLE_128B.
__device__ __forceinline__ uint64_t make_smem_desc(const void* tile) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_u32(tile) & 0x3ffff) >> 4);
  d |= (uint64_t)1 << 16;            // leading byte offset (ignored), in 16 B units
  d |= (uint64_t)(1024 >> 4) << 32;  // stride byte offset
  d |= (uint64_t)1 << 46;            // descriptor version
  d |= (uint64_t)2 << 61;            // SWIZZLE_128B
  return d;
}
// kind::f16 instruction descriptor: D = fp32, A = B = bf16, both K-major, M = 128, N = 256
constexpr uint32_t kInstrDesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) |
                                ((uint32_t)(BM >> 4) << 24);

struct Params {
  void* y;
  const float* bias;
  int M, N, K;
  int out_dtype;
  // fused all-gather
  int fused;
  int gather_ctas;
  int rank, world;
  const char* w_peer[PSB_MAX_PEERS];  // peer-mapped replicas of W
  char* w_local;
  uint32_t* tile_flags;  // [ceil(N/BN)] arrival counters, zeroed before launch
  const uint32_t* version_pad;  // local pad + PSB_PAD_VERSION_OFF + bucket * PSB_MAX_PEERS
  const uint32_t* epoch;
  int skip_wait;
  // CTA-pair kernel only: fused epilogue of the all-gather entry (PSB_EPI_*), Z output, bias type
  int epilogue;
  void* aux_out;
  int bias_bf16;
};

struct TileOrder {
  int num_m, num_n, n_start;
  __device__ __forceinline__ void decode(int t, int& m_blk, int& n_blk) const {
    const int n_ord = t / num_m;
    m_blk = t - n_ord * num_m;
    n_blk = n_start + n_ord;
    if (n_blk >= num_n) n_blk -= num_n;
  }
};

__device__ void gather_role(const Params& p, uint8_t* ring, uint64_t* bars, int gidx) {
  // one thread drives the whole ring
  const size_t slice_bytes = (size_t)(p.N / p.world) * p.K * 2;
  const size_t chunks_per_slice = slice_bytes / G_CHUNK;
  const size_t tile_bytes = (size_t)BN * p.K * 2;
  const size_t n_remote = chunks_per_slice * (size_t)(p.world - 1);
  // items of this CTA: gidx, gidx + G, ...
  const size_t n_items = n_remote > (size_t)gidx
                             ? (n_remote - gidx + p.gather_ctas - 1) / p.gather_ctas
                             : 0;
  auto item_offset = [&](size_t it, int& owner) -> size_t {
    const size_t c = gidx + it * p.gather_ctas;
    const int ord = (int)(c / chunks_per_slice);
    owner = (p.rank + 1 + ord) % p.world;
    return (size_t)owner * slice_bytes + (c - (size_t)ord * chunks_per_slice) * G_CHUNK;
  };
  uint32_t waited_owner_mask = p.skip_wait ? 0xffffffffu : 0u;
  const uint32_t target = p.skip_wait ? 0u : *p.epoch;
  size_t prev_off = 0;
  bool have_prev = false;
  for (size_t i = 0; i < n_items + (G_STAGES - 1); ++i) {
    if (i < n_items) {
      const int s = (int)(i % G_STAGES);
      if (i >= G_STAGES) bulk_wait_read<0>();  // store(i - G_STAGES) has drained this buffer
      int owner;
      const size_t off = item_offset(i, owner);
      if (!((waited_owner_mask >> owner) & 1u)) {
        wait_ge_sys(p.version_pad + owner, target);  // owner finished writing its slice
        fence_proxy_async();
        waited_owner_mask |= 1u << owner;
      }
      mbar_expect_tx(&bars[s], G_CHUNK);
      bulk_load(ring + (size_t)s * G_CHUNK, p.w_peer[owner] + off, G_CHUNK, &bars[s]);
    }
    if (i >= G_STAGES - 1) {
      const size_t j = i - (G_STAGES - 1);
      if (j < n_items) {
        const int s = (int)(j % G_STAGES);
        mbar_wait(&bars[s], (uint32_t)((j / G_STAGES) & 1));
        int owner;
        const size_t off = item_offset(j, owner);
        bulk_store(p.w_local + off, ring + (size_t)s * G_CHUNK, G_CHUNK);
        bulk_commit();
        if (have_prev) {
          bulk_wait<1>();  // everything but the store just issued is globally visible
          red_release_gpu_add(p.tile_flags + prev_off / tile_bytes, 1u);
        }
        prev_off = off;
        have_prev = true;
      }
    }
  }
  if (have_prev) {
    bulk_wait<0>();
    red_release_gpu_add(p.tile_flags + prev_off / tile_bytes, 1u);
  }
}

__global__ void __launch_bounds__(THREADS, 1)
    gemm_bf16_kernel(const __grid_constant__ CUtensorMap map_a,
                     const __grid_constant__ CUtensorMap map_b, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_a = smem;                         // STAGES x 16 KB
  uint8_t* smem_b = smem + STAGES * A_BYTES;      // STAGES x 32 KB
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* gbars = tmem_empty + 2;  // G_STAGES barriers of the gather ring
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(gbars + G_STAGES);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const bool is_gather = p.fused && (int)blockIdx.x < p.gather_ctas;

  if (is_gather) {
    if (threadIdx.x == 0) {
      for (int i = 0; i < G_STAGES; ++i) mbar_init(&gbars[i], 1);
      fence_barrier_init();
      fence_proxy_async();
      gather_role(p, smem, gbars, blockIdx.x);
    }
    return;
  }

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 4);
    }
    fence_barrier_init();
    fence_proxy_async();
  }
  if (warp == 1) tmem_alloc(tmem_base_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;

  TileOrder order;
  order.num_m = (p.M + BM - 1) / BM;
  order.num_n = (p.N + BN - 1) / BN;
  order.n_start = p.fused ? p.rank * (order.num_n / p.world) : 0;
  const int num_tiles = order.num_m * order.num_n;
  const int num_k = (p.K + BK - 1) / BK;
  const int first = p.fused ? (int)blockIdx.x - p.gather_ctas : (int)blockIdx.x;
  const int step = p.fused ? (int)gridDim.x - p.gather_ctas : (int)gridDim.x;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const int n_per_rank = p.fused ? order.num_n / p.world : 0;
      const uint32_t chunks_per_tile = p.fused ? (uint32_t)((size_t)BN * p.K * 2 / G_CHUNK) : 0;
      for (int t = first; t < num_tiles; t += step) {
        int m_blk, n_blk;
        order.decode(t, m_blk, n_blk);
        if (p.fused && n_blk / n_per_rank != p.rank) {
          // gathered tile: wait until every chunk of it has landed in the local replica
          SpinWatch watch;
          while (ld_acquire_gpu(p.tile_flags + n_blk) < chunks_per_tile) {
            __nanosleep(64);
            if (watch.tick()) break;
          }
          fence_proxy_async();
        }
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_expect_tx(&full[stage], STAGE_BYTES);
          tma_load_2d(&map_a, &full[stage], smem_a + stage * A_BYTES, kb * BK, m_blk * BM);
          tma_load_2d(&map_b, &full[stage], smem_b + stage * B_BYTES, kb * BK, n_blk * BN);
          if (++stage == STAGES) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = first; t < num_tiles; t += step, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint64_t da = make_smem_desc(smem_a + stage * A_BYTES);
          const uint64_t db = make_smem_desc(smem_b + stage * B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // advance 16 elements = 32 B inside the swizzle atom: +2 in 16-byte units
            umma_bf16(tmem_d, da + 2 * k, db + 2 * k, kInstrDesc, (kb | k) != 0);
          }
          umma_commit(&empty[stage]);
          if (++stage == STAGES) stage = 0, phase ^= 1;
        }
        umma_commit(&tmem_full[acc]);
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int quarter = warp & 3;  // TMEM lane quarter this warp may access
    int it = 0;
    for (int t = first; t < num_tiles; t += step, ++it) {
      int m_blk, n_blk;
      order.decode(t, m_blk, n_blk);
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const int row = m_blk * BM + quarter * 32 + lane;
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(quarter * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld32(taddr + c * 32, r);
        tmem_ld_wait();
        const int col0 = n_blk * BN + c * 32;
        if (row < p.M && col0 < p.N) {
          float f[32];
#pragma unroll
          for (int k = 0; k < 32; ++k) f[k] = __uint_as_float(r[k]);
          if (p.bias) {
#pragma unroll
            for (int k = 0; k < 32; ++k)
              if (col0 + k < p.N) f[k] += __ldg(p.bias + col0 + k);
          }
          if (p.out_dtype == PSB_BF16) {
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.y) + (size_t)row * p.N + col0;
            if (col0 + 32 <= p.N) {
#pragma unroll
              for (int q = 0; q < 4; ++q) st_stream_16(dst + q * 8, pack8(f + q * 8));
            } else {
#pragma unroll
              for (int k = 0; k < 32; ++k)
                if (col0 + k < p.N) dst[k] = __float2bfloat16_rn(f[k]);
            }
          } else {
            float* dst = reinterpret_cast<float*>(p.y) + (size_t)row * p.N + col0;
            if (col0 + 32 <= p.N) {
#pragma unroll
              for (int q = 0; q < 8; ++q)
                st_stream_f4(dst + q * 4, make_float4(f[q * 4], f[q * 4 + 1], f[q * 4 + 2],
                                                      f[q * 4 + 3]));
            } else {
#pragma unroll
              for (int k = 0; k < 32; ++k)
                if (col0 + k < p.N) dst[k] = f[k];
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ---- host side -----------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

// row-major [rows, cols] bf16 matrix, box = [box_rows, 64 cols], 128-byte swizzle
static int make_map(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                    uint32_t box_rows) {
  EncodeTiledFn enc = get_encode();
  PSB_REQUIRE(enc, PSB_EUNSUPPORTED, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, box_rows};
  cuuint32_t elem[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                   box, elem, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PSB_REQUIRE(r == CUDA_SUCCESS, PSB_EINVAL, "cuTensorMapEncodeTiled failed with CUresult %d",
              (int)r);
  return 0;
}

}  // namespace gemm

// gemm_tcgen05_2cta.cu
int gemm2_launch(const void* x, const void* w, void* y, const float* bias, int M, int N, int K,
                 int out_dtype, int fused, int gather_ctas, int rank, int world,
                 const char* const* w_peer, char* w_local, uint32_t* tile_flags,
                 const uint32_t* version_pad, const uint32_t* epoch, int skip_wait, cudaStream_t s,
                 int epilogue, void* aux_out, int bias_bf16);

namespace gemm {

// 0 = one CTA per tile (cta_group::1, 128x256), 1 = CTA pair per tile (cta_group::2, 256x256)
static std::atomic<int> g_variant{-1};
static int variant() {
  int v = g_variant.load();
  if (v < 0) {
    const char* e = getenv("PSB_GEMM_VARIANT");
    v = (e && e[0] == '1') ? 0 : 1;  // PSB_GEMM_VARIANT=1cta selects the single-CTA kernel
    g_variant.store(v);
  }
  return v;
}

static int launch(const void* x, const void* w, Params& p, cudaStream_t s) {
  PSB_REQUIRE(p.M > 0 && p.N > 0 && p.K > 0, PSB_EINVAL, "gemm: empty problem %dx%dx%d", p.M, p.N,
              p.K);
  PSB_REQUIRE(p.K % 8 == 0, PSB_EINVAL, "gemm: K=%d must be a multiple of 8 (16-byte rows)", p.K);
  PSB_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)w & 15) == 0 && ((uintptr_t)p.y & 15) == 0,
              PSB_EINVAL, "gemm: operands must be 16-byte aligned");
  PSB_REQUIRE(p.out_dtype == PSB_BF16 || p.out_dtype == PSB_F32, PSB_EINVAL,
              "gemm: bad output dtype %d", p.out_dtype);
  PSB_REQUIRE((p.out_dtype == PSB_BF16 ? p.N % 8 : p.N % 4) == 0, PSB_EINVAL,
              "gemm: N=%d breaks 16-byte row alignment of the output", p.N);
  if (variant() == 1)
    return gemm2_launch(x, w, p.y, p.bias, p.M, p.N, p.K, p.out_dtype, p.fused, p.gather_ctas,
                        p.rank, p.world, p.w_peer, p.w_local, p.tile_flags, p.version_pad, p.epoch,
                        p.skip_wait, s, p.epilogue, p.aux_out, p.bias_bf16);
  PSB_REQUIRE(p.epilogue == 0 && !p.bias_bf16, PSB_EUNSUPPORTED,
              "gemm: fused epilogues / bf16 bias need the CTA-pair kernel (psb_gemm_select(1))");
  CUtensorMap map_a, map_b;
  if (int rc = make_map(&map_a, x, p.M, p.K, BM)) return rc;
  if (int rc = make_map(&map_b, w, p.N, p.K, BN)) return rc;
  // per launch, not once per process: the attribute belongs to the current device's context
  PSB_CUDA_OK(cudaFuncSetAttribute(gemm_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   SMEM_BYTES));
  const int num_tiles = ((p.M + BM - 1) / BM) * ((p.N + BN - 1) / BN);
  int grid = num_tiles < kNumSMs ? num_tiles : kNumSMs;
  if (p.fused) {
    grid = kNumSMs;
    if (num_tiles + p.gather_ctas < grid) grid = num_tiles + p.gather_ctas;
  }
  gemm_bf16_kernel<<<grid, THREADS, SMEM_BYTES, s>>>(map_a, map_b, p);
  PSB_LAUNCH_OK("gemm_bf16_tcgen05");
  return 0;
}

}  // namespace gemm
}  // namespace psb

using namespace psb;

extern "C" {

int psb_gemm_select(int variant) {
  PSB_REQUIRE(variant == 0 || variant == 1, PSB_EINVAL, "psb_gemm_select: variant must be 0 or 1");
  gemm::g_variant.store(variant);
  return 0;
}

int psb_gemm_bf16_tn(const void* x, const void* w, const float* bias, void* y, int M, int N, int K,
                     int out_dtype, psb_stream_t stream) {
  PSB_REQUIRE(x && w && y, PSB_EINVAL, "psb_gemm_bf16_tn: null operand");
  gemm::Params p{};
  p.y = y;
  p.bias = bias;
  p.M = M, p.N = N, p.K = K;
  p.out_dtype = out_dtype;
  p.fused = 0;
  return gemm::launch(x, w, p, (cudaStream_t)stream);
}

int psb_gemm_allgather_bf16(const PsbGemmAG* a, psb_stream_t stream) {
  PSB_REQUIRE(a && a->x && a->y, PSB_EINVAL, "psb_gemm_allgather_bf16: null operand");
  const PsbFabric& f = a->fabric;
  PSB_REQUIRE(f.world >= 1 && f.world <= PSB_MAX_PEERS && f.rank >= 0 && f.rank < f.world,
              PSB_ERANGE, "psb_gemm_allgather_bf16: bad fabric rank %d world %d", f.rank, f.world);
  for (int i = 0; i < f.world; ++i)
    PSB_REQUIRE(a->w[i], PSB_EINVAL, "psb_gemm_allgather_bf16: replica of peer %d is null", i);
  gemm::Params p{};
  p.y = a->y;
  p.bias = a->bias;
  p.M = a->M, p.N = a->N, p.K = a->K;
  p.out_dtype = a->out_dtype;
  PSB_REQUIRE(a->epilogue == PSB_EPI_NONE || a->epilogue == PSB_EPI_GELU, PSB_EINVAL,
              "psb_gemm_allgather_bf16: epilogue must be PSB_EPI_NONE or PSB_EPI_GELU");
  PSB_REQUIRE(a->bias_dtype == PSB_F32 || a->bias_dtype == PSB_BF16, PSB_EINVAL,
              "psb_gemm_allgather_bf16: bad bias dtype %d", a->bias_dtype);
  p.epilogue = a->epilogue;
  p.aux_out = a->aux_out;
  p.bias_bf16 = a->bias_dtype == PSB_BF16;
  cudaStream_t s = (cudaStream_t)stream;
  if (f.world == 1) {
    p.fused = 0;
    return gemm::launch(a->x, a->w[f.rank], p, s);
  }
  PSB_REQUIRE(a->N % (gemm::BN * f.world) == 0, PSB_EINVAL,
              "psb_gemm_allgather_bf16: N=%d must be a multiple of %d x world so that every "
              "N-tile has one owner",
              a->N, gemm::BN);
  PSB_REQUIRE(a->K % 64 == 0, PSB_EINVAL,
              "psb_gemm_allgather_bf16: K=%d must be a multiple of 64", a->K);
  PSB_REQUIRE(a->skip_wait || a->epoch, PSB_EINVAL, "psb_gemm_allgather_bf16: epoch missing");
  PSB_REQUIRE(a->bucket >= 0 && a->bucket < PSB_MAX_BUCKETS, PSB_ERANGE,
              "psb_gemm_allgather_bf16: bucket %d outside [0, %d)", a->bucket, PSB_MAX_BUCKETS);
  // arrival counters live in a small per-device scratch allocated on first use
  static uint32_t* flags[64] = {nullptr};
  int dev = 0;
  PSB_CUDA_OK(cudaGetDevice(&dev));
  PSB_REQUIRE(dev >= 0 && dev < 64, PSB_ERANGE, "device ordinal %d unsupported", dev);
  const int num_n = a->N / gemm::BN;
  constexpr int kMaxTiles = 4096;
  PSB_REQUIRE(num_n <= kMaxTiles, PSB_ERANGE, "psb_gemm_allgather_bf16: N too large (%d tiles)",
              num_n);
  if (!flags[dev]) PSB_CUDA_OK(cudaMalloc(&flags[dev], kMaxTiles * sizeof(uint32_t)));
  PSB_CUDA_OK(cudaMemsetAsync(flags[dev], 0, num_n * sizeof(uint32_t), s));
  p.fused = 1;
  {
    // SMs for the gather ring vs the tensor cores, from the two sides of the roofline:
    // time of the math at ~1.5 PFLOP/s vs time of the remote slices at ~700 GB/s
    const double t_gemm = 2.0 * a->M * a->N * a->K / 1.5e15;
    const double t_link = (double)(f.world - 1) / f.world * a->N * a->K * 2.0 / 700e9;
    const double r = t_link / t_gemm;
    p.gather_ctas = r < 0.15 ? 16 : r < 0.4 ? 24 : r < 0.8 ? 40 : 64;
  }
  p.rank = f.rank, p.world = f.world;
  for (int i = 0; i < f.world; ++i) p.w_peer[i] = reinterpret_cast<const char*>(a->w[i]);
  p.w_local = reinterpret_cast<char*>(a->w[f.rank]);
  p.tile_flags = flags[dev];
  p.version_pad = f.pad[f.rank] + PSB_PAD_VERSION_OFF + a->bucket * PSB_MAX_PEERS;
  p.epoch = a->epoch;
  p.skip_wait = a->skip_wait;
  return gemm::launch(a->x, a->w[f.rank], p, s);
}

}  // extern "C"

PSB_EXPORT_SPIN_LIMIT(gemm1)
