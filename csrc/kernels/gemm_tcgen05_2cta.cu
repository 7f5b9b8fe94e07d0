// gemm_tcgen05_2cta.cu — K2f GEMM, CTA-pair version: tcgen05.mma.cta_group::2.
//
//   Y[M,N] = X[M,K] · W[N,K]^T (+ bias), bf16 in, fp32 accumulate in TMEM.
//
// Two CTAs of one cluster (the two SMs of a TPC) work on one 256x256 output tile.  Each CTA
// TMA-loads its own 128 rows of X and HALF of the W tile (128 of the 256 rows), so a k-block
// costs 32 KB of L2->smem traffic per SM instead of 48 KB and the ring holds 6 stages instead
// of 4.  One elected thread of the leader CTA issues 256x256x16 UMMAs that read A from both
// CTAs' shared memory and the two B halves; each CTA keeps the accumulator rows of its own
// 128 output rows in its own TMEM (2 x 256 columns, double-buffered) and runs its own
// epilogue.  Barrier protocol (as in CUTLASS 2-SM kernels): both CTAs' TMA loads complete_tx
// on the LEADER's full barrier; tcgen05.commit multicasts to both CTAs' empty / tmem_full
// barriers; the epilogue warps of both CTAs arrive on the leader's tmem_empty barrier.
// The fused all-gather mode is the same as in gemm_tcgen05.cu (gather clusters + tile flags).
#include <cuda.h>

#include <mutex>

#include "common.cuh"

namespace psb {
namespace gemm2 {

constexpr int BM = 256;        // per cluster; 128 per CTA
constexpr int BN = 256;        // per cluster; each CTA loads 128 rows of W
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int STAGES = 6;
constexpr int A_BYTES = 128 * BK * 2;  // 16 KB per CTA
constexpr int B_BYTES = 128 * BK * 2;  // 16 KB per CTA
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int EPI_WARPS = 8;   // two warps per TMEM lane quarter, each takes half of the columns:
                               // halves the drain time of an accumulator, which matters for the
                               // GELU / dGELU epilogues and for short-K tiles
constexpr int THREADS = 32 * (2 + EPI_WARPS);
constexpr int TMEM_COLS = 512;
constexpr int BAR_BYTES = 1024;                 // mbarriers, TMEM slot, CLC ring
constexpr int EPI_BUF = 32 * 64;                // one staged piece: 32 rows x 32 bf16 columns
constexpr int EPI_STAGE_BYTES = EPI_WARPS * 2 * EPI_BUF;  // two pieces in flight per epilogue warp
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + BAR_BYTES + EPI_STAGE_BYTES;
// Gather (fused all-gather mode): pieces of G_CHUNK bytes travel peer -> shared memory -> local
// replica on the bulk-copy engine, G_STAGES in flight per gather CTA (see gather_role).
constexpr int G_CHUNK = 16384;
constexpr int G_STAGES = 11;   // 176 KB of the CTA's dynamic shared memory
constexpr int G_CHUNK_LDG = 8192;   // register-path gather: 8 KB items, one warp each
constexpr int G_WARPS_LDG = THREADS / 32;
constexpr int G_PEND = 96;     // stores whose completion has not been published yet (batch = half)
constexpr int CLC_SLOTS = 4;   // responses in flight: the producer runs up to ~3 tiles ahead
constexpr int CLC_OFF = 256;   // byte offset of the CLC ring inside the 512-byte barrier area

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t cta) {
  asm volatile(
      "{\n\t"
      ".reg .b32 rem;\n\t"
      "mapa.shared::cluster.u32 rem, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [rem];\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(cta)
      : "memory");
}
// ---- cluster launch control: a running cluster cancels the launch of a not-yet-started one and
// takes over its tile.  The 16-byte response lands in BOTH CTAs' shared memory (multicast) and
// completes 16 bytes on the barrier at the same offset in each of them.
__device__ __forceinline__ void mbar_expect_tx_cluster(uint64_t* bar, uint32_t cta, uint32_t bytes) {
  asm volatile(
      "{\n\t"
      ".reg .b32 rem;\n\t"
      "mapa.shared::cluster.u32 rem, %0, %1;\n\t"
      "mbarrier.arrive.expect_tx.release.cluster.shared::cluster.b64 _, [rem], %2;\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(cta), "r"(bytes)
      : "memory");
}
__device__ __forceinline__ void clc_try_cancel(void* response, uint64_t* bar) {
  asm volatile(
      "clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes"
      ".multicast::cluster::all.b128 [%0], [%1];" ::"r"(smem_u32(response)),
      "r"(smem_u32(bar))
      : "memory");
}
// returns the x index of the first CTA of the cancelled cluster, or -1 when nothing was left
__device__ __forceinline__ int clc_read(const void* response) {
  uint32_t x, valid;
  asm volatile(
      "{\n\t"
      ".reg .pred p1;\n\t"
      ".reg .b128 r;\n\t"
      "ld.shared.b128 r, [%2];\n\t"
      "clusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p1, r;\n\t"
      "selp.u32 %1, 1, 0, p1;\n\t"
      "mov.u32 %0, 0;\n\t"
      "@p1 clusterlaunchcontrol.query_cancel.get_first_ctaid.v4.b32.b128 {%0, _, _, _}, r;\n\t"
      "}"
      : "=r"(x), "=r"(valid)
      : "r"(smem_u32(response))
      : "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  return valid ? (int)x : -1;
}
// 2-SM TMA load: data lands in THIS CTA's shared memory, the transaction bytes are counted on
// the leader CTA's barrier (address with the peer bit cleared)
__device__ __forceinline__ void tma_load_2d_2sm(const CUtensorMap* map, uint64_t* bar, void* smem,
                                                int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem)),
      "l"(map), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t* dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(dst_smem)),
               "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols)
               : "memory");
}
__device__ __forceinline__ void umma2_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                           uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// commit to the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma2_commit_mc(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 "
      "[%0], %1;" ::"r"(smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ uint64_t make_smem_desc(const void* tile) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_u32(tile) & 0x3ffff) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;  // SWIZZLE_128B
  return d;
}
// MN-major operand (the reduction index is the OUTER dimension of the row-major source, e.g.
// W[N_red, K_out] as the B operand of dX = dY * W).  One TMA box per 64-element MN atom:
// {64 MN elements = 128 bytes, BK reduction rows}, swizzle 128B, 8 KB; the two atoms of a
// CTA's 128 rows / columns sit 8 KB apart (leading byte offset), groups of 8 reduction rows
// 1 KB apart (stride byte offset).  One UMMA of K = 16 consumes 16 rows = 2 KB.
constexpr int MN_ATOM_BYTES = 64 * BK * 2;
__device__ __forceinline__ uint64_t make_smem_desc_mn(const void* tile) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_u32(tile) & 0x3ffff) >> 4);
  d |= (uint64_t)(MN_ATOM_BYTES >> 4) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;  // SWIZZLE_128B
  return d;
}
constexpr uint32_t kDescStepK = (UMMA_K * 2) >> 4;           // K-major: 32 bytes inside the atom
constexpr uint32_t kDescStepMN = (UMMA_K * 128) >> 4;        // MN-major: 16 rows of 128 bytes
// D = fp32, A = B = bf16, K-major, M = 256 (pair), N = 256; bit 15 / 16 = A / B is MN-major
constexpr uint32_t kInstrDesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) |
                                ((uint32_t)(BM >> 4) << 24);

struct Params {
  void* y;
  const float* bias;
  int M, N, K;
  int out_dtype;
  int fused, gather_ctas, rank, world;
  int gather_tma;                // gather CTAs use the TMA ring (else 16-byte peer loads)
  int group_m;                   // row blocks per tile-walk group (L2 reuse of the A panels)
  int tma_store;                 // bf16 outputs leave through shared memory + TMA stores (map_y / map_z)
  const char* w_peer[PSB_MAX_PEERS];
  char* w_local;
  uint32_t* tile_flags;
  const uint32_t* version_pad;
  const uint32_t* epoch;
  int skip_wait;
  int bias_bf16;                 // `bias` points at bf16 values
  int clc;                       // tiles come from cluster launch control (grid = one pair per tile)
  int a_mn, b_mn;                // operand is MN-major (stored [K, M] / [K, N] row-major)
  const __nv_bfloat16* aux_in;   // EPI_DGELU: pre-activation Z [M, N]
  __nv_bfloat16* aux_out;        // EPI_GELU: where Z goes (optional)
};

enum { EPI_NONE = 0, EPI_GELU = 1, EPI_DGELU = 2 };

// tanh-form GELU, as torch.nn.functional.gelu(approximate="tanh")
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float gelu_tanh(float z) {
  const float u = 0.7978845608028654f * (z + 0.044715f * z * z * z);
  return 0.5f * z * (1.f + tanh_approx(u));
}
__device__ __forceinline__ float dgelu_tanh(float z) {
  const float z2 = z * z;
  const float t = tanh_approx(0.7978845608028654f * (z + 0.044715f * z * z2));
  return 0.5f * (1.f + t) + 0.5f * z * (1.f - t * t) * 0.7978845608028654f * (1.f + 0.134145f * z2);
}

// Gather role: a CTA that does no math streams remote tiles of W into the local replica through
// a shared-memory ring fed by TMA — peer HBM -> (NVLink) -> smem -> local HBM, no registers
// involved.  A piece is one TMA box of the weight's tensor map, 128 rows x 64 columns = 16 KB
// (the same box the math clusters load): one thread issues cp.async.bulk.tensor loads on the
// OWNER's tensor map (G_STAGES boxes in flight per CTA cover the ~4 us peer latency; measured on
// B200, benchmarks/probe_fabric.cu: 16 CTAs reach 710 GB/s from a single peer, 32 CTAs 755 GB/s —
// 1-D cp.async.bulk copies of the same size stay at ~1 GB/s per copy in flight), a second thread
// writes each box home with a TMA store on the local map and publishes completed boxes by bumping
// the arrival counter of their 256-row N-tile (a tile is complete at 2 * K/64 boxes).  Boxes are
// dealt round-robin over the gather CTAs in owner order, so tiles complete in the order in which
// the math clusters consume them.
struct PeerMaps {
  CUtensorMap m[PSB_MAX_PEERS];
};

__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* smem, int c0,
                                            int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::
          "r"(smem_u32(smem)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* smem, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map),
               "r"(smem_u32(smem)), "r"(c0), "r"(c1)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

__device__ void gather_role(const Params& p, int gidx, uint8_t* smem, const PeerMaps& peers,
                            const CUtensorMap* local_map) {
  uint8_t* ring = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(ring + G_STAGES * G_CHUNK);
  uint64_t* empty = full + G_STAGES;
  if (threadIdx.x == 0) {
    for (int i = 0; i < G_STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
  }
  __syncthreads();
  const int rows_per_slice = p.N / p.world;
  const uint32_t boxes_per_tile = 2u * (uint32_t)(p.K / BK);  // two 128-row halves per k-block
  const size_t boxes_per_slice = (size_t)(rows_per_slice / BN) * boxes_per_tile;
  const size_t n_remote = boxes_per_slice * (size_t)(p.world - 1);
  // box c -> owner, tile, coordinates
  auto locate = [&](size_t c, int& owner, int& row0, int& col0) -> uint32_t {
    const int ord = (int)(c / boxes_per_slice);
    owner = (p.rank + 1 + ord) % p.world;
    const size_t rem = c - (size_t)ord * boxes_per_slice;
    const int tile_in_slice = (int)(rem / boxes_per_tile);
    const uint32_t r2 = (uint32_t)(rem - (size_t)tile_in_slice * boxes_per_tile);
    row0 = owner * rows_per_slice + tile_in_slice * BN + (int)(r2 & 1u) * 128;
    col0 = (int)(r2 >> 1) * BK;
    return (uint32_t)(row0 / BN);
  };
  if (threadIdx.x == 0) {
    // ---- producer: peer -> ring
    uint32_t waited = p.skip_wait ? 0xffffffffu : 0u;
    const uint32_t target = p.skip_wait ? 0u : *p.epoch;
    int stage = 0;
    uint32_t phase = 0;
    for (size_t c = gidx; c < n_remote; c += p.gather_ctas) {
      int owner, row0, col0;
      locate(c, owner, row0, col0);
      if (!((waited >> owner) & 1u)) {
        wait_ge_sys(p.version_pad + owner, target);  // the owner's slice is final
        fence_proxy_async();                          // ... also for the TMA unit
        waited |= 1u << owner;
      }
      mbar_wait(&empty[stage], phase ^ 1);
      mbar_expect_tx(&full[stage], G_CHUNK);
      tma_load_2d(&peers.m[owner], &full[stage], ring + (size_t)stage * G_CHUNK, col0, row0);
      if (++stage == G_STAGES) stage = 0, phase ^= 1;
    }
  } else if (threadIdx.x == 32) {
    // ---- consumer: ring -> local replica, then publish
    int stage = 0;
    uint32_t phase = 0;
    // Publishing costs a proxy fence + a membar: far more than the 0.36 us a 16 KB box may
    // take at the target rate, so completions are published in batches — when G_PEND stores are
    // unpublished, wait until all but the newest G_PEND/2 have completed
    // (cp.async.bulk.wait_group), fence once, and bump the counters of that older half (boxes of
    // one tile merged into one red).
    uint32_t tile_of[G_PEND];
    int pending = 0;
    auto publish = [&](int n) {  // the n oldest
      fence_proxy_async();       // the stores were made by the async proxy
      __threadfence();
      int i = 0;
      while (i < n) {
        int j = i + 1;
        while (j < n && tile_of[j] == tile_of[i]) ++j;
        red_release_gpu_add(p.tile_flags + tile_of[i], (uint32_t)(j - i));
        i = j;
      }
      for (int k = n; k < pending; ++k) tile_of[k - n] = tile_of[k];
      pending -= n;
    };
    for (size_t c = gidx; c < n_remote; c += p.gather_ctas) {
      int owner, row0, col0;
      const uint32_t tile = locate(c, owner, row0, col0);
      mbar_wait(&full[stage], phase);
      tma_store_2d(local_map, ring + (size_t)stage * G_CHUNK, col0, row0);
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // smem of the slot is free
      mbar_arrive(&empty[stage]);
      if (++stage == G_STAGES) stage = 0, phase ^= 1;
      tile_of[pending++] = tile;
      if (pending == G_PEND) {
        asm volatile("cp.async.bulk.wait_group %0;" ::"n"(G_PEND / 2) : "memory");
        publish(G_PEND / 2);
      }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    if (pending) publish(pending);
  }
}

// Register-path gather (the round-1 mechanism, still the faster one from 4 GPUs up, where the
// remote tiles come from several owners at once — fused_gemm_{4,8}gpu.jsonl): items (8 KB pieces of the remote slices, in owner order) are dealt round-robin to all gather
// warps of the grid.  After its stores a warp makes them visible (membar + release) and bumps
// the arrival counter of the piece's N-tile; a tile is complete at tile_bytes / G_CHUNK_LDG.
__device__ void gather_role_ldg(const Params& p, int gidx) {
  const size_t slice_bytes = (size_t)(p.N / p.world) * p.K * 2;
  const size_t chunks_per_slice = slice_bytes / G_CHUNK_LDG;
  const size_t tile_bytes = (size_t)BN * p.K * 2;
  const size_t n_remote = chunks_per_slice * (size_t)(p.world - 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const size_t worker = (size_t)gidx * G_WARPS_LDG + warp;
  const size_t workers = (size_t)p.gather_ctas * G_WARPS_LDG;
  uint32_t waited = p.skip_wait ? 0xffffffffu : 0u;
  const uint32_t target = p.skip_wait ? 0u : *p.epoch;
  constexpr int V = G_CHUNK_LDG / 512;
  auto locate = [&](size_t c) -> size_t {  // byte offset of item c; waits for its owner once
    const int ord = (int)(c / chunks_per_slice);
    const int owner = (p.rank + 1 + ord) % p.world;
    if (!((waited >> owner) & 1u)) {
      if (lane == 0) wait_ge_sys(p.version_pad + owner, target);  // owner's slice is final
      __syncwarp();
      waited |= 1u << owner;
    }
    return (size_t)owner * slice_bytes + (c - (size_t)ord * chunks_per_slice) * G_CHUNK_LDG;
  };
  auto load = [&](size_t off, uint4 (&v)[V]) {
    const char* src = p.w_peer[off / slice_bytes] + off + lane * 16;
#pragma unroll
    for (int k = 0; k < V; ++k) v[k] = ld_stream_16(src + k * 512);
  };
  auto store = [&](size_t off, const uint4 (&v)[V]) {
    char* dst = p.w_local + off + lane * 16;
#pragma unroll
    for (int k = 0; k < V; ++k) st_stream_16(dst + k * 512, v[k]);
    __threadfence();
    __syncwarp();
    if (lane == 0) red_release_gpu_add(p.tile_flags + off / tile_bytes, 1u);
  };
  // two items in flight per warp: the loads of the next one are issued before the stores and
  // the fence of the current one
  uint4 va[V], vb[V];
  size_t c = worker;
  if (c >= n_remote) return;
  size_t off_a = locate(c), off_b = 0;
  load(off_a, va);
  while (true) {
    const bool more_b = c + workers < n_remote;
    if (more_b) {
      off_b = locate(c + workers);
      load(off_b, vb);
    }
    store(off_a, va);
    if (!more_b) break;
    c += workers;
    const bool more_a = c + workers < n_remote;
    if (more_a) {
      off_a = locate(c + workers);
      load(off_a, va);
    }
    store(off_b, vb);
    if (!more_a) break;
    c += workers;
  }
}

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
    gemm2_bf16_kernel(const __grid_constant__ CUtensorMap map_a,
                      const __grid_constant__ CUtensorMap map_b,
                      const __grid_constant__ PeerMaps peers,
                      const __grid_constant__ CUtensorMap map_y,
                      const __grid_constant__ CUtensorMap map_z, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  uint8_t* clc_resp = smem + STAGES * STAGE_BYTES + CLC_OFF;  // CLC_SLOTS x 16 bytes
  uint8_t* epi_stage = smem + STAGES * STAGE_BYTES + BAR_BYTES;  // 1024-aligned
  uint64_t* clc_full = reinterpret_cast<uint64_t*>(clc_resp + CLC_SLOTS * 16);
  uint64_t* clc_empty = clc_full + CLC_SLOTS;
  uint64_t* zbar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES + 512);  // one per epilogue warp

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();  // 0 = leader
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int gather_clusters = p.fused ? p.gather_ctas >> 1 : 0;

  if (cluster_id < gather_clusters) {  // both CTAs of a gather cluster copy, independently
    if (p.gather_tma) gather_role(p, blockIdx.x, smem, peers, &map_b);
    else gather_role_ldg(p, blockIdx.x);
    return;
  }

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 2);   // leader's expect_tx arrive + the peer producer's arrive
      mbar_init(&empty[i], 1);  // one multicast commit
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);   // one multicast commit
      mbar_init(&tmem_empty[i], 2 * EPI_WARPS);  // epilogue warps of both CTAs arrive on the leader's copy
    }
    for (int i = 0; i < CLC_SLOTS; ++i) {
      mbar_init(&clc_full[i], 1);    // the issuer's expect_tx; the response completes the bytes
      mbar_init(&clc_empty[i], 2 + 2 * EPI_WARPS);  // readers besides the issuer: MMA, peer producer, epilogue warps
    }
    for (int i = 0; i < EPI_WARPS; ++i) mbar_init(&zbar[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
  }
  cluster_sync();  // barriers of both CTAs exist before anybody signals across the pair
  if (warp == 1) tmem_alloc2(tmem_base_slot, TMEM_COLS);
  tc_fence_before();
  cluster_sync();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;

  const int num_m = (p.M + BM - 1) / BM;
  const int num_n = (p.N + BN - 1) / BN;
  const int n_start = p.fused ? p.rank * (num_n / p.world) : 0;
  const int num_tiles = num_m * num_n;
  const int num_k = (p.K + BK - 1) / BK;
  const int first = cluster_id - gather_clusters;
  const int step = num_clusters - gather_clusters;
  // Tile walk.  Static: t = first, first + step, ...  CLC (plain GEMM): the grid has one cluster
  // per tile, a cluster starts on its own tile and then takes over the tiles of clusters whose
  // launch it cancels, so the kernel neither waits for SMs that another stream's kernel holds
  // nor leaves a tail when some clusters start late.
  auto clc_next = [&](int slot, uint32_t phase) -> int {
    mbar_wait(&clc_full[slot], phase);
    const int x = clc_read(clc_resp + slot * 16);
    return x < 0 ? num_tiles : (x >> 1);
  };
  // Tile order: regions of N-tiles by owner (the locally owned slice first, so the gather of the
  // next owner's slice overlaps the math on this one); inside a region tiles are walked in
  // groups of kGroupM row-blocks so that one wave of 74 clusters touches ~8 A panels and ~9 W
  // panels (L2-resident) instead of streaming all of X once per N-tile.
  const int kGroupM = p.group_m;
  const int region_n = p.fused ? num_n / p.world : num_n;
  const int tiles_per_region = num_m * region_n;
  auto decode = [&](int t, int& m_blk, int& n_blk) {
    const int region = t / tiles_per_region;
    const int local = t - region * tiles_per_region;
    const int per_group = kGroupM * region_n;
    const int g = local / per_group;
    const int first_m = g * kGroupM;
    const int gm = num_m - first_m < kGroupM ? num_m - first_m : kGroupM;
    const int in_group = local - g * per_group;
    m_blk = first_m + in_group % gm;
    n_blk = n_start + region * region_n + in_group / gm;
    if (n_blk >= num_n) n_blk -= num_n;
  };

  if (warp == 0) {
    // ===================== TMA producer (one per CTA) =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const int n_per_rank = p.fused ? num_n / p.world : 0;
      const uint32_t chunks_per_tile =  // arrivals that complete a tile: TMA boxes or 8 KB items
          !p.fused ? 0u : p.gather_tma ? 2u * (uint32_t)(p.K / BK)
                                       : (uint32_t)((size_t)BN * p.K * 2 / G_CHUNK_LDG);
      int cs = 0, issued = 0;
      uint32_t cph = 0;
      auto clc_issue = [&](int q) {  // leader only: query number q goes to slot q % CLC_SLOTS
        const int slot = q % CLC_SLOTS;
        mbar_wait(&clc_empty[slot], ((uint32_t)(q / CLC_SLOTS) & 1u) ^ 1u);
        mbar_expect_tx(&clc_full[slot], 16);
        mbar_expect_tx_cluster(&clc_full[slot], 1, 16);
        clc_try_cancel(clc_resp + slot * 16, &clc_full[slot]);
      };
      if (p.clc && cta == 0) clc_issue(0);
      for (int t = first; t < num_tiles;) {
        int t_next = t + step;
        if (p.clc) {
          // the answer asked for one tile ago; a new query only after a successful one
          t_next = clc_next(cs, cph);
          if (cta != 0) mbar_arrive_cluster(&clc_empty[cs], 0);
          if (++cs == CLC_SLOTS) cs = 0, cph ^= 1;
          if (cta == 0 && t_next < num_tiles) clc_issue(++issued);
        }
        int m_blk, n_blk;
        decode(t, m_blk, n_blk);
        if (p.fused && n_blk / n_per_rank != p.rank) {
          SpinWatch watch;  // the gather behind this flag waits on peers: same watchdog
          while (ld_acquire_gpu(p.tile_flags + n_blk) < chunks_per_tile) {
            __nanosleep(64);
            if (watch.tick()) break;
          }
          fence_proxy_async();
        }
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          if (cta == 0) mbar_expect_tx(&full[stage], 2 * STAGE_BYTES);  // both CTAs' bytes
          else mbar_arrive_cluster(&full[stage], 0);
          uint8_t* sa = smem_a + stage * A_BYTES;
          uint8_t* sb = smem_b + stage * B_BYTES;
          const int m0 = m_blk * BM + (int)cta * 128, n0 = n_blk * BN + (int)cta * 128;
          if (!p.a_mn) {
            tma_load_2d_2sm(&map_a, &full[stage], sa, kb * BK, m0);
          } else {
            tma_load_2d_2sm(&map_a, &full[stage], sa, m0, kb * BK);
            tma_load_2d_2sm(&map_a, &full[stage], sa + MN_ATOM_BYTES, m0 + 64, kb * BK);
          }
          if (!p.b_mn) {
            tma_load_2d_2sm(&map_b, &full[stage], sb, kb * BK, n0);
          } else {
            tma_load_2d_2sm(&map_b, &full[stage], sb, n0, kb * BK);
            tma_load_2d_2sm(&map_b, &full[stage], sb + MN_ATOM_BYTES, n0 + 64, kb * BK);
          }
          if (++stage == STAGES) stage = 0, phase ^= 1;
        }
        t = t_next;
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (cta == 0 && lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      const uint32_t idesc = kInstrDesc | (p.a_mn ? 1u << 15 : 0u) | (p.b_mn ? 1u << 16 : 0u);
      const uint32_t step_a = p.a_mn ? kDescStepMN : kDescStepK;
      const uint32_t step_b = p.b_mn ? kDescStepMN : kDescStepK;
      int cs = 0;
      uint32_t cph = 0;
      for (int t = first; t < num_tiles; ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint64_t da = p.a_mn ? make_smem_desc_mn(smem_a + stage * A_BYTES)
                                     : make_smem_desc(smem_a + stage * A_BYTES);
          const uint64_t db = p.b_mn ? make_smem_desc_mn(smem_b + stage * B_BYTES)
                                     : make_smem_desc(smem_b + stage * B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            umma2_bf16(tmem_d, da + step_a * k, db + step_b * k, idesc, (kb | k) != 0);
          umma2_commit_mc(&empty[stage]);  // frees the slot in both CTAs
          if (++stage == STAGES) stage = 0, phase ^= 1;
        }
        umma2_commit_mc(&tmem_full[acc]);  // accumulator ready in both CTAs
        if (p.clc) {
          t = clc_next(cs, cph);
          mbar_arrive_cluster(&clc_empty[cs], 0);
          if (++cs == CLC_SLOTS) cs = 0, cph ^= 1;
        } else {
          t += step;
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2.. of each CTA) =====================
    const int quarter = warp & 3;            // TMEM lanes this warp may read
    const int half = (warp - 2) / 4;         // which half of the tile's columns it drains
    constexpr int kChunks = BN / 32 / (EPI_WARPS / 4);
    int it = 0;
    int cs = 0;
    uint32_t cph = 0;
    int kbuf = 0;  // staged pieces issued by this warp (buffer = kbuf & 1)
    uint32_t zphase = 0;  // parity of the warp's Z barrier (dGELU)
    for (int t = first; t < num_tiles; ++it) {
      int m_blk, n_blk;
      decode(t, m_blk, n_blk);
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      const int row0w = m_blk * BM + (int)cta * 128 + quarter * 32;  // first row of this warp
      const int row = row0w + lane;
      const bool row_ok = row < p.M;
      // dGELU: the pre-activation Z of a piece does not depend on the accumulator.  With the
      // staged epilogue it arrives by TMA (the aux_in tensor map rides in the map_z slot): the
      // first piece's load is issued before the wait for the MMA, each next one as soon as the
      // lanes have taken the current piece out of the warp's Z buffer — a DRAM round trip hidden
      // behind the math and the store of the piece before.  (Register prefetch with LDG is not
      // an option: the proxy fence of the store path waits for every outstanding load.)
      uint8_t* const zbuf = epi_stage + ((warp - 2) * 2) * EPI_BUF;        // dGELU: Z in
      uint8_t* const obuf = epi_stage + ((warp - 2) * 2 + 1) * EPI_BUF;    // dGELU: result out
      const bool z_tma = EPI == EPI_DGELU && p.tma_store;
      auto z_issue = [&](int c) {
        if (lane == 0) {
          mbar_expect_tx(&zbar[warp - 2], EPI_BUF);
          tma_load_2d(&map_z, &zbar[warp - 2], zbuf, n_blk * BN + c * 32, row0w);
        }
      };
      const bool tile_rows = row0w < p.M;
      if (z_tma && tile_rows && n_blk * BN + half * kChunks * 32 < p.N) z_issue(half * kChunks);
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(quarter * 32) << 16);
#pragma unroll 1
      for (int c = half * kChunks; c < (half + 1) * kChunks; ++c) {
        uint32_t r[32];
        tmem_ld32(taddr + c * 32, r);
        tmem_ld_wait();
        const int col0 = n_blk * BN + c * 32;
        // bf16 output through shared memory: the lane's 32 values (one row) go into a 64-byte row
        // of a 2 KB piece (16-byte chunks XOR-swizzled with the row, the TMA 64B pattern, so the
        // eight lanes of a store phase hit distinct banks), then ONE TMA store writes the 32 x 32
        // piece; rows / columns beyond M / N are clipped by the tensor map.  Direct stores from
        // registers cost 32 different cache lines per warp instruction.
        auto stage_store = [&](const CUtensorMap* map, const float* v) {
          uint8_t* buf = z_tma ? obuf : epi_stage + ((warp - 2) * 2 + (kbuf & 1)) * EPI_BUF;
          ++kbuf;
          if (lane == 0) {
            if (z_tma) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            else asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          }
          __syncwarp();  // the store that used this buffer before has read it
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint4 pk = pack8(v + q * 8);
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(
                             smem_u32(buf + lane * 64 + ((q ^ ((lane >> 1) & 3)) << 4))),
                         "r"(pk.x), "r"(pk.y), "r"(pk.z), "r"(pk.w)
                         : "memory");
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) tma_store_2d(map, buf, col0, row0w);
        };
        const bool active = p.tma_store ? (row0w < p.M && col0 < p.N) : (row_ok && col0 < p.N);
        if (active) {
          float f[32];
#pragma unroll
          for (int k = 0; k < 32; ++k) f[k] = __uint_as_float(r[k]);
          if (p.bias) {
            if (p.bias_bf16) {
              const __nv_bfloat16* bb = reinterpret_cast<const __nv_bfloat16*>(p.bias);
#pragma unroll
              for (int k = 0; k < 32; ++k)
                if (col0 + k < p.N) f[k] += __bfloat162float(bb[col0 + k]);
            } else {
#pragma unroll
              for (int k = 0; k < 32; ++k)
                if (col0 + k < p.N) f[k] += __ldg(p.bias + col0 + k);
            }
          }
          const bool full_cols = col0 + 32 <= p.N;
          if constexpr (EPI == EPI_GELU) {
            // Z is rounded to bf16 first: that is the value the backward pass will see
#pragma unroll
            for (int k = 0; k < 32; ++k) f[k] = __bfloat162float(__float2bfloat16_rn(f[k]));
            if (p.aux_out) {
              if (p.tma_store) {
                stage_store(&map_z, f);
              } else {
                __nv_bfloat16* z = p.aux_out + (size_t)row * p.N + col0;
                if (full_cols) {
#pragma unroll
                  for (int q = 0; q < 4; ++q) st_stream_16(z + q * 8, pack8(f + q * 8));
                } else {
#pragma unroll
                  for (int k = 0; k < 32; ++k)
                    if (col0 + k < p.N) z[k] = __float2bfloat16_rn(f[k]);
                }
              }
            }
#pragma unroll
            for (int k = 0; k < 32; ++k) f[k] = gelu_tanh(f[k]);
          }
          if constexpr (EPI == EPI_DGELU) {
            if (z_tma) {
              // rows / columns beyond the matrix arrive as zeros and leave clipped
              mbar_wait(&zbar[warp - 2], zphase);
              zphase ^= 1u;
              uint4 zq[4];
#pragma unroll
              for (int q = 0; q < 4; ++q)
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                             : "=r"(zq[q].x), "=r"(zq[q].y), "=r"(zq[q].z), "=r"(zq[q].w)
                             : "r"(smem_u32(zbuf + lane * 64 + ((q ^ ((lane >> 1) & 3)) << 4))));
              __syncwarp();  // every lane has its row: the buffer may be refilled
              const int cn = c + 1;
              if (cn < (half + 1) * kChunks && n_blk * BN + cn * 32 < p.N) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                z_issue(cn);
              }
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                float zz[8];
                unpack8(zq[q], zz);
#pragma unroll
                for (int k = 0; k < 8; ++k) f[q * 8 + k] *= dgelu_tanh(zz[k]);
              }
            } else if (row_ok) {
              const __nv_bfloat16* z = p.aux_in + (size_t)row * p.N + col0;
              if (full_cols) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  float zz[8];
                  unpack8(ld_stream_16(z + q * 8), zz);
#pragma unroll
                  for (int k = 0; k < 8; ++k) f[q * 8 + k] *= dgelu_tanh(zz[k]);
                }
              } else {
#pragma unroll
                for (int k = 0; k < 32; ++k)
                  if (col0 + k < p.N) f[k] *= dgelu_tanh(__bfloat162float(z[k]));
              }
            }
          }
          if (p.tma_store) {
            stage_store(&map_y, f);
          } else if (p.out_dtype == PSB_BF16) {
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
                st_stream_f4(dst + q * 4,
                             make_float4(f[q * 4], f[q * 4 + 1], f[q * 4 + 2], f[q * 4 + 3]));
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
      if (lane == 0) mbar_arrive_cluster(&tmem_empty[acc], 0);  // leader's barrier
      if (p.clc) {
        t = clc_next(cs, cph);
        __syncwarp();  // every lane has read the response before the slot is handed back
        if (lane == 0) mbar_arrive_cluster(&clc_empty[cs], 0);
        if (++cs == CLC_SLOTS) cs = 0, cph ^= 1;
      } else {
        t += step;
      }
    }
    // the staged pieces must have left shared memory (and landed) before the CTA goes away
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc_fence_before();
  cluster_sync();  // nobody touches the peer's smem / TMEM after this
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc2(tmem_base, TMEM_COLS);
  }
}

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

// `ld`: elements between two rows of the source (0 = packed)
static int make_map(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                    uint64_t ld = 0) {
  EncodeTiledFn enc = get_encode();
  PSB_REQUIRE(enc, PSB_EUNSUPPORTED, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {(ld ? ld : cols) * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, 128};
  cuuint32_t elem[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                   box, elem, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PSB_REQUIRE(r == CUDA_SUCCESS, PSB_EINVAL, "cuTensorMapEncodeTiled failed with CUresult %d",
              (int)r);
  return 0;
}

// MN-major source: `red` rows of `mn` contiguous elements; one box = one 64-element atom
static int make_map_mn(CUtensorMap* map, const void* base, uint64_t red, uint64_t mn,
                       uint64_t ld = 0) {
  EncodeTiledFn enc = get_encode();
  PSB_REQUIRE(enc, PSB_EUNSUPPORTED, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t dims[2] = {mn, red};
  cuuint64_t strides[1] = {(ld ? ld : mn) * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)BK};
  cuuint32_t elem[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                   box, elem, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PSB_REQUIRE(r == CUDA_SUCCESS, PSB_EINVAL, "cuTensorMapEncodeTiled (MN-major) failed with CUresult %d",
              (int)r);
  return 0;
}

// Output (or Z) as the epilogue stores it: 32 x 32 bf16 pieces, 64-byte rows, 64B swizzle
static int make_map_out(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols) {
  EncodeTiledFn enc = get_encode();
  PSB_REQUIRE(enc, PSB_EUNSUPPORTED, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {32, 32};
  cuuint32_t elem[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                   box, elem, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B,
                   CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PSB_REQUIRE(r == CUDA_SUCCESS, PSB_EINVAL, "cuTensorMapEncodeTiled (output) failed with CUresult %d",
              (int)r);
  return 0;
}

// PSB_GEMM_TMA_STORE=0: bf16 outputs are stored straight from registers (the round-1 epilogue)
static bool tma_store_enabled() {
  static const bool on = [] {
    const char* e = getenv("PSB_GEMM_TMA_STORE");
    return !(e && e[0] == '0');
  }();
  return on;
}

template <int EPI>
static int launch_epi(const CUtensorMap& map_a, const CUtensorMap& map_b, const PeerMaps& peers,
                      const CUtensorMap& map_y, const CUtensorMap& map_z, const Params& p,
                      int clusters, cudaStream_t s) {
  // per launch, not once per process: the attribute belongs to the current device's context
  PSB_CUDA_OK(cudaFuncSetAttribute(gemm2_bf16_kernel<EPI>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
  gemm2_bf16_kernel<EPI><<<clusters * 2, THREADS, SMEM_BYTES, s>>>(map_a, map_b, peers, map_y, map_z, p);
  PSB_LAUNCH_OK(EPI == EPI_NONE ? "gemm_bf16_tcgen05_2cta"
                                : EPI == EPI_GELU ? "gemm_bf16_tcgen05_2cta_gelu"
                                                  : "gemm_bf16_tcgen05_2cta_dgelu");
  return 0;
}

// PSB_GEMM_SCHED=static keeps the persistent static tile walk for plain GEMMs as well
static bool clc_enabled() {
  static const bool on = [] {
    const char* e = getenv("PSB_GEMM_SCHED");
    return !(e && e[0] == 's');
  }();
  return on;
}

static int launch(const void* a, const void* b, const Params& p, int epi, cudaStream_t s,
                  int lda = 0, int ldb = 0) {
  CUtensorMap map_a, map_b;
  if (int rc = p.a_mn ? make_map_mn(&map_a, a, p.K, p.M, lda) : make_map(&map_a, a, p.M, p.K, lda))
    return rc;
  if (int rc = p.b_mn ? make_map_mn(&map_b, b, p.K, p.N, ldb) : make_map(&map_b, b, p.N, p.K, ldb))
    return rc;
  const int num_tiles = ((p.M + BM - 1) / BM) * ((p.N + BN - 1) / BN);
  int clusters = num_tiles < kNumSMs / 2 ? num_tiles : kNumSMs / 2;
  Params q = p;
  static const PeerMaps kNoPeers = {};
  PeerMaps fused_peers;
  if (p.fused) {  // the owners' replicas of W under the same box as the local one
    for (int r = 0; r < p.world; ++r)
      if (int rc = make_map(&fused_peers.m[r], p.w_peer[r], p.N, p.K)) return rc;
    for (int r = p.world; r < PSB_MAX_PEERS; ++r) fused_peers.m[r] = fused_peers.m[0];
  }
  const PeerMaps& peers = p.fused ? fused_peers : kNoPeers;
  if (p.fused) {
    clusters = kNumSMs / 2;
    if (num_tiles + p.gather_ctas / 2 < clusters) clusters = num_tiles + p.gather_ctas / 2;
  } else if (clc_enabled()) {
    q.clc = 1;
    clusters = num_tiles;  // one pair per tile; running pairs cancel and absorb the rest
  }
  q.group_m = 8;
  if (const char* e = getenv("PSB_GEMM_GROUPM")) {
    const int v = atoi(e);
    if (v >= 1 && v <= 64) q.group_m = v;
  }
  CUtensorMap map_y = map_a, map_z = map_a;  // placeholders unless the TMA-store epilogue is on
  q.tma_store = p.out_dtype == PSB_BF16 && tma_store_enabled();
  if (q.tma_store) {
    if (int rc = make_map_out(&map_y, p.y, p.M, p.N)) return rc;
    if (epi == EPI_GELU && p.aux_out)
      if (int rc = make_map_out(&map_z, p.aux_out, p.M, p.N)) return rc;
    if (epi == EPI_DGELU)  // the epilogue's Z input travels by TMA through the same slot
      if (int rc = make_map_out(&map_z, p.aux_in, p.M, p.N)) return rc;
  }
  switch (epi) {
    case EPI_NONE: return launch_epi<EPI_NONE>(map_a, map_b, peers, map_y, map_z, q, clusters, s);
    case EPI_GELU: return launch_epi<EPI_GELU>(map_a, map_b, peers, map_y, map_z, q, clusters, s);
    case EPI_DGELU: return launch_epi<EPI_DGELU>(map_a, map_b, peers, map_y, map_z, q, clusters, s);
  }
  PSB_REQUIRE(false, PSB_EINVAL, "gemm: unknown epilogue %d", epi);
  return 0;
}

}  // namespace gemm2

// Entry used by gemm_tcgen05.cu when the pair kernel is selected.  `fused` fields of the
// caller's parameter block are passed through unchanged.
int gemm2_launch(const void* x, const void* w, void* y, const float* bias, int M, int N, int K,
                 int out_dtype, int fused, int gather_ctas, int rank, int world,
                 const char* const* w_peer, char* w_local, uint32_t* tile_flags,
                 const uint32_t* version_pad, const uint32_t* epoch, int skip_wait,
                 cudaStream_t s, int epilogue, void* aux_out, int bias_bf16) {
  using namespace gemm2;
  Params p{};
  p.aux_out = reinterpret_cast<__nv_bfloat16*>(aux_out);
  p.bias_bf16 = bias_bf16;
  p.y = y, p.bias = bias, p.M = M, p.N = N, p.K = K, p.out_dtype = out_dtype;
  p.fused = fused, p.gather_ctas = gather_ctas, p.rank = rank, p.world = world;
  for (int i = 0; i < world && i < PSB_MAX_PEERS; ++i) p.w_peer[i] = w_peer ? w_peer[i] : nullptr;
  p.w_local = w_local, p.tile_flags = tile_flags, p.version_pad = version_pad, p.epoch = epoch;
  p.skip_wait = skip_wait;
  if (fused) {
    // Gather mechanism and SM split, from measurements (profiles/r2/fused_gemm_*): the register-
    // path copiers are the default — level with the TMA ring at 2 GPUs (4096x8192x4096: 0.213 vs
    // 0.207 ms; 1024 and 16384 rows: 0.100 / 0.762 vs 0.113 / 0.790) and faster from 4 GPUs up
    // (8 GPUs: 0.221 vs 0.279 ms).  PSB_GATHER=tma selects the TMA ring.
    const double t_gemm = 2.0 * M * N * (double)K / 1.5e15;
    const double t_link = (double)(world - 1) / world * N * (double)K * 2.0 / 700e9;
    const double r = t_link / t_gemm;
    p.gather_tma = 0;
    if (const char* e = getenv("PSB_GATHER")) p.gather_tma = e[0] == 't';
    if (p.gather_tma) p.gather_ctas = r < 0.15 ? 8 : r < 0.3 ? 16 : 32;
    else p.gather_ctas = r < 0.15 ? 16 : r < 0.4 ? 24 : r < 0.8 ? 40 : 64;
    if (const char* e = getenv("PSB_GATHER_CTAS")) {
      const int v = atoi(e);
      if (v >= 2 && v <= 64) p.gather_ctas = v & ~1;
    }
  }
  return launch(x, w, p, epilogue, s);
}

}  // namespace psb

extern "C" int psb_gemm_bf16(const PsbGemm* g, psb_stream_t stream) {
  using namespace psb;
  using namespace psb::gemm2;
  PSB_REQUIRE(g && g->a && g->b && g->y, PSB_EINVAL, "psb_gemm_bf16: null operand");
  PSB_REQUIRE(g->M > 0 && g->N > 0 && g->K > 0, PSB_EINVAL, "psb_gemm_bf16: empty problem %dx%dx%d",
              g->M, g->N, g->K);
  PSB_REQUIRE(g->out_dtype == PSB_BF16 || g->out_dtype == PSB_F32, PSB_EINVAL,
              "psb_gemm_bf16: bad output dtype %d", g->out_dtype);
  // every row of every operand starts on a 16-byte boundary (TMA strides, vector epilogue)
  PSB_REQUIRE((g->trans_a ? g->M : g->K) % 8 == 0, PSB_EINVAL,
              "psb_gemm_bf16: the contiguous dimension of A (%d) must be a multiple of 8",
              g->trans_a ? g->M : g->K);
  PSB_REQUIRE((g->trans_b ? g->N : g->K) % 8 == 0, PSB_EINVAL,
              "psb_gemm_bf16: the contiguous dimension of B (%d) must be a multiple of 8",
              g->trans_b ? g->N : g->K);
  PSB_REQUIRE(g->N % 8 == 0, PSB_EINVAL, "psb_gemm_bf16: N=%d must be a multiple of 8", g->N);
  PSB_REQUIRE(((uintptr_t)g->a & 15) == 0 && ((uintptr_t)g->b & 15) == 0 &&
                  ((uintptr_t)g->y & 15) == 0 && ((uintptr_t)g->aux_in & 15) == 0 &&
                  ((uintptr_t)g->aux_out & 15) == 0,
              PSB_EINVAL, "psb_gemm_bf16: operands must be 16-byte aligned");
  PSB_REQUIRE(g->epilogue >= PSB_EPI_NONE && g->epilogue <= PSB_EPI_DGELU, PSB_EINVAL,
              "psb_gemm_bf16: unknown epilogue %d", g->epilogue);
  PSB_REQUIRE(g->epilogue != PSB_EPI_DGELU || g->aux_in, PSB_EINVAL,
              "psb_gemm_bf16: the dGELU epilogue needs the pre-activation in aux_in");
  Params p{};
  PSB_REQUIRE(g->bias_dtype == PSB_F32 || g->bias_dtype == PSB_BF16, PSB_EINVAL,
              "psb_gemm_bf16: bad bias dtype %d", g->bias_dtype);
  p.bias_bf16 = g->bias_dtype == PSB_BF16;
  p.y = g->y, p.bias = reinterpret_cast<const float*>(g->bias), p.M = g->M, p.N = g->N, p.K = g->K, p.out_dtype = g->out_dtype;
  p.a_mn = g->trans_a != 0, p.b_mn = g->trans_b != 0;
  p.aux_in = reinterpret_cast<const __nv_bfloat16*>(g->aux_in);
  p.aux_out = reinterpret_cast<__nv_bfloat16*>(g->aux_out);
  PSB_REQUIRE(g->lda >= 0 && g->ldb >= 0 && g->lda % 8 == 0 && g->ldb % 8 == 0, PSB_EINVAL,
              "psb_gemm_bf16: lda=%d / ldb=%d must be non-negative multiples of 8", g->lda, g->ldb);
  PSB_REQUIRE(g->lda == 0 || g->lda >= (g->trans_a ? g->M : g->K), PSB_EINVAL,
              "psb_gemm_bf16: lda=%d is shorter than a row of A", g->lda);
  PSB_REQUIRE(g->ldb == 0 || g->ldb >= (g->trans_b ? g->N : g->K), PSB_EINVAL,
              "psb_gemm_bf16: ldb=%d is shorter than a row of B", g->ldb);
  return launch(g->a, g->b, p, g->epilogue, (cudaStream_t)stream, g->lda, g->ldb);
}

PSB_EXPORT_SPIN_LIMIT(gemm2)
