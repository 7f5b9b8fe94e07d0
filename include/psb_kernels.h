// psb_kernels.h — C ABI of libpsb_kernels.so, the sm_100a device library of the
// B200-native parameter-server framework.
//
// Every entry point is `extern "C"`, takes raw device pointers plus a CUDA
// stream, launches asynchronously on that stream and returns 0 on success or a
// non-zero code (a cudaError_t value, or a negative PSB_E* code); the message
// of the last failure on the calling thread is available from
// psb_last_error().  The Python package binds these with ctypes, the C++
// parameter_server / worker_main binaries link them directly.
//
// Kernel IDs (K1..K9) refer to SURVEY.md §8.  Reference citations are paths
// below /root/reference.
#pragma once

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* psb_stream_t;  // cudaStream_t

#define PSB_MAX_PEERS 8
#define PSB_MAX_BUCKETS 512
#define PSB_MAX_GATHER 96

// error codes (negative so they never collide with cudaError_t)
#define PSB_EINVAL (-1)
#define PSB_ERANGE (-2)
#define PSB_EUNSUPPORTED (-3)

// dtypes of gradient payloads
#define PSB_F32 0
#define PSB_BF16 1

// optimizer kinds (server-side update; the reference applies only `p -= g`,
// src/parameter_server.cpp:83-90 — PSB_OPT_SGD with lr = 1 reproduces it)
#define PSB_OPT_SGD 0
#define PSB_OPT_MOMENTUM 1
#define PSB_OPT_ADAM 2   // L2-coupled weight decay
#define PSB_OPT_ADAMW 3  // decoupled weight decay

const char* psb_last_error(void);
int psb_version(void);
// number of kernel launches issued through this library by the calling process
uint64_t psb_launch_count(void);

// ---------------------------------------------------------------------------
// Hyper-parameters of the server-side update.
// ---------------------------------------------------------------------------
typedef struct PsbOptim {
  int32_t kind;        // PSB_OPT_*
  float lr;
  float momentum;      // PSB_OPT_MOMENTUM
  float beta1, beta2;  // Adam
  float eps;
  float weight_decay;
  float grad_scale;    // multiplies the *summed* gradient (1/num_workers for the mean,
                       // src/parameter_server.cpp:57-60)
} PsbOptim;

// ---------------------------------------------------------------------------
// K7 — constant fill (replaces the host loop src/worker.cpp:317-320)
// ---------------------------------------------------------------------------
int psb_fill_f32(float* dst, size_t n, float value, psb_stream_t stream);
int psb_fill_bf16(void* dst, size_t n, float value, psb_stream_t stream);

// ---------------------------------------------------------------------------
// K6 — scale / cast (replaces the host divide src/worker.cpp:436-438)
// ---------------------------------------------------------------------------
int psb_cast_f32_to_bf16(const float* src, void* dst, size_t n, float scale, psb_stream_t stream);
int psb_cast_bf16_to_f32(const void* src, float* dst, size_t n, float scale, psb_stream_t stream);

// Multi-tensor gather: copies `count` separately allocated gradient tensors into
// one flat bucket (a view of the symmetric heap) in a single launch, casting to
// the bucket dtype and scaling.  Each source must be 16-byte aligned; each
// dst_offset (in elements) a multiple of 8.
typedef struct PsbGatherList {
  const void* src[PSB_MAX_GATHER];
  uint64_t dst_offset[PSB_MAX_GATHER];  // elements into the bucket
  uint64_t numel[PSB_MAX_GATHER];
  int32_t count;
  int32_t src_dtype;  // PSB_F32 / PSB_BF16
  int32_t dst_dtype;
  float scale;
} PsbGatherList;
int psb_gather_to_bucket(const PsbGatherList* list, void* bucket, psb_stream_t stream);

// ---------------------------------------------------------------------------
// K3 — stand-alone flat-shard optimizer step (H3 of the reference generalised):
// fp32 master weights (+ optional m, v), gradient in fp32 or bf16, optional bf16
// copy of the updated weights.  `step` is a device counter holding the number of
// completed steps (Adam bias correction uses step+1); it is NOT advanced here —
// call psb_counter_add afterwards (once per optimizer step, not per bucket).
// `clip_coef` may be NULL, else a device float that multiplies the gradient.
// ---------------------------------------------------------------------------
int psb_optim_step(const PsbOptim* opt, float* master, float* m, float* v, const void* grad,
                   int grad_dtype, void* param_bf16_out, size_t n, const float* clip_coef,
                   const uint32_t* step, psb_stream_t stream);
int psb_counter_add(uint32_t* counter, uint32_t delta, psb_stream_t stream);

// ---------------------------------------------------------------------------
// K4 — gradient norm / clipping.
// psb_sumsq: out[0] = Σ x², two-phase reduction.  `partials` is caller-provided
// scratch of at least psb_sumsq_scratch_floats() floats; it need not be zeroed.
// psb_clip_coef: coef = min(1, max_norm / (sqrt(Σ parts) + eps)).
// ---------------------------------------------------------------------------
size_t psb_sumsq_scratch_floats(void);
int psb_sumsq(const void* x, int dtype, size_t n, float* partials, float* out, psb_stream_t stream);
int psb_clip_coef(const float* sumsq_parts, int n_parts, float max_norm, float eps, float* coef_out,
                  float* norm_out, psb_stream_t stream);

// psb_sumsq_noalloc — Σx² without any caller-provided or freshly allocated scratch.
// No counterpart in the reference (it has no device code and no gradient norm at all).
//
// Who needs it: code that is not allowed to allocate GPU memory at the point where it wants a
// gradient norm — a service that hands this library a fixed, fully pre-carved memory pool; a
// monitoring hook that runs while the framework's allocator is locked or already exhausted; a
// plain C host program that only passes pointers it was given.  psb_sumsq asks such callers
// for a scratch buffer they cannot make; this entry point brings its own.
//
// The partial sums live in ONE static device array per GPU: PSB_SUMSQ_NOALLOC_SLOTS floats
// (4 KB), resident on every GPU that loads the library for the life of the process.  Each slot
// covers at most PSB_SUMSQ_NOALLOC_ELEMS_PER_SLOT elements, so inputs longer than
// PSB_SUMSQ_NOALLOC_MAX_ELEMS (2^31) are rejected with PSB_ERANGE and a message — use
// psb_sumsq for those.  The function never calls cudaMalloc / cudaFree or any allocating
// routine.
//
// First call on a device: looks up the array's device address (cudaGetSymbolAddress, with
// that device current) and remembers it.  Make that first call during set-up, before any
// stream capture begins.
// Sharing: all calls on a GPU use the same array, which keeps whatever the last call left in
// it.  Nothing needs to start at zero (each call overwrites every slot it later reads), but
// two calls that overlap on the GPU — different streams, or different host threads without
// ordering — race on it: serialise them on one stream or with events.
// `x` must be 16-byte aligned; `out` receives one float; dtype is PSB_F32 or PSB_BF16.
#define PSB_SUMSQ_NOALLOC_SLOTS 1024
#define PSB_SUMSQ_NOALLOC_ELEMS_PER_SLOT (1 << 21)
#define PSB_SUMSQ_NOALLOC_MAX_ELEMS (1ull << 31)
int psb_sumsq_noalloc(const void* x, int dtype, size_t n, float* out, psb_stream_t stream);

// ---------------------------------------------------------------------------
// Fabric — symmetric heap over the GPUs of one NVSwitch box.  Replaces the
// reference's per-iteration cudaMalloc + host staging (src/worker.cpp:416-444)
// and its NCCL clique (src/nccl_manager.cpp:28-85).
// ---------------------------------------------------------------------------
#define PSB_IPC_HANDLE_BYTES 64
int psb_device_count(int* count);
int psb_set_device(int dev);
int psb_heap_alloc(size_t bytes, void** ptr);  // cudaMalloc + zero
int psb_heap_free(void* ptr);
int psb_ipc_export(void* ptr, unsigned char handle[PSB_IPC_HANDLE_BYTES]);
int psb_ipc_open(const unsigned char handle[PSB_IPC_HANDLE_BYTES], void** ptr);
int psb_ipc_close(void* ptr);
int psb_can_access_peer(int dev, int peer, int* ok);
int psb_enable_peer_access(int peer);  // from the current device
int psb_memcpy_d2h(void* dst_host, const void* src_dev, size_t bytes, psb_stream_t stream);
int psb_memcpy_h2d(void* dst_dev, const void* src_host, size_t bytes, psb_stream_t stream);
int psb_memset(void* dst_dev, int value, size_t bytes, psb_stream_t stream);
int psb_stream_sync(psb_stream_t stream);
int psb_device_sync(void);

// Signal pad layout (uint32 words, one pad per GPU, peer-mapped):
//   ready  [PSB_MAX_BUCKETS][PSB_MAX_PEERS]   written by worker w on the owner's pad
//   version[PSB_MAX_BUCKETS][PSB_MAX_PEERS]   written by owner r on every worker's pad
//   barrier[64][PSB_MAX_PEERS]                generic device barriers
//   clock  [PSB_MAX_PEERS]                    SSP clocks (one word per worker)
//   norm   [2][PSB_MAX_PEERS] (64-bit words)   {seq, float bits} of per-shard Σg², double-buffered
//   ack    [PSB_MAX_BUCKETS][PSB_MAX_PEERS]   async mode: owner r consumed worker's clock
//   done   [PSB_MAX_PEERS]                    async mode: final clock of a worker that finished
#define PSB_PAD_READY_OFF 0
#define PSB_PAD_VERSION_OFF (PSB_MAX_BUCKETS * PSB_MAX_PEERS)
#define PSB_PAD_BARRIER_OFF (2 * PSB_MAX_BUCKETS * PSB_MAX_PEERS)
#define PSB_PAD_CLOCK_OFF (PSB_PAD_BARRIER_OFF + 64 * PSB_MAX_PEERS)
#define PSB_PAD_NORM_OFF (PSB_PAD_CLOCK_OFF + PSB_MAX_PEERS)
#define PSB_PAD_ACK_OFF (PSB_PAD_NORM_OFF + 4 * PSB_MAX_PEERS)
#define PSB_PAD_DONE_OFF (PSB_PAD_ACK_OFF + PSB_MAX_BUCKETS * PSB_MAX_PEERS)
#define PSB_PAD_WORDS (PSB_PAD_DONE_OFF + PSB_MAX_PEERS)
size_t psb_pad_bytes(void);

typedef struct PsbFabric {
  int32_t rank;
  int32_t world;
  uint32_t* pad[PSB_MAX_PEERS];  // peer-mapped signal pads; pad[rank] is local
} PsbFabric;

// ---------------------------------------------------------------------------
// K1 — fused gradient push: reduce-scatter over NVLink peer loads into the
// owning shard + server-side optimizer update + (optionally) push-mode
// all-gather of the updated bf16 slice into every replica, in ONE kernel and
// without NCCL.  Replaces N1 + H1 + H2 + H3 (+H4) of the reference
// (src/nccl_manager.cpp:111, src/parameter_server.cpp:49-60,83-90,
//  src/worker.cpp:436-438).
//
// Protocol for bucket b at epoch e = *epoch + 1 (device word, advanced by the
// kernel itself so a captured CUDA graph can be replayed):
//   1. every rank w stores ready[b][w] = e on every owner's pad (release.sys)
//   2. owner r spins until ready[b][0..N) >= e on its own pad (acquire.sys)
//   3. for its slice: g = Σ_w grad_w (fp32, fixed order w = 0..N-1), scaled,
//      optimizer update of master/m/v, bf16 slice written locally and, with
//      PSB_PUSH_BROADCAST, into every peer replica
//   4. last CTA stores version[b][r] = e on every pad and *epoch = e
// flags: see PSB_PUSH_* below.
// ---------------------------------------------------------------------------
#define PSB_PUSH_BROADCAST 1    // write the updated slice into all replicas (push-mode all-gather)
#define PSB_PUSH_NO_WAIT 2      // skip steps 1-2 (caller already synchronised; async single-pusher)
#define PSB_PUSH_REDUCE_ONLY 4  // stop after the reduction: write fp32 sum to `reduced`, Σg² to norm pads
#define PSB_PUSH_FROM_REDUCED 8 // skip the reduction: gradient comes from `reduced` (after clipping)
#define PSB_PUSH_MASKED 16      // only sum peers whose bit is set in *peer_mask (async / SSP);
                                // an empty mask makes the launch a no-op (nothing published)
#define PSB_PUSH_TMA 64         // bulk-copy (cp.async.bulk) smem-ring variant: fills NVLink from ~24 SMs
                                // instead of ~148, so backward kernels keep the other SMs
#define PSB_PUSH_NVLS 128       // NVLS variant (push_nvls.cu): multimem.ld_reduce over grad_mc, multimem.st
                                // into param_mc — the switch reduces and replicates; bf16 gradients, VMM
                                // heap bound to a multicast object (psb_mc_*); bf16-rounded sum
#define PSB_PUSH_SEQLOCK 32     // publish 2*epoch (even) instead of epoch: readers pair it with the
                                // odd mark written by psb_async_plan before the update started

typedef struct PsbPush {
  PsbFabric fabric;
  const void* grad[PSB_MAX_PEERS];  // peer-mapped base of this bucket's gradient buffer
  void* param[PSB_MAX_PEERS];       // peer-mapped base of this bucket's bf16 replica
  float* master;                    // local shard state, slice_elems each
  float* m;
  float* v;
  float* reduced;                   // optional fp32 scratch, slice_elems (clip path)
  float* sumsq_acc;                 // optional local float: += Σ(reduced²) (PSB_PUSH_REDUCE_ONLY)
  uint64_t slice_elems;             // bucket = world * slice_elems elements
  int32_t bucket;
  int32_t grad_dtype;
  uint32_t* epoch;                  // local device word, one per bucket
  const uint32_t* step;             // local device word, optimizer step counter
  const float* clip_coef;           // optional
  const uint32_t* peer_mask;        // optional (PSB_PUSH_MASKED)
  uint32_t* done_counter;           // local device word, zero-initialised, one per bucket
  int32_t flags;
  int32_t grid;                     // CTAs (0 = auto)
  PsbOptim opt;
  const uint64_t* grad_off;         // optional device array [world]: byte offset added to grad[w]
                                    // (async mode: which slot of worker w's gradient ring to read)
  int32_t block;                    // threads per CTA of the register-path kernel (0 = 512).  256
                                    // lets a K1 CTA share an SM with a CTA of the tcgen05 GEMM
                                    // (32 K + 24 K registers); measured slower than taking turns
                                    // (profiles/overlap_matrix_1gpu.txt), kept as an experiment knob
  const void* grad_mc;              // PSB_PUSH_NVLS: multicast address of this bucket's gradient buffer
  void* param_mc;                   // PSB_PUSH_NVLS: multicast address of this bucket's bf16 replica
} PsbPush;
int psb_push_fused(const PsbPush* args, psb_stream_t stream);
// Steps 1-2 of the protocol alone (one warp): announce READY for args->bucket at epoch
// *epoch + 1 and wait for every rank's.  Followed on the same stream by psb_push_fused with
// PSB_PUSH_NO_WAIT it gives the same result as one waiting launch, but the data kernel's CTAs
// occupy SMs only once their inputs exist.
int psb_push_ready(const PsbPush* args, psb_stream_t stream);

// ---------------------------------------------------------------------------
// K2 — parameter pull: wait for version[b][*] >= *epoch, then gather the N
// slices of the bucket from their owners into the local replica (peer loads).
// With PSB_PUSH_BROADCAST the data is already local and only the wait remains
// (psb_wait_versions).  Replaces ServeParameters + H6/H7
// (src/parameter_server.cpp:93-97, src/worker.cpp:240-252).
// ---------------------------------------------------------------------------
typedef struct PsbPull {
  PsbFabric fabric;
  void* param[PSB_MAX_PEERS];  // peer-mapped replicas of this bucket
  uint64_t slice_bytes;
  int32_t bucket;
  const uint32_t* epoch;       // local device word of this bucket
  int32_t grid;
  int32_t skip_wait;
  int32_t seqlock;             // async mode: copy each owner slice under a seqlock (retry on change)
} PsbPull;
int psb_pull_allgather(const PsbPull* args, psb_stream_t stream);

// K2 over NVLS: the owner stores its slice of the bucket (param_local + rank * slice_bytes)
// into every replica with multimem.st through `param_mc`; with `epoch` it then publishes
// version[bucket][rank] = *epoch (consumers: psb_wait_versions).  Push-mode all-gather in which
// each GPU sends its slice once and the switch replicates it.
int psb_bcast_slice_nvls(const PsbFabric* fabric, const void* param_local, void* param_mc,
                         uint64_t slice_bytes, int bucket, const uint32_t* epoch,
                         uint32_t* done_counter, int grid, psb_stream_t stream);

// VMM heap and multicast (NVLS) plumbing — csrc/fabric/fabric.cu.  A VMM segment is what the
// IPC heap above is, but allocated with cuMemCreate so that (a) it travels between processes as
// a POSIX file descriptor and (b) it can be bound to a multicast object.
int psb_vmm_support(int* vmm, int* multicast);
int psb_vmm_granularity(int world, int multicast, int recommended, size_t* gran);
int psb_vmm_alloc(size_t bytes, void** ptr, uint64_t* handle, int* fd);
int psb_vmm_import(int fd, size_t bytes, void** ptr, uint64_t* handle);
int psb_vmm_release(void* ptr, size_t bytes, uint64_t handle);
int psb_mc_create(size_t bytes, int world, uint64_t* mc_handle, int* fd);
int psb_mc_import(int fd, uint64_t* mc_handle);
int psb_mc_add_device(uint64_t mc_handle);                 // every rank, then a barrier
int psb_mc_bind_map(uint64_t mc_handle, uint64_t local_handle, size_t bytes, void** mc_ptr);
int psb_mc_unbind(uint64_t mc_handle, size_t bytes);

// K9 — device-side waits / barriers (replace the 50 ms CheckSyncStatus poll and
// the 500 ms sleeps, src/worker.cpp:359-375).
// wait until version[b][r] >= epoch_words[b] for b in [bucket_lo, bucket_hi), r in [0, world)
int psb_wait_versions(const PsbFabric* fabric, const uint32_t* epoch_words, int bucket_lo,
                      int bucket_hi, psb_stream_t stream);
// all-rank barrier on slot `slot` (0..63); `seq` is a local device word advanced by the kernel
int psb_device_barrier(const PsbFabric* fabric, int slot, uint32_t* seq, psb_stream_t stream);
// cross-shard sum of the N per-shard Σg² written by PSB_PUSH_REDUCE_ONLY, then clip coefficient
// `local_sumsq` is consumed: the kernel zeroes it after reading.
int psb_norm_allreduce(const PsbFabric* fabric, float* local_sumsq, uint32_t* seq, float max_norm,
                       float eps, float* coef_out, float* norm_out, psb_stream_t stream);

// ---------------------------------------------------------------------------
// Asynchronous / bounded-staleness mode (no counterpart in the reference, which is BSP only;
// PullRequest.iteration / ParameterUpdate.ready are the wire carriers on the host path).
// Every rank is a worker with its own clock and the owner of one slice per bucket.
//   worker:  stages clock c into slot c % ring of its gradient ring, then psb_async_signal
//            raises ready[b][w] = c on every owner's pad and publishes clock[w] = c
//   owner:   psb_async_plan picks, per bucket, the workers with an unapplied clock (one clock
//            per worker per call, in order), writes the peer mask + ring offsets, advances
//            applied[b][w] and marks version[b][r] odd on every pad; K1 (MASKED | SEQLOCK)
//            applies the picked gradients and publishes the even version; psb_async_ack tells
//            each worker which of its clocks have been consumed (ring slot reusable)
//   gate:    psb_async_gate blocks the stream until clock c may start: every owner consumed
//            clock c - ring (slot free) and, with staleness >= 0, no worker is more than
//            `staleness` clocks behind c - 1
// ---------------------------------------------------------------------------
typedef struct PsbAsync {
  PsbFabric fabric;
  int32_t num_buckets;
  int32_t ring;                 // slots in the gradient ring
  uint64_t ring_stride_bytes;   // bytes between two slots
  uint32_t* applied;            // local [num_buckets][PSB_MAX_PEERS]: last applied clock per worker
  uint32_t* masks;              // local [num_buckets]: peer mask chosen by the plan
  uint64_t* grad_off;           // local [num_buckets][PSB_MAX_PEERS]: ring offsets chosen by the plan
  const uint32_t* epoch;        // local [num_buckets]: K1 epoch words
} PsbAsync;
int psb_async_signal(const PsbAsync* a, uint32_t clock, psb_stream_t stream);
int psb_async_plan(const PsbAsync* a, psb_stream_t stream);
int psb_async_ack(const PsbAsync* a, psb_stream_t stream);
int psb_async_gate(const PsbAsync* a, uint32_t clock, int32_t staleness, psb_stream_t stream);
// store `value` into word [word_base + rank] of every rank's pad (release.sys)
int psb_pad_post(const PsbFabric* fabric, uint32_t word_base, uint32_t value, psb_stream_t stream);
// Watchdog of every spin-wait on a word written by a peer (K1 ready flags, K2 / K9 version
// flags, barriers, norm all-reduce, async clocks, the gather of K2f): a kernel that has been
// waiting longer than this traps instead of spinning for ever on a dead peer; the launch then
// fails with a CUDA error and the GPU is released.  Default 300 s; per device (call it with
// the device current).  No counterpart in the reference, whose host-side waits poll for ever
// (src/worker.cpp:293-306).
int psb_set_spin_timeout_ms(uint64_t ms);
// Abort word: the soft counterpart of the watchdog.  psb_abort_word_create allocates one 32-bit
// word of host-mapped memory, points the spin-waits of the CURRENT device at it and returns its
// host address.  Whoever stores 1 there — the coordinator client's heartbeat thread when a peer
// was evicted (psbc_watch, include/psb_control.h), or the application — makes every kernel that
// is waiting on a peer's flag return within ~0.1 ms instead of trapping after the watchdog
// limit: the CUDA context survives, so the rank can leave cleanly (the results of kernels that
// gave up are undefined; training resumes from the last checkpoint with the smaller world).
// The reference has no counterpart: its workers wait for a dead peer's iteration for ever
// (src/worker.cpp:372-397).
int psb_abort_word_create(uint32_t** host_word);
int psb_abort_word_destroy(uint32_t* host_word);

// ---------------------------------------------------------------------------
// K2f — bf16 GEMM on tcgen05 tensor cores with TMEM accumulators and TMA-fed
// shared-memory tiles:  Y[M,N] = X[M,K] · W[N,K]^T (+ bias), fp32 accumulate.
// psb_gemm_bf16_tn: all operands local.
// psb_gemm_allgather_bf16: W is row-sharded over the peers (rank r owns rows
// [r*N/world, (r+1)*N/world)).  The all-gather of W runs INSIDE the GEMM kernel: a few CTAs
// do no math and stream the remote row-slices from the owners' replicas over NVLink into the
// local replica, bumping an arrival counter per 256-row N-tile; the math clusters start on the
// locally owned tiles and acquire a tile's counter before their TMA loads touch it.  Two gather
// mechanisms (PSB_GATHER=ldg|tma): warp copiers with 16-byte peer loads, two 8 KB items in
// flight per warp — the default, as fast or faster at every GPU count measured — and a
// shared-memory ring on the TMA unit: tensor-map loads on the OWNER's map (peer -> smem,
// 11 x 16 KB in flight per CTA) and TMA stores into the local replica (level at 2 GPUs, slower
// from 4 up, profiles/r2/fused_gemm_*).  The
// gathered W stays in the local replica for the backward pass.  (B tiles are not fed to the
// MMA straight from the peer: a peer TMA load takes ~4 us, the ring of the MMA pipeline covers
// ~1.6 us, and every W tile is needed by M/256 row blocks — fetched once, shared through L2.)
// Optional GELU epilogue as in psb_gemm_bf16.  No counterpart in the reference (north-star
// item).
// ---------------------------------------------------------------------------
// kernel variant: 1 (default) = CTA pair per 256x256 tile (tcgen05.mma.cta_group::2),
// 0 = one CTA per 128x256 tile (cta_group::1).  Also PSB_GEMM_VARIANT=1cta|2cta.
int psb_gemm_select(int variant);
int psb_gemm_bf16_tn(const void* x, const void* w, const float* bias, void* y, int M, int N, int K,
                     int out_dtype, psb_stream_t stream);
enum { PSB_EPI_NONE = 0, PSB_EPI_GELU = 1, PSB_EPI_DGELU = 2 };  // fused epilogues, see psb_gemm_bf16
typedef struct PsbGemmAG {
  PsbFabric fabric;
  void* w[PSB_MAX_PEERS];  // peer-mapped replicas of W [N,K] bf16 (full-size on every rank)
  const void* x;           // [M,K] bf16, local
  const float* bias;       // optional [N] fp32
  void* y;                 // [M,N]
  int32_t M, N, K;
  int32_t out_dtype;       // PSB_BF16 / PSB_F32
  int32_t bucket;          // version flags of this bucket gate the remote loads
  const uint32_t* epoch;
  int32_t skip_wait;
  int32_t epilogue;        // PSB_EPI_NONE / PSB_EPI_GELU (Y = gelu(Z), Z = X·W^T + bias)
  void* aux_out;           // PSB_EPI_GELU: where Z goes (bf16 [M,N], optional)
  int32_t bias_dtype;      // PSB_F32 / PSB_BF16
} PsbGemmAG;
int psb_gemm_allgather_bf16(const PsbGemmAG* args, psb_stream_t stream);

// General form used by the training layers (forward, dX and dW of a Linear without any
// transposed copy):  Y[M,N] = op(A) · op(B) (+ bias) with an optional fused epilogue.
//   trans_a = 0: A stored [M,K] row-major (K-major operand)   1: A stored [K,M] (MN-major)
//   trans_b = 0: B stored [N,K] row-major (K-major operand)   1: B stored [K,N] (MN-major)
// forward  Y  = X · W^T      : trans_a 0, trans_b 0
// dX           = dY · W       : trans_a 0, trans_b 1
// dW           = dY^T · X     : trans_a 1, trans_b 1   (A = dY [batch,N_out], B = X [batch,K_in])
// bf16 outputs (Y and Z) leave the kernel through shared memory: every epilogue warp packs its
// 32 x 32 piece into a swizzled 2 KB buffer and one lane issues a TMA store; the dGELU input Z
// arrives by TMA one piece ahead (PSB_GEMM_TMA_STORE=0: the older register-path epilogue; fp32
// outputs always use it).
// Epilogues (tanh-form GELU): PSB_EPI_GELU stores Z = acc + bias rounded to bf16 into aux_out
// (optional) and gelu(Z) into Y; PSB_EPI_DGELU stores acc * gelu'(aux_in) — the activation
// gradient of the PREVIOUS layer is applied by the dX GEMM of this one.  CTA-pair kernel only.
typedef struct PsbGemm {
  const void* a;
  const void* b;
  void* y;             // [M,N] row-major, out_dtype
  const void* bias;    // optional [N], fp32 or bf16 (bias_dtype)
  const void* aux_in;  // PSB_EPI_DGELU: Z [M,N] bf16
  void* aux_out;       // PSB_EPI_GELU: Z [M,N] bf16 (optional)
  int32_t M, N, K;
  int32_t trans_a, trans_b;
  int32_t out_dtype;   // PSB_BF16 / PSB_F32
  int32_t epilogue;
  int32_t bias_dtype;  // PSB_F32 (0, default) or PSB_BF16: element type behind `bias`
  int32_t lda, ldb;    // elements between two stored rows of A / B (0 = packed).  A column block
                       // of a wider matrix is an operand in place: dW of rows [r0, r1) of a
                       // weight = dY[:, r0:r1]^T · X takes a = dY + r0, lda = N_out
} PsbGemm;
int psb_gemm_bf16(const PsbGemm* g, psb_stream_t stream);

// ---------------------------------------------------------------------------
// Memory-bound pieces of the training step between the GEMMs (train_ops.cu).
// psb_colsum_bf16: out[n] = sum_m x[m,n] (bias gradient), x bf16 [M,N], fp32 accumulation in a
//   fixed order; scratch of psb_colsum_scratch_bytes(M,N) bytes, zero-filled before first use
//   (it holds arrival counters that the kernel leaves at zero), one launch in flight per scratch.
// psb_mse_fwd_bf16 / psb_mse_bwd_bf16: loss = mean((a-b)^2) and da = 2(a-b)/n * (*gout);
//   scratch of PSB_MSE_SCRATCH_FLOATS floats, zero-filled before first use.
// ---------------------------------------------------------------------------
#define PSB_MSE_SCRATCH_FLOATS 304
int psb_colsum_slabs(int M, int N);
size_t psb_colsum_scratch_bytes(int M, int N);
int psb_colsum_bf16(const void* x, void* out, int out_dtype, int M, int N, void* scratch,
                    psb_stream_t stream);
int psb_mse_fwd_bf16(const void* a, const void* b, size_t n, float* scratch, float* loss,
                     psb_stream_t stream);
int psb_mse_bwd_bf16(const void* a, const void* b, const float* gout, size_t n, void* da,
                     psb_stream_t stream);

// ---------------------------------------------------------------------------
// K5 — sparse embedding push / pull (sparse.cu).  No counterpart in the reference: its Tensor
// is a dense `repeated float` (proto/parameter_server.proto:19-24) and every push carries the
// whole model (src/worker.cpp:253-275).  Here an embedding table of `vocab` rows x `dim`
// columns is row-sharded over the ranks (rank r owns rows [r*rows_per_rank, (r+1)*rows_per_rank)):
// fp32 master rows (+ m, v) stay with the owner, a bf16 copy of the shard lives in the symmetric
// heap.  Workers move only the rows a batch touches.
//
//   pull   out[i, :] = table[owner(idx[i])][idx[i] - owner*rows_per_rank, :]   (bf16, peer loads;
//          an index outside [0, vocab) yields a zero row — padding)
//   push   every worker w stages {count, idx[count], grad[count, dim]} in its own heap range;
//          between two device barriers (caller: psb_device_barrier) the owner runs
//          psb_sparse_apply: (1) mark — every staged index it owns gets a slot in a compact
//          accumulator (atomicCAS on slot[row], atomic counter), (2) accumulate — the staged
//          gradient rows are read from the peers (only the rows this rank owns cross NVLink)
//          and added in fp32 (float4 atomics: the order of a row's duplicates is not fixed, so
//          sums of >2 contributions can differ in the last bit between runs), (3) apply — one
//          warp per touched row: g * grad_scale through the same update as K1/K3
//          (SGD / momentum / Adam(W), *lazy*: rows without a gradient keep their state),
//          master + bf16 row written, accumulator row and slot cleared for the next step.
// dim % 8 == 0 (16-byte vectors).  acc holds world*cap rows of dim floats and starts zeroed,
// slot holds rows_per_rank int32 and starts at -1 (psb_sparse_init), n_touched starts at 0.
// ---------------------------------------------------------------------------
typedef struct PsbSparse {
  PsbFabric fabric;
  const void* table[PSB_MAX_PEERS];       // peer-mapped bf16 shard [rows_per_rank, dim]
  const void* stage_idx[PSB_MAX_PEERS];   // peer-mapped int64 [cap]
  const void* stage_grad[PSB_MAX_PEERS];  // peer-mapped [cap, dim] gradient rows (grad_dtype)
  const void* stage_cnt[PSB_MAX_PEERS];   // peer-mapped uint32: entries staged by that worker
  int64_t vocab;
  int64_t rows_per_rank;
  int32_t dim;
  int32_t cap;          // staging capacity per worker (entries)
  int32_t grad_dtype;   // PSB_F32 / PSB_BF16
  int32_t reserved;
  float* master;        // owner-local [rows_per_rank, dim]
  float* m;             // optional (momentum / Adam)
  float* v;             // optional (Adam)
  int32_t* slot;        // [rows_per_rank], -1 = untouched
  int32_t* touched;     // [world * cap] local rows that received a gradient this step
  uint32_t* n_touched;  // device word
  float* acc;           // [world * cap, dim] fp32 accumulator, zero between steps
  const uint32_t* step; // device word: completed optimizer steps (Adam bias correction)
  PsbOptim opt;
} PsbSparse;
int psb_sparse_init(int32_t* slot, int64_t rows, psb_stream_t stream);  // slot[:] = -1
int psb_sparse_pull(const PsbSparse* s, const int64_t* idx, int64_t n, void* out_bf16,
                    psb_stream_t stream);
// worker half of a push in one launch: ids, gradient rows (fp32 or bf16, cast to s->grad_dtype) and
// the count into this rank's staging range; n <= s->cap, grad 16-byte aligned
int psb_sparse_stage(const PsbSparse* s, const int64_t* idx, const void* grad, int grad_dtype,
                     int64_t n, psb_stream_t stream);
int psb_sparse_apply(const PsbSparse* s, psb_stream_t stream);

#ifdef __cplusplus
}
#endif
