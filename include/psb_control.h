// psb_control.h — C ABI of libpsb_control.so, the native client of the coordinator.
//
// One handle per worker process.  It does what the reference worker does against the
// coordinator (/root/reference/src/worker.cpp:108-122 initialize, :141-157 discovery,
// :159-186 registration, :216-238 heartbeat thread) over the same wire protocol
// (proto/coordinator.proto), plus the extension service of this framework
// (proto/psb_ext.proto: shard map, named barriers, blob exchange) that the GPU data plane
// needs to come up: the ranks of a job find each other, exchange the handles of their
// symmetric-heap segments and synchronise on the host THROUGH THE COORDINATOR — no
// torch.distributed / NCCL process group is involved.
//
// The library has no CUDA dependency (it loads on a CPU-only box); Python binds it with ctypes
// (parameter_server_distributed_b200/control/client.py).
// All functions return 0 on success and a negative PSBC_E* code on failure unless stated
// otherwise; psbc_last_error() holds the message of the calling thread's last failure.
#pragma once

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PSBC_EINVAL (-1)
#define PSBC_EUNAVAILABLE (-2)  // coordinator not reachable / RPC failed
#define PSBC_ETIMEOUT (-3)      // barrier or gather did not complete in time
#define PSBC_ERANGE (-4)        // caller's buffer too small

typedef struct PsbCoord PsbCoord;

const char* psbc_last_error(void);

// Connect, discover the parameter server (GetParameterServerAddress), register this worker
// (RegisterWorker, exponential back-off like src/worker.cpp:129-139 until `connect_timeout_ms`
// has passed) and start the heartbeat thread (`heartbeat_ms`, reference: 5000).
// Returns NULL on failure.
PsbCoord* psbc_connect(const char* coordinator_address, int32_t worker_id, const char* self_address,
                       int32_t self_port, const char* hostname, int32_t heartbeat_ms,
                       int32_t connect_timeout_ms);
void psbc_close(PsbCoord* c);

// "host:port" of the parameter server as repaired from the coordinator's answers.
const char* psbc_ps_address(PsbCoord* c);
// WorkerStatus carried by the next heartbeats (0 IDLE, 1 TRAINING, 2 CHECKPOINTING, 3 ERROR).
int psbc_set_status(PsbCoord* c, int32_t status);

// ---- membership
// ids of the registered workers (ListWorkers); *count receives how many exist even when
// `cap` is smaller.
int psbc_list_workers(PsbCoord* c, int32_t* ids, int32_t cap, int32_t* count);
// Number of live workers the heartbeat thread saw last (-1 before its first poll).
int psbc_live_workers(PsbCoord* c);
// Times this worker found itself evicted (Heartbeat success=false) and re-registered.
int psbc_evictions(PsbCoord* c);
// Failure detection for the GPU plane.  From now on the heartbeat thread compares the live
// worker count with `expected_world` after every heartbeat; when it drops below (a peer was
// evicted) or this worker itself was evicted, the thread sets *abort_word = 1 (if non-NULL)
// and psbc_peer_lost() starts returning 1.  `abort_word` is typically the host-mapped word
// that the spin-waits of the kernels poll (psb_abort_word in libpsb_kernels), so kernels
// waiting on the dead peer's flags return instead of trapping.
int psbc_watch(PsbCoord* c, int32_t expected_world, volatile uint32_t* abort_word);
int psbc_peer_lost(PsbCoord* c);

// ---- shard map (psb.CoordinatorExt)
int psbc_register_shard(PsbCoord* c, int32_t shard_id, const char* address, int32_t device);
// Fills up to `cap` entries; addresses are written NUL-separated into addr_buf.
int psbc_shard_map(PsbCoord* c, int32_t* shard_ids, int32_t* devices, int32_t cap, char* addr_buf,
                   size_t addr_cap, int32_t* count, int32_t* total_workers);

// ---- host synchronisation and small-object exchange
// Named barrier over `expected` participants (0 = every live registered worker: an evicted
// worker stops blocking it).  Generations are counted per name inside the handle.
// Returns 0 when released, PSBC_ETIMEOUT otherwise.
int psbc_barrier(PsbCoord* c, const char* name, int32_t expected, int32_t timeout_ms);
// Every rank contributes `n` bytes under `key`; on return out[r*n .. r*n+n) holds rank r's
// contribution for r in [0, world).  Collective over ranks 0..world-1 (rank = worker_id).
int psbc_allgather(PsbCoord* c, const char* key, const void* blob, size_t n, int32_t world,
                   void* out, int32_t timeout_ms);
// Non-collective pieces of the above (a blob of length 0 erases the entry).
int psbc_put(PsbCoord* c, const char* key, const void* blob, size_t n);
// Fetches rank `rank`'s blob of `key` into out (cap bytes); *n receives its length, 0 = absent.
int psbc_get(PsbCoord* c, const char* key, int32_t rank, void* out, size_t cap, size_t* n);
// SSP clock (UpdateClock): publishes this worker's clock, returns the slowest live worker's.
int psbc_update_clock(PsbCoord* c, int32_t clock, int32_t* min_clock);

#ifdef __cplusplus
}
#endif
