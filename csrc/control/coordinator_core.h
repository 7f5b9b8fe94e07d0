// coordinator_core.h — worker registry, liveness, barriers and shard map (no I/O).
//
// Capabilities of the reference's CoordinatorCore (/root/reference/include/coordinator.h:19-38,
// src/coordinator.cpp:7-67): register (overwrite allowed), heartbeat stamp + status, list,
// PS address, eviction of workers silent for longer than a timeout.  Added for the
// north-star: a shard map (one PS shard per GPU), named barriers whose
// membership follows the registry (an evicted worker no longer blocks a barrier — fixes
// SURVEY D9), SSP clocks and an opaque handle exchange for the symmetric heap.
#pragma once

#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace psb {

struct WorkerEntry {
  int32_t worker_id = 0;
  std::string address;
  int32_t port = 0;
  std::string hostname;
  int32_t status = 0;
  std::chrono::steady_clock::time_point last_heartbeat;
};

struct ShardEntry {
  int32_t shard_id = 0;
  std::string address;  // dialable host:port of the shard's gRPC service
  int32_t device = -1;  // GPU ordinal on the box, -1 = host shard
};

class CoordinatorCore {
 public:
  CoordinatorCore(const std::string& ps_host, int32_t ps_port);

  // Returns the dialable "host:port" of the (first) parameter server — the reference returns
  // the host part only and breaks its own workers.
  bool register_worker(const WorkerEntry& w, std::string& ps_address, int32_t& total_workers);
  bool update_heartbeat(int32_t worker_id, int32_t status);
  std::vector<WorkerEntry> list_workers();
  void get_parameter_server_address(std::string& host, int32_t& port);
  // Removes workers silent for more than `timeout_seconds`; returns the evicted ids.
  std::vector<int32_t> remove_stale_workers(int64_t timeout_seconds = 30);
  size_t num_workers();

  // ---- extensions
  void register_shard(const ShardEntry& s);
  std::vector<ShardEntry> shard_map();
  // Blocks until `expected` participants (0 = every registered worker) arrived at barrier
  // `name` for generation `generation`, the timeout expires, or membership shrank below the
  // arrivals.  Returns the number of arrivals seen; `released` tells whether it opened.
  int32_t barrier(const std::string& name, int32_t worker_id, int32_t generation, int32_t expected,
                  int64_t timeout_ms, bool& released);
  // SSP clocks: worker publishes its clock; returns the minimum over live workers.
  int32_t update_clock(int32_t worker_id, int32_t clock);
  // Opaque blobs keyed by (key, rank): the symmetric-heap IPC handles of the C++ binaries.
  void put_handle(const std::string& key, int32_t rank, const std::string& blob);
  std::vector<std::pair<int32_t, std::string>> get_handles(const std::string& key);

 private:
  std::string ps_host_;
  int32_t ps_port_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::map<int32_t, WorkerEntry> workers_;
  std::map<int32_t, ShardEntry> shards_;
  struct BarrierState {
    int32_t generation = -1;
    std::map<int32_t, bool> arrived;
    bool open = false;
  };
  std::map<std::string, BarrierState> barriers_;
  std::map<int32_t, int32_t> clocks_;
  std::map<std::string, std::map<int32_t, std::string>> handles_;
};

}  // namespace psb
