// ps_service.h — gRPC adapter of ParameterServerCore + periodic checkpoint thread.
// Counterpart of /root/reference/src/parameter_server_service.cpp:15-191.
#pragma once

#include <atomic>
#include <string>
#include <memory>
#include <mutex>
#include <thread>

#include "../wire/h2.h"
#include "ps_core.h"
#include "ps_sparse.h"

namespace psb {

class ParameterServerService {
 public:
  ParameterServerService(int total_workers, int checkpoint_interval, const PsOptions& opt,
                         int checkpoint_scan_ms = 5000);
  // Follow the coordinator's worker registry (ListWorkers every `poll_ms`): once the job has
  // been complete, an eviction shrinks the sync barrier to the live workers and pending
  // iterations complete with their mean; a worker that re-registers grows it back (never above
  // the configured count).  The reference's coordinator and parameter server never talk
  // (SURVEY D9); this uses the reference's own ListWorkers RPC, so it also works against a
  // reference coordinator.
  void follow_membership(const std::string& coordinator_address, int poll_ms = 1000);
  // Host-plane sharding: announce this process as shard `shard_id` at `advertise` (dialable
  // host:port) in the coordinator's shard map (psb.CoordinatorExt/RegisterShard) and keep the
  // entry fresh; workers started with --shards=N split the model's tensors over the N shards
  // they find there.  The reference has exactly one parameter server
  // (src/coordinator_main.cpp:9-21 knows a single address).
  void announce_shard(const std::string& coordinator_address, int shard_id,
                      const std::string& advertise, int poll_ms = 1000);
  ~ParameterServerService();
  int listen(const std::string& address);
  void serve_forever();
  void start();
  void shutdown();
  ParameterServerCore& core() { return core_; }
  // Sparse embedding tables (service psb.ParameterServerExt: CreateTable / PushSparse /
  // PullRows); they follow the same membership and are saved next to every checkpoint
  // (`<path>.tables`).  No counterpart in the reference.
  SparseTables& sparse() { return sparse_; }
  // Observability (the reference prints nothing but its start-up lines): one JSON
  // line on stdout every `every_s` seconds — iteration, model version, live worker count, RPC
  // and byte counters, mean time of a push inside the core.  --stats-every-s / PSB_STATS_EVERY_S.
  void log_stats_every(int every_s);
  std::string stats_json();

 private:
  void periodic_checkpoint();
  void membership_loop(std::string coordinator_address, int poll_ms);
  // `repeated Tensor parameters = 2` of ParameterUpdate, encoded for the current model version.
  // In sync mode every worker pulls the same weights between two updates: they are encoded
  // once (straight from the core's tensors) and every response is sent from this buffer.  The
  // reference converts float by float into a fresh proto for each pull
  // (src/parameter_server_service.cpp:61-83).
  std::shared_ptr<const std::string> encoded_parameters();
  std::mutex serve_mu_;
  std::shared_ptr<const std::string> serve_cache_;
  int64_t serve_cache_version_ = -1;
  ParameterServerCore core_;
  SparseTables sparse_;
  grpc_lite::Server server_;
  int checkpoint_interval_;
  int scan_ms_;
  std::thread checkpoint_thread_;
  std::thread membership_thread_;
  std::thread shard_thread_;
  std::thread stats_thread_;
  std::atomic<uint64_t> pushes_{0}, pulls_{0}, status_polls_{0}, bytes_in_{0}, bytes_out_{0},
      push_us_{0}, sparse_pushes_{0}, row_pulls_{0}, rows_pushed_{0}, rows_pulled_{0};
  int configured_workers_;
  std::atomic<bool> running_{true};
};

// Blocking entry point with the reference's signature and log lines
// (include/parameter_server_service.h:5, src/parameter_server_service.cpp:177-191).
void run_server(const std::string& server_address, int total_workers, int checkpoint_interval,
                const PsOptions& opt = PsOptions(), const std::string& coordinator_address = "",
                int membership_poll_ms = 1000, const std::string& resume = "", int shard_id = -1,
                const std::string& advertise = "");

// Newest periodic checkpoint ("checkpoint_epoch_<E>.ckpt", largest E) in `dir`; "" if none.
std::string latest_checkpoint(const std::string& dir);

}  // namespace psb
