// worker.h — worker client runtime and training state machine.
//
// Same public surface as the reference Worker (/root/reference/include/worker.h:21-69):
// initialize / reconnect / run_iteration / load_checkpoint_from_server, a heartbeat thread,
// discovery through the coordinator.  Differences by design:
//  * one persistent channel per peer instead of a new TCP + HTTP/2 handshake per RPC (D7);
//  * no fixed 500 ms sleeps: a push that does not complete the iteration is followed by an
//    adaptive CheckSyncStatus poll (1 ms doubling to 50 ms), never by a re-pull + re-push;
//  * `initialized_` is atomic, the heartbeat thread wakes up immediately on shutdown (D6);
//  * a host-only parameter-server address in RegisterResponse is repaired with the port
//    learnt from GetParameterServerAddress (tolerates the reference coordinator, D1);
//  * an evicted worker (Heartbeat success=false) re-registers (D9);
//  * optional --device: synthetic gradients are produced on the GPU by kernel K7.
#pragma once

#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../wire/h2.h"

namespace psb {

struct TensorLite {
  std::string name;
  std::vector<int32_t> shape;
  std::vector<float> data;
  int32_t dtype = 0;  // 0 = float32, 1 = float64
};

struct WorkerOptions {
  std::string mode = "sync";      // sync | async | ssp
  int heartbeat_ms = 5000;        // reference: 5 s (src/worker.cpp:233)
  int sync_timeout_ms = 10000;    // reference: 200 polls x 50 ms (src/worker.cpp:372-375)
  int max_retries = 3;            // reference: 3 (src/worker.cpp:335)
  int device = -1;                // >= 0: synthetic gradients filled on this GPU (K7)
  float synthetic_grad = 0.01f;   // reference: 0.01f (src/worker.cpp:316-320)
  // Simulated compute time of one iteration (benchmarks/straggler_host.sh): the gradient step
  // takes compute_ms, and with probability straggle_prob an iteration takes straggle_factor
  // times as long (a transient straggler; drawn from a per-worker deterministic sequence).
  double compute_ms = 0.0;
  double straggle_prob = 0.0;
  double straggle_factor = 1.0;
  uint64_t sim_seed = 0;
  int shards = 0;                 // > 1: the model's tensors are split over this many PS shards
  bool trace = false;             // per-iteration pull / compute / push / wait times on stdout
};

class Worker {
 public:
  Worker(int worker_id, const std::string& coordinator_address,
         const std::string& worker_address = "", int32_t worker_port = 0,
         const WorkerOptions& opt = WorkerOptions());
  ~Worker();

  bool initialize();
  bool reconnect();
  // One synchronous iteration: pull -> compute -> push -> wait until aggregated.
  bool run_iteration(int iteration);
  bool load_checkpoint_from_server(const std::string& checkpoint_path, int32_t& epoch);

  // Lower-level pieces (public for tests and for Python-driven workers).
  std::vector<TensorLite> pull_parameters(int iteration, bool* ready = nullptr);
  bool push_gradients(int iteration, const std::vector<TensorLite>& grads, int& workers_received,
                      int& total_workers);
  bool push_gradients(int iteration, std::vector<TensorLite>&& grads, int& workers_received,
                      int& total_workers);
  bool check_sync_ready(int iteration, int& workers_received, int& total_workers);
  // The newest iteration any worker has pushed to the server (shard 0), -1 if it cannot be
  // asked: where a restarted or late worker joins the job (worker_main --join).
  int server_iteration();
  // Sparse embedding tables (psb.ParameterServerExt; no counterpart in the reference, whose
  // Tensor is dense).  Single parameter server: the table lives on the server this worker
  // pushes to (shard 0 of a sharded job).
  bool create_table(const std::string& name, int64_t rows, int32_t dim, float init_std = 0.0f,
                    int64_t seed = 0);
  // One (rows, gradients) push for `iteration`; *complete = every live worker has pushed.
  bool push_sparse(int iteration, const std::string& name, int32_t dim,
                   const std::vector<int64_t>& indices, const std::vector<float>& values,
                   bool* complete = nullptr);
  // Rows once the sparse pushes of `after_iteration` are applied (-1: at once).
  bool pull_rows(const std::string& name, const std::vector<int64_t>& indices, int after_iteration,
                 std::vector<float>& out, int timeout_ms = 10000);
  std::vector<std::string> discover_peer_workers();
  const std::string& ps_address() const { return ps_address_; }
  // Replace the synthetic gradient source (default: every element = 0.01).
  void set_gradient_fn(std::function<std::vector<TensorLite>(const std::vector<TensorLite>&)> fn) {
    grad_fn_ = std::move(fn);
  }

 private:
  bool discover_parameter_server();
  bool discover_shards();
  int num_shards() const { return shard_addrs_.empty() ? 1 : static_cast<int>(shard_addrs_.size()); }
  grpc_lite::Channel& shard_channel(int s);
  // run fn(shard) for every shard (in parallel when there are several); true iff all returned true
  bool for_each_shard(const std::function<bool(int)>& fn);
  bool register_with_coordinator();
  bool query_with_retry(const std::function<bool()>& query, int max_retries = 5);
  void send_heartbeat();
  void heartbeat_loop();
  std::vector<TensorLite> compute_gradients(std::vector<TensorLite>&& params);
  grpc_lite::Channel& ps_channel();

  int worker_id_;
  std::string coordinator_address_;
  std::string ps_address_;
  int ps_port_hint_ = 0;
  std::string worker_address_;
  int32_t worker_port_;
  WorkerOptions opt_;
  std::atomic<bool> initialized_{false};
  std::atomic<bool> running_{true};
  std::atomic<int32_t> current_status_{0};
  std::thread heartbeat_thread_;
  std::mutex hb_mu_;
  std::condition_variable hb_cv_;
  std::unique_ptr<grpc_lite::Channel> coord_;
  std::unique_ptr<grpc_lite::Channel> coord_hb_;  // heartbeats use their own connection
  std::unique_ptr<grpc_lite::Channel> ps_;
  std::string ps_channel_target_;
  std::vector<std::string> shard_addrs_;                      // by shard id; empty = single server
  std::vector<std::unique_ptr<grpc_lite::Channel>> shard_ch_;
  bool last_push_ok_ = false;
  uint64_t sim_rng_ = 0;
  std::function<std::vector<TensorLite>(const std::vector<TensorLite>&)> grad_fn_;
  struct DeviceFill;
  DeviceFill* dev_ = nullptr;
};

}  // namespace psb
