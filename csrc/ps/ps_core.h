// ps_core.h — host-side parameter-server shard: staging, aggregation, update, checkpoints.
//
// Compat mode reproduces the reference's ParameterServerCore bit for bit
// (/root/reference/include/parameter_server.h:17-54, src/parameter_server.cpp:18-188):
// per-iteration staging by worker id, completion when #distinct pushers >= total_workers,
// element-wise mean, `p -= mean` (lr = 1), bootstrap-from-first-gradient,
// idempotent re-push, positional name+shape matching, App. B checkpoint bytes.
// Beyond the reference: lr / momentum / Adam(W) server updates, async and bounded-staleness
// (SSP) modes with versioned parameters, bounded memory (staged gradients are dropped after
// aggregation — fixes D5), atomic checkpoint writes plus a sidecar for optimizer state, one
// mutex discipline (fixes the races of D6), optional HBM residency through libpsb_kernels.
#pragma once

#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace psb {

struct HostTensor {
  std::string name;
  std::vector<int32_t> shape;
  std::vector<float> data;
  int32_t dtype = 0;
};

enum class PsMode { kSync, kAsync, kSsp };

struct PsOptions {
  PsMode mode = PsMode::kSync;
  int staleness = 0;            // SSP bound s: a worker may run at most s clocks ahead
  std::string optimizer = "sgd";  // sgd | momentum | adam | adamw
  float lr = 1.0f;              // compat default: p -= mean(g)
  float momentum = 0.9f;
  float beta1 = 0.9f, beta2 = 0.999f, eps = 1e-8f;
  float weight_decay = 0.0f;
  int keep_iterations = 1024;   // iteration states retained for status queries
  int device = -1;              // >= 0: keep master weights in HBM and update with K3
  int threads = 0;              // host aggregation / update threads (0 = hardware, capped at 16)
  // Status queries that would answer "not yet" are held this long for the answer to change
  // (CheckSyncStatus in sync mode, the staleness gate of ServeParameters in SSP mode), so a
  // waiting worker learns of the completion when it happens instead of at its next poll.
  // 0 = answer at once like the reference (src/parameter_server_service.cpp:85-96).
  int long_poll_ms = 25;
};

class DeviceShard;  // ps_device.cpp (dlopen of libpsb_kernels.so)

class ParameterServerCore {
 public:
  explicit ParameterServerCore(int total_workers, const PsOptions& opt = PsOptions());
  ~ParameterServerCore();

  void initialize_parameters(const std::vector<HostTensor>& initial);
  // Returns true when `iteration` is (now or already) aggregated.
  bool receive_gradients(int32_t worker_id, int32_t iteration, const std::vector<HostTensor>& grads);
  // The same, taking ownership: the service hands over the tensors it decoded, so staging a
  // push is a pointer move (the reference copies twice: proto -> tensor -> staging map,
  // src/parameter_server_service.cpp:35-42, src/parameter_server.cpp:32).
  bool receive_gradients(int32_t worker_id, int32_t iteration, std::vector<HostTensor>&& grads);
  std::vector<HostTensor> serve_parameters(int32_t iteration);
  // Runs `fn(parameters, version)` on the current model under the lock, without copying it
  // (the service encodes a pull response straight from the weights, once per version).
  void with_parameters(const std::function<void(const std::vector<HostTensor>&, int64_t)>& fn);
  bool check_sync_status(int32_t iteration, int32_t& workers_received);
  // The same answers, but a negative one is held up to options().long_poll_ms for the state
  // to change (an iteration aggregates, a clock advances, the membership changes).
  bool wait_sync_status(int32_t iteration, int32_t& workers_received);
  bool wait_staleness_ok(int32_t worker_id, int32_t iteration);
  // SSP / async: is a worker allowed to start clock `iteration`?
  bool staleness_ok(int32_t worker_id, int32_t iteration);
  bool save_checkpoint(int32_t epoch, const std::string& path);
  bool load_checkpoint(const std::string& path, int32_t& epoch);

  int get_total_workers() const { return total_workers_.load(); }
  // Follow the coordinator's registry: pending sync iterations that are complete for the new
  // count aggregate at once.  Returns how many iterations that completed.
  int set_total_workers(int n);
  int32_t get_current_iteration() const { return current_iteration_.load(); }
  int64_t version() const { return version_.load(); }
  const PsOptions& options() const { return opt_; }

 private:
  struct IterState;
  void apply_locked(const std::vector<HostTensor>& mean_grads);
  void aggregate_locked(IterState& st, int32_t shape_from);  // names / shapes of that worker's push
  void prune_locked();
  void update_tensor(size_t idx, const std::vector<float>& g);
  bool write_file(int32_t epoch, const std::string& path);

  std::atomic<int> total_workers_;
  PsOptions opt_;
  bool sync_status_locked(int32_t iteration, int32_t& workers_received);
  bool staleness_ok_locked(int32_t iteration);

  std::mutex mu_;  // guards everything below (single lock: no ordering hazards)
  std::condition_variable changed_;  // an iteration aggregated / a clock or the membership moved
  std::vector<HostTensor> params_;
  std::vector<std::vector<float>> m_, v_;
  int64_t opt_step_ = 0;
  struct IterState {
    std::map<int32_t, std::vector<HostTensor>> staged;  // ordered by worker id: fixed sum order
    std::map<int32_t, bool> seen;
    bool aggregated = false;
  };
  std::map<int32_t, IterState> iters_;
  std::map<int32_t, int32_t> clocks_;  // async / ssp: last clock pushed per worker
  std::atomic<int32_t> current_iteration_{0};
  std::atomic<int64_t> version_{0};
  DeviceShard* dev_ = nullptr;
};

}  // namespace psb
