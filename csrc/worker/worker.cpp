// worker.cpp — see worker.h.  Reference counterpart: /root/reference/src/worker.cpp:69-406.
#include "worker.h"

#include <dlfcn.h>
#include <unistd.h>

#include <chrono>
#include <cmath>
#include <cstdlib>
#include <iostream>

#include "../../include/psb_kernels.h"
#include "../wire/pb.h"
#include "../wire/pb_ext.h"

namespace psb {

using grpc_lite::Channel;
using grpc_lite::Status;

namespace {
TensorLite from_pb(pb::Tensor&& p) {
  TensorLite t;
  t.name = std::move(p.name), t.shape = std::move(p.shape), t.data = std::move(p.data);
  t.dtype = p.dtype;
  return t;
}
void sleep_ms(int ms) { std::this_thread::sleep_for(std::chrono::milliseconds(ms)); }
}  // namespace

// Synthetic gradients on the GPU (kernel K7), through dlopen like the PS device shard.
struct Worker::DeviceFill {
  void* lib = nullptr;
  decltype(&psb_set_device) set_device = nullptr;
  decltype(&psb_heap_alloc) heap_alloc = nullptr;
  decltype(&psb_heap_free) heap_free = nullptr;
  decltype(&psb_fill_f32) fill = nullptr;
  decltype(&psb_memcpy_d2h) d2h = nullptr;
  decltype(&psb_stream_sync) sync = nullptr;
  int device = 0;
  float* buf = nullptr;
  size_t cap = 0;

  static DeviceFill* create(int device) {
    std::string path = "libpsb_kernels.so";
    if (const char* e = std::getenv("PSB_KERNELS_LIB")) path = e;
    else {
      char exe[4096];
      ssize_t n = readlink("/proc/self/exe", exe, sizeof(exe) - 1);
      if (n > 0) {
        exe[n] = 0;
        std::string p(exe);
        size_t cut = p.rfind("/build/bin/");
        if (cut != std::string::npos)
          path = p.substr(0, cut) + "/parameter_server_distributed_b200/_native/libpsb_kernels.so";
      }
    }
    void* lib = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!lib) return nullptr;
    auto* d = new DeviceFill;
    d->lib = lib;
    d->device = device;
#define PSB_SYM(field, name) d->field = reinterpret_cast<decltype(d->field)>(dlsym(lib, #name))
    PSB_SYM(set_device, psb_set_device);
    PSB_SYM(heap_alloc, psb_heap_alloc);
    PSB_SYM(heap_free, psb_heap_free);
    PSB_SYM(fill, psb_fill_f32);
    PSB_SYM(d2h, psb_memcpy_d2h);
    PSB_SYM(sync, psb_stream_sync);
#undef PSB_SYM
    if (!d->fill || d->set_device(device) != 0) {
      dlclose(lib);
      delete d;
      return nullptr;
    }
    return d;
  }
  bool run(std::vector<float>& out, float value) {
    if (out.empty()) return true;
    set_device(device);
    if (out.size() > cap) {
      if (buf) heap_free(buf);
      void* p = nullptr;
      if (heap_alloc(out.size() * sizeof(float), &p) != 0) return false;
      buf = static_cast<float*>(p);
      cap = out.size();
    }
    return fill(buf, out.size(), value, nullptr) == 0 &&
           d2h(out.data(), buf, out.size() * sizeof(float), nullptr) == 0 && sync(nullptr) == 0;
  }
  ~DeviceFill() {
    if (buf) heap_free(buf);
    if (lib) dlclose(lib);
  }
};

Worker::Worker(int worker_id, const std::string& coordinator_address,
               const std::string& worker_address, int32_t worker_port, const WorkerOptions& opt)
    : worker_id_(worker_id), coordinator_address_(coordinator_address),
      worker_address_(worker_address), worker_port_(worker_port), opt_(opt) {
  coord_.reset(new Channel(coordinator_address_));
  coord_hb_.reset(new Channel(coordinator_address_));
  sim_rng_ = 0x9E3779B97F4A7C15ull * static_cast<uint64_t>(worker_id + 1) + opt_.sim_seed;
  if (opt_.device >= 0) {
    dev_ = DeviceFill::create(opt_.device);
    if (!dev_) std::cerr << "worker " << worker_id_ << ": GPU fill unavailable, using host" << std::endl;
  }
  heartbeat_thread_ = std::thread(&Worker::heartbeat_loop, this);
}

Worker::~Worker() {
  {
    std::lock_guard<std::mutex> lk(hb_mu_);
    running_ = false;
  }
  hb_cv_.notify_all();
  if (heartbeat_thread_.joinable()) heartbeat_thread_.join();
  delete dev_;
}

bool Worker::initialize() {
  if (initialized_) return true;
  if (!discover_parameter_server()) return false;
  if (!register_with_coordinator()) return false;
  if (opt_.shards > 1 && !discover_shards()) return false;
  initialized_ = true;
  current_status_ = pb::IDLE;
  return true;
}

bool Worker::reconnect() {
  initialized_ = false;
  coord_.reset(new Channel(coordinator_address_));
  ps_.reset();
  return initialize();
}

// exponential back-off 100 * 2^k ms, 5 tries (reference: src/worker.cpp:129-139)
bool Worker::query_with_retry(const std::function<bool()>& query, int max_retries) {
  for (int attempt = 0; attempt < max_retries; ++attempt) {
    if (query()) return true;
    if (attempt + 1 < max_retries) sleep_ms(static_cast<int>(std::pow(2, attempt) * 100));
  }
  return false;
}

bool Worker::discover_parameter_server() {
  return query_with_retry([this]() {
    std::string resp;
    Status s = coord_->call("/coordinator.Coordinator/GetParameterServerAddress",
                            pb::GetPSAddressRequest().encode(), resp, 5000);
    pb::GetPSAddressResponse out;
    if (!s.ok() || !out.decode(resp)) return false;
    ps_port_hint_ = out.port;
    // tolerate an address that already carries a port ("localhost:50051" + 50051)
    std::string host;
    int port = 0;
    if (grpc_lite::split_host_port(out.address, host, port)) ps_address_ = out.address;
    else ps_address_ = out.address + ":" + std::to_string(out.port);
    return true;
  });
}

// Host-plane sharding: the coordinator's shard map (psb.CoordinatorExt/GetShardMap) names the
// parameter servers started with --shard-id; tensor k of the model lives on shard k % N.  Waits
// until all N have registered (they may still be starting).
bool Worker::discover_shards() {
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(30);
  while (std::chrono::steady_clock::now() < deadline) {
    std::string resp;
    Status s = coord_->call("/psb.CoordinatorExt/GetShardMap", std::string(), resp, 5000);
    pb::ShardMapResponse out;
    if (s.ok() && out.decode(resp)) {
      std::vector<std::string> addrs(static_cast<size_t>(opt_.shards));
      int found = 0;
      for (const auto& sh : out.shards) {
        if (sh.shard_id < 0 || sh.shard_id >= opt_.shards) continue;
        if (sh.address.rfind("hbm://", 0) == 0) continue;  // a GPU-plane shard, not a server
        if (addrs[static_cast<size_t>(sh.shard_id)].empty()) ++found;
        addrs[static_cast<size_t>(sh.shard_id)] = sh.address;
      }
      if (found == opt_.shards) {
        shard_addrs_ = std::move(addrs);
        shard_ch_.clear();
        for (const auto& a : shard_addrs_) shard_ch_.emplace_back(new Channel(a));
        return true;
      }
    }
    sleep_ms(100);
  }
  std::cerr << "worker " << worker_id_ << ": fewer than " << opt_.shards
            << " parameter-server shards registered with the coordinator" << std::endl;
  return false;
}

grpc_lite::Channel& Worker::shard_channel(int s) {
  return shard_addrs_.empty() ? ps_channel() : *shard_ch_[static_cast<size_t>(s)];
}

bool Worker::for_each_shard(const std::function<bool(int)>& fn) {
  const int n = num_shards();
  if (n == 1) return fn(0);
  std::vector<char> ok(static_cast<size_t>(n), 0);
  std::vector<std::thread> th;
  for (int s = 1; s < n; ++s) th.emplace_back([&, s] { ok[static_cast<size_t>(s)] = fn(s); });
  ok[0] = fn(0);
  for (auto& t : th) t.join();
  for (char c : ok)
    if (!c) return false;
  return true;
}

bool Worker::register_with_coordinator() {
  return query_with_retry([this]() {
    pb::WorkerInfo req;
    req.worker_id = worker_id_;
    req.address = worker_address_.empty() ? "localhost" : worker_address_;
    req.port = worker_port_;
    req.hostname = "worker-" + std::to_string(worker_id_);
    std::string resp;
    Status s = coord_->call("/coordinator.Coordinator/RegisterWorker", req.encode(), resp, 5000);
    pb::RegisterResponse out;
    if (!s.ok() || !out.decode(resp) || !out.success) return false;
    if (!out.parameter_server_address.empty()) {
      std::string host;
      int port = 0;
      if (grpc_lite::split_host_port(out.parameter_server_address, host, port)) {
        ps_address_ = out.parameter_server_address;
      } else if (ps_port_hint_ > 0) {
        // host-only answer (what the reference coordinator sends): keep the discovered port
        ps_address_ = out.parameter_server_address + ":" + std::to_string(ps_port_hint_);
      }
    }
    return true;
  });
}

std::vector<std::string> Worker::discover_peer_workers() {
  std::vector<std::string> peers;
  query_with_retry([this, &peers]() {
    std::string resp;
    Status s = coord_->call("/coordinator.Coordinator/ListWorkers",
                            pb::ListWorkersRequest().encode(), resp, 5000);
    pb::ListWorkersResponse out;
    if (!s.ok() || !out.decode(resp)) return false;
    peers.clear();
    for (const auto& w : out.workers)
      if (w.worker_id != worker_id_) peers.push_back(w.address + ":" + std::to_string(w.port));
    return true;
  });
  return peers;
}

void Worker::send_heartbeat() {
  if (!initialized_) return;
  pb::HeartbeatRequest req;
  req.worker_id = worker_id_;
  req.status = current_status_.load();
  std::string resp;
  Status s = coord_hb_->call("/coordinator.Coordinator/Heartbeat", req.encode(), resp, 3000);
  pb::HeartbeatResponse out;
  if (s.ok() && out.decode(resp) && !out.success) {
    // the coordinator evicted us (e.g. after a stall): announce ourselves again
    pb::WorkerInfo info;
    info.worker_id = worker_id_;
    info.address = worker_address_.empty() ? "localhost" : worker_address_;
    info.port = worker_port_;
    info.hostname = "worker-" + std::to_string(worker_id_);
    coord_hb_->call("/coordinator.Coordinator/RegisterWorker", info.encode(), resp, 3000);
  }
}

void Worker::heartbeat_loop() {
  std::unique_lock<std::mutex> lk(hb_mu_);
  while (running_) {
    hb_cv_.wait_for(lk, std::chrono::milliseconds(opt_.heartbeat_ms));
    if (!running_) break;
    lk.unlock();
    send_heartbeat();
    lk.lock();
  }
}

grpc_lite::Channel& Worker::ps_channel() {
  if (!ps_ || ps_channel_target_ != ps_address_) {
    ps_.reset(new Channel(ps_address_));
    ps_channel_target_ = ps_address_;
  }
  return *ps_;
}

std::vector<TensorLite> Worker::pull_parameters(int iteration, bool* ready) {
  pb::PullRequest req;
  req.worker_id = worker_id_;
  req.iteration = iteration;
  const std::string encoded = req.encode();
  const int n = num_shards();
  std::vector<std::vector<TensorLite>> parts(static_cast<size_t>(n));
  std::vector<char> part_ready(static_cast<size_t>(n), 0);
  const bool ok = for_each_shard([&](int s) {
    std::string resp;
    Status st = shard_channel(s).call("/parameter_server.ParameterServer/ServeParameters", encoded,
                                      resp, 30000);
    pb::ParameterUpdate out;
    if (!st.ok() || !out.decode(resp)) return false;
    part_ready[static_cast<size_t>(s)] = out.ready;
    auto& mine = parts[static_cast<size_t>(s)];
    mine.reserve(out.parameters.size());
    for (auto& t : out.parameters) mine.push_back(from_pb(std::move(t)));
    return true;
  });
  if (!ok) return {};
  if (ready) {
    *ready = true;
    for (char r : part_ready) *ready = *ready && r;
  }
  if (n == 1) return std::move(parts[0]);
  // tensor k lives on shard k % n as that shard's tensor k / n: interleave them back
  size_t total = 0;
  for (const auto& p : parts) total += p.size();
  std::vector<TensorLite> params;
  params.reserve(total);
  for (size_t k = 0; k < total; ++k) {
    auto& p = parts[k % static_cast<size_t>(n)];
    if (k / static_cast<size_t>(n) >= p.size()) return {};  // shards disagree about the model
    params.push_back(std::move(p[k / static_cast<size_t>(n)]));
  }
  return params;
}

bool Worker::push_gradients(int iteration, const std::vector<TensorLite>& grads,
                            int& workers_received, int& total_workers) {
  return push_gradients(iteration, std::vector<TensorLite>(grads), workers_received,
                        total_workers);
}

// Takes the gradients over: they go into the request without a copy, and the request is encoded
// into one buffer reserved once (the reference appends float by float, src/worker.cpp:40-52).
bool Worker::push_gradients(int iteration, std::vector<TensorLite>&& grads,
                            int& workers_received, int& total_workers) {
  const int n = num_shards();
  std::vector<pb::GradientUpdate> reqs(static_cast<size_t>(n));
  for (auto& r : reqs) {
    r.worker_id = worker_id_;
    r.iteration = iteration;
  }
  for (size_t k = 0; k < grads.size(); ++k) {
    TensorLite& g = grads[k];
    pb::Tensor p;
    p.name = std::move(g.name), p.shape = std::move(g.shape), p.data = std::move(g.data);
    p.dtype = g.dtype;
    reqs[k % static_cast<size_t>(n)].gradients.push_back(std::move(p));
  }
  std::vector<pb::PushResponse> outs(static_cast<size_t>(n));
  last_push_ok_ = for_each_shard([&](int s) {
    std::string resp;
    Status st = shard_channel(s).call("/parameter_server.ParameterServer/ReceiveGradients",
                                      reqs[static_cast<size_t>(s)].encode(), resp, 30000);
    return st.ok() && outs[static_cast<size_t>(s)].decode(resp);
  });
  if (!last_push_ok_) return false;
  bool complete = true;
  workers_received = outs[0].workers_received;
  total_workers = outs[0].total_workers;
  for (const auto& o : outs) {  // the iteration is done when every shard has aggregated it
    complete = complete && o.aggregation_complete;
    workers_received = std::min(workers_received, o.workers_received);
  }
  return complete;
}

int Worker::server_iteration() {
  pb::PullRequest req;
  req.worker_id = worker_id_;
  req.iteration = -1;
  std::string resp;
  Status st = shard_channel(0).call("/parameter_server.ParameterServer/ServeParameters",
                                    req.encode(), resp, 30000);
  pb::ParameterUpdate out;
  if (!st.ok() || !out.decode(resp)) return -1;
  return out.iteration;
}

bool Worker::check_sync_ready(int iteration, int& workers_received, int& total_workers) {
  pb::SyncStatusRequest req;
  req.iteration = iteration;
  const std::string encoded = req.encode();
  const int n = num_shards();
  bool ready = true;
  for (int s = 0; s < n && ready; ++s) {  // each answer is long-polled by its server
    std::string resp;
    Status st = shard_channel(s).call("/parameter_server.ParameterServer/CheckSyncStatus", encoded,
                                      resp, 5000);
    pb::SyncStatusResponse out;
    if (!st.ok() || !out.decode(resp)) return false;
    workers_received = s == 0 ? out.workers_received
                              : std::min(workers_received, out.workers_received);
    total_workers = out.total_workers;
    ready = out.ready;
  }
  return ready;
}

bool Worker::create_table(const std::string& name, int64_t rows, int32_t dim, float init_std,
                          int64_t seed) {
  pb::TableSpec req;
  req.name = name, req.rows = rows, req.dim = dim, req.init_std = init_std, req.seed = seed;
  std::string resp;
  Status st = shard_channel(0).call("/psb.ParameterServerExt/CreateTable", req.encode(), resp, 10000);
  pb::Ack out;
  return st.ok() && out.decode(resp) && out.success;
}

bool Worker::push_sparse(int iteration, const std::string& name, int32_t dim,
                         const std::vector<int64_t>& indices, const std::vector<float>& values,
                         bool* complete) {
  pb::SparsePush req;
  req.worker_id = worker_id_, req.iteration = iteration;
  pb::SparseRows rows;
  rows.name = name, rows.dim = dim, rows.indices = indices, rows.values = values;
  req.tables.push_back(std::move(rows));
  std::string resp;
  Status st = shard_channel(0).call("/psb.ParameterServerExt/PushSparse", req.encode(), resp, 10000);
  pb::SparsePushResponse out;
  if (!st.ok() || !out.decode(resp) || !out.success) return false;
  if (complete) *complete = out.aggregation_complete;
  return true;
}

bool Worker::pull_rows(const std::string& name, const std::vector<int64_t>& indices,
                       int after_iteration, std::vector<float>& out, int timeout_ms) {
  pb::RowsRequest req;
  req.worker_id = worker_id_, req.iteration = after_iteration, req.name = name;
  req.indices = indices;
  const std::string encoded = req.encode();
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  do {  // a "not yet" answer was already held by the server for its long-poll time
    std::string resp;
    Status st = shard_channel(0).call("/psb.ParameterServerExt/PullRows", encoded, resp, 10000);
    pb::RowsResponse r;
    if (!st.ok() || !r.decode(resp) || !r.message.empty()) return false;
    if (r.ready) {
      out = std::move(r.values);
      return true;
    }
  } while (std::chrono::steady_clock::now() < deadline);
  return false;
}

bool Worker::load_checkpoint_from_server(const std::string& checkpoint_path, int32_t& epoch) {
  if (!initialized_ && ps_address_.empty() && !discover_parameter_server()) return false;
  if (opt_.shards > 1 && shard_addrs_.empty() && !discover_shards()) return false;
  pb::LoadCheckpointRequest req;
  req.path = checkpoint_path;  // every shard resolves it in its own working directory
  const std::string encoded = req.encode();
  std::vector<int32_t> epochs(static_cast<size_t>(num_shards()), 0);
  const bool ok = for_each_shard([&](int sh) {
    std::string resp;
    Status s = shard_channel(sh).call("/parameter_server.ParameterServer/LoadCheckpoint", encoded,
                                      resp, 60000);
    pb::LoadCheckpointResponse out;
    if (!s.ok() || !out.decode(resp) || !out.success) return false;
    epochs[static_cast<size_t>(sh)] = out.epoch;
    return true;
  });
  if (!ok) return false;
  epoch = epochs[0];  // the parameters now live in the servers; the next pull returns them
  return true;
}

std::vector<TensorLite> Worker::compute_gradients(std::vector<TensorLite>&& params) {
  if (opt_.compute_ms > 0.0) {
    // xorshift per worker: the same run sees the same stragglers in every mode
    sim_rng_ ^= sim_rng_ << 13;
    sim_rng_ ^= sim_rng_ >> 7;
    sim_rng_ ^= sim_rng_ << 17;
    const double u = static_cast<double>(sim_rng_ >> 11) / 9007199254740992.0;
    const double ms = opt_.compute_ms * (u < opt_.straggle_prob ? opt_.straggle_factor : 1.0);
    // sleeping, not spinning: the host thread of a GPU worker is idle while the device works
    std::this_thread::sleep_for(std::chrono::microseconds(static_cast<int64_t>(ms * 1000.0)));
  }
  if (grad_fn_) return grad_fn_(params);
  std::vector<TensorLite> grads = std::move(params);  // same names and shapes: filled in place
  for (auto& t : grads) {
    if (dev_ && dev_->run(t.data, opt_.synthetic_grad)) continue;
    for (auto& v : t.data) v = opt_.synthetic_grad;
  }
  return grads;
}

bool Worker::run_iteration(int iteration) {
  current_status_ = pb::TRAINING;
  using clk = std::chrono::steady_clock;
  auto since = [](clk::time_point t) {
    return std::chrono::duration<double, std::milli>(clk::now() - t).count();
  };
  for (int attempt = 0; attempt < opt_.max_retries; ++attempt) {
    bool ready = true;
    const auto t_pull = clk::now();
    std::vector<TensorLite> params = pull_parameters(iteration, &ready);
    if (opt_.mode == "ssp") {
      // bounded staleness: wait until the slowest worker is within the bound
      const auto deadline = std::chrono::steady_clock::now() +
                            std::chrono::milliseconds(opt_.sync_timeout_ms);
      int backoff = 1;
      while (!ready && std::chrono::steady_clock::now() < deadline) {
        sleep_ms(backoff);
        backoff = std::min(backoff * 2, 50);
        params = pull_parameters(iteration, &ready);
      }
    }
    if (params.empty()) {
      // empty server: same fabricated model as the reference (src/worker.cpp:346-353)
      TensorLite w;
      w.name = "weight";
      w.shape = {10, 10};
      w.dtype = 0;
      w.data.assign(100, 0.0f);
      params.push_back(std::move(w));
    }
    const double pull_ms = since(t_pull);
    const auto t_compute = clk::now();
    std::vector<TensorLite> grads = compute_gradients(std::move(params));
    const double compute_ms = since(t_compute);
    int received = 0, total = 0;
    const auto t_push = clk::now();
    bool complete = push_gradients(iteration, std::move(grads), received, total);
    const double push_ms = since(t_push);
    const auto t_wait = clk::now();
    auto trace = [&](bool done) {
      if (!opt_.trace) return;
      std::cout << "worker " << worker_id_ << " trace iter " << iteration << " pull_ms=" << pull_ms
                << " compute_ms=" << compute_ms << " push_ms=" << push_ms
                << " wait_ms=" << since(t_wait) << " done=" << (done ? "true" : "false")
                << std::endl;
    };
    if (complete) {
      current_status_ = pb::IDLE;
      trace(true);
      return true;
    }
    if (!last_push_ok_) {  // transport failure: back off and try the whole step again
      if (attempt + 1 < opt_.max_retries) sleep_ms(500);
      continue;
    }
    // not the last pusher: poll the barrier with a short adaptive back-off
    const auto deadline =
        std::chrono::steady_clock::now() + std::chrono::milliseconds(opt_.sync_timeout_ms);
    int backoff = 1;
    while (std::chrono::steady_clock::now() < deadline) {
      if (check_sync_ready(iteration, received, total)) {
        current_status_ = pb::IDLE;
        trace(true);
        return true;
      }
      sleep_ms(backoff);
      backoff = std::min(backoff * 2, 50);
    }
    break;  // the barrier did not open in time: give up on this iteration like the reference
  }
  current_status_ = pb::IDLE;
  return false;
}

}  // namespace psb
