// ps_service.cpp — see ps_service.h.
#include "ps_service.h"

#include <dirent.h>
#include <pthread.h>
#include <signal.h>

#include <chrono>
#include <cstdlib>
#include <iostream>
#include <sstream>

#include "../wire/pb.h"
#include "../wire/pb_ext.h"

namespace psb {

using grpc_lite::Status;

namespace {
HostTensor from_pb(pb::Tensor&& t) {
  HostTensor h;
  h.name = std::move(t.name);
  h.shape = std::move(t.shape);
  h.data = std::move(t.data);
  h.dtype = t.dtype;
  return h;
}
pb::Tensor to_pb(HostTensor&& h) {
  pb::Tensor t;
  t.name = std::move(h.name);
  t.shape = std::move(h.shape);
  t.data = std::move(h.data);
  t.dtype = h.dtype;
  return t;
}
}  // namespace

ParameterServerService::ParameterServerService(int total_workers, int checkpoint_interval,
                                               const PsOptions& opt, int checkpoint_scan_ms)
    : core_(total_workers, opt), sparse_(opt), checkpoint_interval_(checkpoint_interval),
      scan_ms_(checkpoint_scan_ms), configured_workers_(total_workers) {
  const std::string base = "/parameter_server.ParameterServer/";
  server_.add(base + "ReceiveGradients", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::GradientUpdate in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed GradientUpdate"};
    std::vector<HostTensor> grads;
    grads.reserve(in.gradients.size());
    for (auto& t : in.gradients) grads.push_back(from_pb(std::move(t)));  // one move, no copies
    pb::PushResponse out;
    const auto t0 = std::chrono::steady_clock::now();
    out.aggregation_complete =
        core_.receive_gradients(in.worker_id, in.iteration, std::move(grads));
    push_us_ += static_cast<uint64_t>(std::chrono::duration_cast<std::chrono::microseconds>(
                                          std::chrono::steady_clock::now() - t0).count());
    pushes_++;
    bytes_in_ += req.size();
    out.success = true;
    out.message = "gradients received";
    out.iteration = in.iteration;
    core_.check_sync_status(in.iteration, out.workers_received);
    out.total_workers = core_.get_total_workers();
    resp = out.encode();
    return Status{};
  });
  server_.add(base + "ServeParameters", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::PullRequest in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed PullRequest"};
    pb::ParameterUpdate out;
    out.iteration = in.iteration;
    // A negative iteration asks "where is the job?": the answer carries the newest iteration any
    // worker has pushed, so a worker that was restarted (or joins late) continues there instead
    // of at 0 (the reference's workers always start at 0, scripts/scale_workers.sh:113).
    if (in.iteration < 0) out.iteration = core_.get_current_iteration();
    if (core_.options().mode == PsMode::kSync) {
      int32_t received = 0;
      out.ready = core_.check_sync_status(in.iteration, received);
    } else {
      out.ready = core_.wait_staleness_ok(in.worker_id, in.iteration);
    }
    resp = out.encode();  // iteration + ready; the tensors follow from the shared encoding
    resp.shared_tail = encoded_parameters();
    pulls_++;
    bytes_out_ += resp.body.size() + resp.shared_tail->size();
    return Status{};
  });
  server_.add(base + "CheckSyncStatus", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::SyncStatusRequest in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed SyncStatusRequest"};
    pb::SyncStatusResponse out;
    out.iteration = in.iteration;
    out.ready = core_.wait_sync_status(in.iteration, out.workers_received);
    status_polls_++;
    out.total_workers = core_.get_total_workers();
    resp = out.encode();
    return Status{};
  });
  server_.add(base + "SaveCheckpoint", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::SaveCheckpointRequest in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed SaveCheckpointRequest"};
    std::string path = in.path;
    if (path.empty()) path = "checkpoint_epoch_" + std::to_string(in.epoch) + ".ckpt";
    pb::SaveCheckpointResponse out;
    out.success = core_.save_checkpoint(in.epoch, path);
    if (out.success && !sparse_.empty()) out.success = sparse_.save(path + ".tables");
    if (out.success) {
      out.message = "checkpoint saved";
      out.checkpoint_path = path;
    } else {
      out.message = "failed to save checkpoint";
    }
    resp = out.encode();
    return Status{};
  });
  server_.add(base + "LoadCheckpoint", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::LoadCheckpointRequest in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed LoadCheckpointRequest"};
    pb::LoadCheckpointResponse out;
    int32_t epoch = 0;
    out.success = core_.load_checkpoint(in.path, epoch);
    if (out.success) sparse_.load(in.path + ".tables");  // absent when the job has no tables
    if (out.success) {
      out.message = "checkpoint loaded";
      out.epoch = epoch;
      for (auto& t : core_.serve_parameters(0)) out.parameters.push_back(to_pb(std::move(t)));
    } else {
      out.message = "failed to load checkpoint";
    }
    resp = out.encode();
    return Status{};
  });
  // ---- sparse embedding tables (proto/psb_ext.proto: service psb.ParameterServerExt)
  const std::string ext = "/psb.ParameterServerExt/";
  server_.add(ext + "CreateTable", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::TableSpec in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed TableSpec"};
    pb::Ack out;
    out.success = sparse_.create(in.name, in.rows, in.dim, in.init_std, in.seed);
    resp = out.encode();
    return Status{};
  });
  server_.add(ext + "PushSparse", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::SparsePush in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed SparsePush"};
    std::vector<SparseGrad> grads;
    grads.reserve(in.tables.size());
    for (auto& t : in.tables) {
      SparseGrad g;
      g.name = std::move(t.name), g.dim = t.dim;
      g.indices = std::move(t.indices), g.values = std::move(t.values);
      grads.push_back(std::move(g));
    }
    pb::SparsePushResponse out;
    std::string err;
    sparse_pushes_++;
    for (const auto& g : grads) rows_pushed_ += g.indices.size();
    out.aggregation_complete = sparse_.push(in.worker_id, in.iteration, std::move(grads),
                                            core_.get_total_workers(), out.workers_received, err);
    out.success = err.empty();
    out.message = err.empty() ? "sparse gradients received" : err;
    out.total_workers = core_.get_total_workers();
    bytes_in_ += req.size();
    resp = out.encode();
    return Status{};
  });
  server_.add(ext + "PullRows", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::RowsRequest in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed RowsRequest"};
    pb::RowsResponse out;
    row_pulls_++;
    rows_pulled_ += in.indices.size();
    out.ready = sparse_.ready(in.iteration, core_.options().long_poll_ms);
    if (!sparse_.rows(in.name, in.indices, out.values, out.dim)) {
      out.ready = false;
      out.message = "unknown table '" + in.name + "' (or more than 2^30 values asked for at once)";
    }
    resp = out.encode();
    bytes_out_ += resp.body.size();
    return Status{};
  });
  if (checkpoint_interval_ > 0)
    checkpoint_thread_ = std::thread(&ParameterServerService::periodic_checkpoint, this);
}

ParameterServerService::~ParameterServerService() { shutdown(); }

// Same policy as the reference (src/parameter_server_service.cpp:150-169): every scan,
// epoch = current_iteration / interval; save when a new epoch started and iteration > 0.
std::shared_ptr<const std::string> ParameterServerService::encoded_parameters() {
  std::lock_guard<std::mutex> lk(serve_mu_);  // single flight: concurrent pulls wait for one encode
  if (serve_cache_ && serve_cache_version_ == core_.version()) return serve_cache_;
  auto bytes = std::make_shared<std::string>();
  int64_t version = -1;
  core_.with_parameters([&](const std::vector<HostTensor>& params, int64_t v) {
    version = v;
    size_t total = 0;
    std::vector<size_t> sizes(params.size());
    pb::Tensor shell;  // for the size / header arithmetic only; the floats are never copied into it
    for (size_t i = 0; i < params.size(); ++i) {
      shell.name = params[i].name;
      shell.shape = params[i].shape;
      shell.dtype = params[i].dtype;
      const size_t n = params[i].data.size();
      const size_t data_field = n ? 1 + pb::varint_size(n * 4) + n * 4 : 0;
      sizes[i] = shell.encoded_size() + data_field;
      total += 1 + pb::varint_size(sizes[i]) + sizes[i];
    }
    pb::Writer w;
    w.buf.reserve(total);
    for (size_t i = 0; i < params.size(); ++i) {
      w.tag(2, pb::kLen);
      w.varint(sizes[i]);
      w.str(1, params[i].name);
      w.packed_i32(2, params[i].shape);
      w.packed_f32(3, params[i].data.data(), params[i].data.size());
      w.i32(4, params[i].dtype);
    }
    *bytes = std::move(w.buf);
  });
  static const bool keep = [] {  // PSB_SERVE_CACHE=0: do not hold a second copy of the model
    const char* e = std::getenv("PSB_SERVE_CACHE");
    return !(e && e[0] == '0');
  }();
  if (!keep) return bytes;
  serve_cache_ = std::move(bytes);
  serve_cache_version_ = version;
  return serve_cache_;
}

void ParameterServerService::periodic_checkpoint() {
  int32_t last_epoch = -1;
  int waited = 0;
  while (running_) {
    std::this_thread::sleep_for(std::chrono::milliseconds(50));
    waited += 50;
    if (waited < scan_ms_) continue;
    waited = 0;
    const int32_t iter = core_.get_current_iteration();
    const int32_t epoch = iter / checkpoint_interval_;
    if (epoch > last_epoch && iter > 0) {
      const std::string path = "checkpoint_epoch_" + std::to_string(epoch) + ".ckpt";
      if (core_.save_checkpoint(epoch, path)) {
        if (!sparse_.empty()) sparse_.save(path + ".tables");
        last_epoch = epoch;
        std::cout << "saved checkpoint: " << path << " (epoch " << epoch << ")" << std::endl;
      }
    }
  }
}

void ParameterServerService::follow_membership(const std::string& coordinator_address,
                                               int poll_ms) {
  if (coordinator_address.empty() || membership_thread_.joinable()) return;
  membership_thread_ = std::thread(&ParameterServerService::membership_loop, this,
                                   coordinator_address, poll_ms > 0 ? poll_ms : 1000);
}

void ParameterServerService::membership_loop(std::string coordinator_address, int poll_ms) {
  grpc_lite::Channel coord(coordinator_address);
  bool seen_full = false;  // never shrink before every configured worker has registered once
  int waited = poll_ms;
  while (running_) {
    if (waited < poll_ms) {
      std::this_thread::sleep_for(std::chrono::milliseconds(20));
      waited += 20;
      continue;
    }
    waited = 0;
    std::string resp;
    Status s = coord.call("/coordinator.Coordinator/ListWorkers", std::string(), resp, 2000);
    pb::ListWorkersResponse lw;
    if (!s.ok() || !lw.decode(resp)) continue;  // coordinator away: keep the current count
    const int live = static_cast<int>(lw.workers.size());
    if (live >= configured_workers_) seen_full = true;
    int target = configured_workers_;
    if (seen_full && live >= 1 && live < configured_workers_) target = live;
    const int cur = core_.get_total_workers();
    if (target != cur) {
      const int completed = core_.set_total_workers(target) + sparse_.set_total_workers(target);
      std::cout << "membership: total_workers " << cur << " -> " << target << " (" << live
                << " live), " << completed << " pending iteration(s) completed" << std::endl;
    }
  }
}

void ParameterServerService::announce_shard(const std::string& coordinator_address, int shard_id,
                                            const std::string& advertise, int poll_ms) {
  if (coordinator_address.empty() || shard_id < 0 || shard_thread_.joinable()) return;
  shard_thread_ = std::thread([this, coordinator_address, shard_id, advertise, poll_ms] {
    grpc_lite::Channel coord(coordinator_address);
    pb::ShardInfo me;
    me.shard_id = shard_id;
    me.address = advertise;
    me.device = 0;  // host shard
    bool announced = false;
    int waited = poll_ms;
    while (running_) {
      if (waited < poll_ms) {
        std::this_thread::sleep_for(std::chrono::milliseconds(20));
        waited += 20;
        continue;
      }
      waited = 0;
      std::string resp;
      Status s = coord.call("/psb.CoordinatorExt/RegisterShard", me.encode(), resp, 2000);
      pb::Ack ack;
      const bool ok = s.ok() && ack.decode(resp) && ack.success;
      if (ok && !announced)
        std::cout << "registered as shard " << shard_id << " at " << advertise << std::endl;
      announced = announced || ok;  // repeated: a restarted coordinator learns the map again
    }
  });
}

std::string ParameterServerService::stats_json() {
  const uint64_t pushes = pushes_.load();
  std::ostringstream o;
  o << "{\"iteration\": " << core_.get_current_iteration() << ", \"version\": " << core_.version()
    << ", \"total_workers\": " << core_.get_total_workers() << ", \"pushes\": " << pushes
    << ", \"pulls\": " << pulls_.load() << ", \"status_polls\": " << status_polls_.load()
    << ", \"bytes_in\": " << bytes_in_.load() << ", \"bytes_out\": " << bytes_out_.load()
    << ", \"push_ms_mean\": " << (pushes ? push_us_.load() / 1000.0 / pushes : 0.0)
    << ", \"sparse_pushes\": " << sparse_pushes_.load() << ", \"rows_pushed\": " << rows_pushed_.load()
    << ", \"row_pulls\": " << row_pulls_.load() << ", \"rows_pulled\": " << rows_pulled_.load() << "}";
  return o.str();
}

void ParameterServerService::log_stats_every(int every_s) {
  if (every_s <= 0 || stats_thread_.joinable()) return;
  stats_thread_ = std::thread([this, every_s] {
    int waited = 0;
    while (running_) {
      std::this_thread::sleep_for(std::chrono::milliseconds(50));
      waited += 50;
      if (waited < every_s * 1000) continue;
      waited = 0;
      std::cout << "stats " << stats_json() << std::endl;
    }
  });
}

int ParameterServerService::listen(const std::string& address) { return server_.listen(address); }
void ParameterServerService::serve_forever() { server_.serve_forever(); }
void ParameterServerService::start() { server_.start(); }

void ParameterServerService::shutdown() {
  if (running_.exchange(false)) {
    server_.shutdown();
    if (checkpoint_thread_.joinable()) checkpoint_thread_.join();
    if (membership_thread_.joinable()) membership_thread_.join();
    if (shard_thread_.joinable()) shard_thread_.join();
    if (stats_thread_.joinable()) stats_thread_.join();
  }
}

std::string latest_checkpoint(const std::string& dir) {
  std::string best;
  long best_epoch = -1;
  DIR* d = ::opendir(dir.empty() ? "." : dir.c_str());
  if (!d) return best;
  const std::string pre = "checkpoint_epoch_", suf = ".ckpt";
  while (dirent* e = ::readdir(d)) {
    const std::string n = e->d_name;
    if (n.size() <= pre.size() + suf.size() || n.compare(0, pre.size(), pre) != 0 ||
        n.compare(n.size() - suf.size(), suf.size(), suf) != 0)
      continue;
    const std::string num = n.substr(pre.size(), n.size() - pre.size() - suf.size());
    if (num.find_first_not_of("0123456789") != std::string::npos) continue;
    const long epoch = std::strtol(num.c_str(), nullptr, 10);
    if (epoch > best_epoch) {
      best_epoch = epoch;
      best = (dir.empty() ? std::string() : dir + "/") + n;
    }
  }
  ::closedir(d);
  return best;
}

void run_server(const std::string& server_address, int total_workers, int checkpoint_interval,
                const PsOptions& opt, const std::string& coordinator_address,
                int membership_poll_ms, const std::string& resume, int shard_id,
                const std::string& advertise) {
  // SIGTERM / SIGINT are taken by one dedicated thread (below): blocked here, before any other
  // thread exists, so that every thread of the process inherits the mask.
  sigset_t stop_signals;
  sigemptyset(&stop_signals);
  sigaddset(&stop_signals, SIGTERM);
  sigaddset(&stop_signals, SIGINT);
  pthread_sigmask(SIG_BLOCK, &stop_signals, nullptr);
  int scan_ms = 5000;
  if (const char* e = std::getenv("PSB_CHECKPOINT_SCAN_MS")) scan_ms = std::atoi(e);
  ParameterServerService svc(total_workers, checkpoint_interval, opt, scan_ms);
  if (svc.listen(server_address) < 0) {
    std::cerr << "parameter server failed to listen on " << server_address << std::endl;
    std::exit(1);
  }
  std::cout << "parameter server listening on " << server_address << std::endl;
  if (!resume.empty()) {
    // A restarted server (supervisor, scale_workers.sh) picks its model up again by itself; the
    // reference comes back empty unless a worker passes CHECKPOINT_PATH
    // (/root/reference/scripts/scale_workers.sh:137-144, scripts/start_worker.sh:22-23).
    const std::string path = resume == "latest" ? latest_checkpoint("") : resume;
    int32_t epoch = 0;
    if (path.empty()) {
      std::cout << "resume: no checkpoint_epoch_*.ckpt here, starting empty" << std::endl;
    } else if (svc.core().load_checkpoint(path, epoch)) {
      svc.sparse().load(path + ".tables");
      std::cout << "resumed from " << path << " (epoch " << epoch << ", iteration "
                << svc.core().get_current_iteration() << ")" << std::endl;
    } else {
      std::cerr << "resume: failed to load " << path << std::endl;
      std::exit(1);
    }
  }
  if (checkpoint_interval > 0)
    std::cout << "periodic checkpointing every " << checkpoint_interval << " iterations"
              << std::endl;
  if (const char* e = std::getenv("PSB_STATS_EVERY_S")) svc.log_stats_every(std::atoi(e));
  if (shard_id >= 0) {
    if (coordinator_address.empty()) {
      std::cerr << "--shard-id needs --coordinator" << std::endl;
      std::exit(2);
    }
    std::string adv = advertise;
    if (adv.empty()) {  // the listen address, made dialable
      std::string host;
      int port = 0;
      grpc_lite::split_host_port(server_address, host, port);
      if (host.empty() || host == "0.0.0.0" || host == "[::]") host = "127.0.0.1";
      adv = host + ":" + std::to_string(port);
    }
    svc.announce_shard(coordinator_address, shard_id, adv, membership_poll_ms);
  }
  if (!coordinator_address.empty()) {
    svc.follow_membership(coordinator_address, membership_poll_ms);
    std::cout << "following worker membership of coordinator " << coordinator_address << std::endl;
  }
  // A planned stop (systemctl stop, scale_workers.sh, Ctrl-C) writes a last checkpoint where the
  // periodic ones go, so that --resume=latest continues from the very last update; the reference
  // loses everything since its last periodic file (src/parameter_server_service.cpp:150-169).
  std::thread stopper([&svc, stop_signals, checkpoint_interval] {
    int sig = 0;
    if (sigwait(&stop_signals, &sig) != 0) return;
    std::cout << "signal " << sig << ": shutting down" << std::endl;
    const int32_t iter = svc.core().get_current_iteration();
    if (checkpoint_interval > 0 && svc.core().version() > 0) {
      const int32_t epoch = iter / checkpoint_interval;
      const std::string path = "checkpoint_epoch_" + std::to_string(epoch) + ".ckpt";
      if (!svc.sparse().empty()) svc.sparse().save(path + ".tables");
      if (svc.core().save_checkpoint(epoch, path))
        std::cout << "saved checkpoint: " << path << " (epoch " << epoch << ")" << std::endl;
    }
    svc.shutdown();
  });
  svc.serve_forever();  // returns once the stopper has shut the service down
  stopper.join();
}

}  // namespace psb
