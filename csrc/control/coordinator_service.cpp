// coordinator_service.cpp — see coordinator_service.h.
#include "coordinator_service.h"

#include <chrono>
#include <cstdlib>
#include <iostream>

#include "../wire/pb.h"
#include "../wire/pb_ext.h"

namespace psb {

using grpc_lite::Status;

CoordinatorService::CoordinatorService(const std::string& ps_host, int32_t ps_port,
                                       int scan_period_ms, int timeout_s)
    : core_(ps_host, ps_port), scan_period_ms_(scan_period_ms), timeout_s_(timeout_s) {
  const std::string base = "/coordinator.Coordinator/";
  server_.add(base + "RegisterWorker", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::WorkerInfo in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed WorkerInfo"};
    WorkerEntry e;
    e.worker_id = in.worker_id, e.address = in.address, e.port = in.port, e.hostname = in.hostname;
    pb::RegisterResponse out;
    out.success = core_.register_worker(e, out.parameter_server_address, out.total_workers);
    out.message = out.success ? "worker registered" : "registration failed";
    resp = out.encode();
    return Status{};
  });
  server_.add(base + "Heartbeat", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::HeartbeatRequest in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed HeartbeatRequest"};
    pb::HeartbeatResponse out;
    out.success = core_.update_heartbeat(in.worker_id, in.status);
    out.timestamp = std::chrono::duration_cast<std::chrono::seconds>(
                        std::chrono::system_clock::now().time_since_epoch())
                        .count();
    resp = out.encode();
    return Status{};
  });
  server_.add(base + "ListWorkers", [this](std::string_view, grpc_lite::Reply& resp) {
    pb::ListWorkersResponse out;
    for (const auto& w : core_.list_workers()) {
      pb::WorkerInfo wi;
      wi.worker_id = w.worker_id, wi.address = w.address, wi.port = w.port, wi.hostname = w.hostname;
      out.workers.push_back(std::move(wi));
    }
    out.total_workers = static_cast<int32_t>(out.workers.size());
    resp = out.encode();
    return Status{};
  });
  server_.add(base + "GetParameterServerAddress", [this](std::string_view, grpc_lite::Reply& resp) {
    pb::GetPSAddressResponse out;
    core_.get_parameter_server_address(out.address, out.port);
    resp = out.encode();
    return Status{};
  });

  // ---- extension service (proto/psb_ext.proto)
  const std::string ext = "/psb.CoordinatorExt/";
  server_.add(ext + "RegisterShard", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::ShardInfo in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed ShardInfo"};
    core_.register_shard({in.shard_id, in.address, in.device - 1});
    pb::Ack ack;
    ack.success = true;
    resp = ack.encode();
    return Status{};
  });
  server_.add(ext + "GetShardMap", [this](std::string_view, grpc_lite::Reply& resp) {
    pb::ShardMapResponse out;
    for (const auto& s : core_.shard_map()) {
      pb::ShardInfo si;
      si.shard_id = s.shard_id, si.address = s.address, si.device = s.device + 1;
      out.shards.push_back(std::move(si));
    }
    out.total_workers = static_cast<int32_t>(core_.num_workers());
    resp = out.encode();
    return Status{};
  });
  server_.add(ext + "Barrier", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::BarrierRequest in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed BarrierRequest"};
    pb::BarrierResponse out;
    int64_t to = in.timeout_ms > 0 ? in.timeout_ms : 10000;
    out.arrived = core_.barrier(in.name, in.worker_id, in.generation, in.expected, to, out.released);
    resp = out.encode();
    return Status{};
  });
  server_.add(ext + "UpdateClock", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::ClockRequest in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed ClockRequest"};
    pb::ClockResponse out;
    out.min_clock = core_.update_clock(in.worker_id, in.clock);
    resp = out.encode();
    return Status{};
  });
  server_.add(ext + "PutHandle", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::HandleBlob in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed HandleBlob"};
    core_.put_handle(in.key, in.rank, in.blob);
    pb::Ack ack;
    ack.success = true;
    resp = ack.encode();
    return Status{};
  });
  server_.add(ext + "GetHandles", [this](std::string_view req, grpc_lite::Reply& resp) {
    pb::HandleQuery in;
    if (!in.decode(req)) return Status{grpc_lite::INTERNAL, "malformed HandleQuery"};
    pb::HandleList out;
    for (auto& kv : core_.get_handles(in.key)) {
      pb::HandleBlob h;
      h.key = in.key, h.rank = kv.first, h.blob = std::move(kv.second);
      out.handles.push_back(std::move(h));
    }
    resp = out.encode();
    return Status{};
  });
  cleanup_thread_ = std::thread(&CoordinatorService::cleanup_loop, this);
}

CoordinatorService::~CoordinatorService() { shutdown(); }

void CoordinatorService::cleanup_loop() {
  int waited = 0;
  while (running_) {
    std::this_thread::sleep_for(std::chrono::milliseconds(50));
    waited += 50;
    if (waited < scan_period_ms_) continue;
    waited = 0;
    for (int32_t id : core_.remove_stale_workers(timeout_s_))
      std::cout << "evicted stale worker " << id << std::endl;
  }
}

int CoordinatorService::listen(const std::string& address) { return server_.listen(address); }
void CoordinatorService::serve_forever() { server_.serve_forever(); }
void CoordinatorService::start() { server_.start(); }

void CoordinatorService::shutdown() {
  if (running_.exchange(false)) {
    server_.shutdown();
    if (cleanup_thread_.joinable()) cleanup_thread_.join();
  }
}

void run_coordinator_server(const std::string& server_address, const std::string& ps_address,
                            int32_t ps_port) {
  int scan_ms = 10000, timeout_s = 30;
  if (const char* e = std::getenv("PSB_STALE_SCAN_MS")) scan_ms = std::atoi(e);
  if (const char* e = std::getenv("PSB_STALE_TIMEOUT_S")) timeout_s = std::atoi(e);
  CoordinatorService svc(ps_address, ps_port, scan_ms, timeout_s);
  const int bound = svc.listen(server_address);
  if (bound < 0) {
    std::cerr << "coordinator failed to listen on " << server_address << std::endl;
    std::exit(1);
  }
  std::string shown = server_address;  // port 0 = "pick one": show the port that was picked
  {
    std::string h;
    int p = -1;
    if (grpc_lite::split_host_port(server_address, h, p) && p == 0)
      shown = h + ":" + std::to_string(bound);
  }
  std::cout << "coordinator listening on " << shown << std::endl;
  std::cout << "parameter server: " << ps_address << ":" << ps_port << std::endl;
  svc.serve_forever();
}

}  // namespace psb
