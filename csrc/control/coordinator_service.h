// coordinator_service.h — gRPC adapter of CoordinatorCore + stale-worker cleanup thread.
// Counterpart of /root/reference/src/coordinator_service.cpp:26-126.
#pragma once

#include <atomic>
#include <memory>
#include <string>
#include <thread>

#include "../wire/h2.h"
#include "coordinator_core.h"

namespace psb {

class CoordinatorService {
 public:
  // scan_period_s / timeout_s default to the reference's 10 s / 30 s
  // (src/coordinator_service.cpp:104-105); tests shorten them.
  CoordinatorService(const std::string& ps_host, int32_t ps_port, int scan_period_ms = 10000,
                     int timeout_s = 30);
  ~CoordinatorService();
  int listen(const std::string& address);  // returns the bound port or -1
  void serve_forever();
  void start();
  void shutdown();
  CoordinatorCore& core() { return core_; }

 private:
  void cleanup_loop();
  CoordinatorCore core_;
  grpc_lite::Server server_;
  std::thread cleanup_thread_;
  std::atomic<bool> running_{true};
  int scan_period_ms_;
  int timeout_s_;
};

// Blocking entry point with the reference's signature and log lines
// (include/coordinator_service.h:5, src/coordinator_service.cpp:114-126).
void run_coordinator_server(const std::string& server_address, const std::string& ps_address,
                            int32_t ps_port);

}  // namespace psb
