// coordinator — argv-compatible with the reference binary
// (/root/reference/src/coordinator_main.cpp:5-24):
//   coordinator [listen=0.0.0.0:50052] [ps=host:port (localhost:50051)]
#include <cstring>
#include <iostream>
#include <string>

#include "coordinator_service.h"

static const char kUsage[] =
    "usage: coordinator [listen=0.0.0.0:50052] [parameter_server=localhost:50051]\n"
    "  Worker registry, heartbeats and eviction (coordinator.Coordinator) plus shard map, named\n"
    "  barriers, SSP clocks and handle exchange (psb.CoordinatorExt) over gRPC.\n"
    "  environment: PSB_STALE_TIMEOUT_S (30)  evict a worker after this many seconds of silence\n"
    "               PSB_STALE_SCAN_MS (10000) how often the registry is scanned\n";

int main(int argc, char** argv) {
  if (argc > 1 && (!std::strcmp(argv[1], "--help") || !std::strcmp(argv[1], "-h"))) {
    std::cout << kUsage;
    return 0;
  }
  std::string listen = "0.0.0.0:50052";
  std::string ps_host = "localhost";
  int ps_port = 50051;
  if (argc > 1) listen = argv[1];
  if (argc > 2) {
    std::string ps = argv[2];
    if (!psb::grpc_lite::split_host_port(ps, ps_host, ps_port)) ps_host = ps;
  }
  psb::run_coordinator_server(listen, ps_host, ps_port);
  return 0;
}
