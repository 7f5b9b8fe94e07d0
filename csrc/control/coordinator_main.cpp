// coordinator — argv-compatible with the reference binary
// (/root/reference/src/coordinator_main.cpp:5-24):
//   coordinator [listen=0.0.0.0:50052] [ps=host:port (localhost:50051)]
#include <string>

#include "coordinator_service.h"

int main(int argc, char** argv) {
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
