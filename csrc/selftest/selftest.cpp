// psb_selftest — in-process smoke test of the C++ control plane (no Python, no GPU):
// coordinator + parameter server + two workers over real sockets on 127.0.0.1, checked
// against the golden behavioural trace and the checkpoint bytes (App. B).
// Also the program the TSAN / ASAN builds run (`make tsan && build/bin-thread/psb_selftest`).
#include <cmath>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <thread>

#include "../wire/hpack.h"
#include "../control/coordinator_service.h"
#include "../ps/ps_service.h"
#include "../worker/worker.h"

#include <string>
#include <vector>

#define CHECK(cond)                                                              \
  do {                                                                           \
    if (!(cond)) {                                                               \
      std::cerr << "FAILED " #cond " at " << __FILE__ << ":" << __LINE__ << "\n"; \
      return 1;                                                                  \
    }                                                                            \
  } while (0)

static std::string unhex(const char* h) {
  std::string out;
  for (; h[0] && h[1]; h += 2) out.push_back(static_cast<char>(std::stoi(std::string(h, 2), nullptr, 16)));
  return out;
}

// RFC 7541 Appendix C.4: three consecutive requests with Huffman strings and a dynamic table
// that carries over from one header block to the next.
static int hpack_rfc_vectors() {
  using psb::hpack::Header;
  psb::hpack::Decoder dec;
  std::vector<Header> h;
  CHECK(dec.decode(unhex("828684418cf1e3c2e5f23a6ba0ab90f4ff"), h));
  CHECK(h.size() == 4 && h[0] == Header(":method", "GET") && h[1] == Header(":scheme", "http") &&
        h[2] == Header(":path", "/") && h[3] == Header(":authority", "www.example.com"));
  h.clear();
  CHECK(dec.decode(unhex("828684be5886a8eb10649cbf"), h));
  CHECK(h.size() == 5 && h[3] == Header(":authority", "www.example.com") &&
        h[4] == Header("cache-control", "no-cache"));
  h.clear();
  CHECK(dec.decode(unhex("828785bf408825a849e95ba97d7f8925a849e95bb8e8b4bf"), h));
  CHECK(h.size() == 5 && h[1] == Header(":scheme", "https") && h[2] == Header(":path", "/index.html") &&
        h[4] == Header("custom-key", "custom-value"));
  // what our own encoder emits must decode to the same list
  psb::hpack::Encoder enc;
  std::string block;
  std::vector<Header> in = {{":status", "200"}, {"content-type", "application/grpc"},
                            {"grpc-status", "0"}, {"x-empty", ""}};
  enc.encode(in, block);
  std::vector<Header> out;
  CHECK(psb::hpack::Decoder().decode(block, out) && out == in);
  std::string s;
  CHECK(!psb::hpack::huffman_decode(std::string("\xff\xff\xff\xff", 4), s));  // EOS in a string
  return 0;
}

int main() {
  using namespace psb;
  if (hpack_rfc_vectors() != 0) return 1;
  ParameterServerService ps(2, 0, PsOptions());
  int ps_port = ps.listen("127.0.0.1:0");
  CHECK(ps_port > 0);
  ps.start();
  CoordinatorService coord("127.0.0.1", ps_port, 100, 30);
  int coord_port = coord.listen("127.0.0.1:0");
  CHECK(coord_port > 0);
  coord.start();
  const std::string caddr = "127.0.0.1:" + std::to_string(coord_port);

  bool ok[2] = {true, true};
  auto run = [&](int id) {
    WorkerOptions wo;
    wo.heartbeat_ms = 50;
    Worker w(id, caddr, "", 0, wo);
    if (!w.initialize()) {
      ok[id] = false;
      return;
    }
    for (int it = 0; it < 5; ++it) ok[id] = ok[id] && w.run_iteration(it);
  };
  std::thread t0(run, 0), t1(run, 1);
  t0.join();
  t1.join();
  CHECK(ok[0] && ok[1]);
  CHECK(coord.core().num_workers() == 2);

  auto params = ps.core().serve_parameters(0);
  CHECK(params.size() == 1 && params[0].name == "weight" && params[0].data.size() == 100);
  for (float v : params[0].data) CHECK(std::fabs(v - (-0.03f)) < 1e-6f);  // App. C
  CHECK(ps.core().get_current_iteration() == 4);

  const std::string path = "/tmp/psb_selftest.ckpt";
  CHECK(ps.core().save_checkpoint(0, path));
  std::ifstream f(path, std::ios::binary | std::ios::ate);
  CHECK(f.is_open() && f.tellg() == 458);  // App. B golden size
  int32_t epoch = -1;
  ParameterServerCore other(2);
  CHECK(other.load_checkpoint(path, epoch) && epoch == 0 && other.get_current_iteration() == 4);
  std::remove(path.c_str());

  coord.shutdown();
  ps.shutdown();
  std::cout << "psb_selftest ok" << std::endl;
  return 0;
}
