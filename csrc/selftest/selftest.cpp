// psb_selftest — in-process smoke test of the C++ control plane (no Python, no GPU):
// coordinator + parameter server + two workers over real sockets on 127.0.0.1, checked
// against the golden behavioural trace and the checkpoint bytes (App. B).
// Also the program the TSAN / ASAN builds run (`make tsan && build/bin-thread/psb_selftest`).
#include <cmath>
#include <cstdio>
#include <cstring>
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

// The threaded aggregation / optimizer loops of ps_core.cpp give the same bits for every thread
// count: three workers push 300 k-element tensors (above the parallel grain) for three
// iterations, once with 1 thread and once with 5, for every optimizer.
static int threads_do_not_change_results() {
  using namespace psb;
  for (const char* kind : {"sgd", "momentum", "adam", "adamw"}) {
    std::vector<std::vector<HostTensor>> results;
    for (int threads : {1, 5}) {
      PsOptions opt;
      opt.optimizer = kind;
      opt.lr = 0.01f;
      opt.weight_decay = 0.1f;
      opt.threads = threads;
      ParameterServerCore core(3, opt);
      uint32_t seed = 12345;
      for (int it = 0; it < 4; ++it) {
        for (int w = 0; w < 3; ++w) {
          std::vector<HostTensor> g(2);
          g[0].name = "big", g[0].shape = {300, 1000};
          g[1].name = "small", g[1].shape = {17};
          g[0].data.resize(300000), g[1].data.resize(17);
          for (auto& t : g)
            for (auto& x : t.data) {
              seed = seed * 1664525u + 1013904223u;
              x = (static_cast<int32_t>(seed >> 9) - (1 << 22)) * (1.0f / (1 << 22));
            }
          const bool done = core.receive_gradients(w, it, std::move(g));
          CHECK(done == (w == 2));
        }
      }
      results.push_back(core.serve_parameters(0));
    }
    CHECK(results[0].size() == 2 && results[1].size() == 2);
    for (size_t i = 0; i < 2; ++i)
      CHECK(results[0][i].data.size() == results[1][i].data.size() &&
            memcmp(results[0][i].data.data(), results[1][i].data.data(),
                   results[0][i].data.size() * sizeof(float)) == 0);
  }
  return 0;
}

// Two parameter servers register as shards 0 and 1; two workers with shards = 2 split a
// three-tensor model over them (tensor k on shard k % 2) and run five iterations.
static int sharded_parameter_servers() {
  using namespace psb;
  CoordinatorService coord("127.0.0.1", 1, 100, 30);
  const int coord_port = coord.listen("127.0.0.1:0");
  CHECK(coord_port > 0);
  coord.start();
  const std::string caddr = "127.0.0.1:" + std::to_string(coord_port);
  ParameterServerService s0(2, 0, PsOptions()), s1(2, 0, PsOptions());
  const int p0 = s0.listen("127.0.0.1:0"), p1 = s1.listen("127.0.0.1:0");
  CHECK(p0 > 0 && p1 > 0);
  s0.start();
  s1.start();
  s0.announce_shard(caddr, 0, "127.0.0.1:" + std::to_string(p0), 50);
  s1.announce_shard(caddr, 1, "127.0.0.1:" + std::to_string(p1), 50);
  auto model = [](float v) {
    std::vector<HostTensor> m(3);
    const char* names[3] = {"a", "b", "c"};
    for (int k = 0; k < 3; ++k) {
      m[k].name = names[k];
      m[k].shape = {4 + k};
      m[k].data.assign(static_cast<size_t>(4 + k), v + k);
    }
    return m;
  };
  {  // tensors a, c on shard 0; b on shard 1
    std::vector<HostTensor> m = model(1.0f);
    s0.core().initialize_parameters({m[0], m[2]});
    s1.core().initialize_parameters({m[1]});
  }
  bool ok[2] = {true, true};
  auto run = [&](int id) {
    WorkerOptions wo;
    wo.heartbeat_ms = 50;
    wo.shards = 2;
    Worker w(id, caddr, "", 0, wo);
    if (!w.initialize()) {
      ok[id] = false;
      return;
    }
    bool ready = false;
    std::vector<TensorLite> pulled = w.pull_parameters(0, &ready);
    ok[id] = pulled.size() == 3 && pulled[0].name == "a" && pulled[1].name == "b" &&
             pulled[2].name == "c" && pulled[1].data.size() == 5;
    for (int it = 0; it < 5; ++it) ok[id] = ok[id] && w.run_iteration(it);
  };
  std::thread t0(run, 0), t1(run, 1);
  t0.join();
  t1.join();
  CHECK(ok[0] && ok[1]);
  auto a = s0.core().serve_parameters(0), b = s1.core().serve_parameters(0);
  CHECK(a.size() == 2 && b.size() == 1 && a[1].name == "c" && b[0].name == "b");
  for (float v : a[0].data) CHECK(std::fabs(v - (1.0f - 0.05f)) < 1e-6f);
  for (float v : b[0].data) CHECK(std::fabs(v - (2.0f - 0.05f)) < 1e-6f);
  for (float v : a[1].data) CHECK(std::fabs(v - (3.0f - 0.05f)) < 1e-6f);
  coord.shutdown();
  s0.shutdown();
  s1.shutdown();
  return 0;
}

// Sparse tables (ps_sparse.h): sync staging and completion, membership shrink completing a
// pending iteration with the live worker's gradient alone, lazy updates, async once-per-clock,
// save / load with optimizer state.
static int sparse_tables() {
  using namespace psb;
  auto grad = [](std::vector<int64_t> idx, float v, int dim) {
    SparseGrad g;
    g.name = "emb", g.dim = dim, g.indices = std::move(idx);
    g.values.assign(g.indices.size() * dim, v);
    return std::vector<SparseGrad>{g};
  };
  PsOptions o;
  o.optimizer = "momentum", o.lr = 0.5f, o.momentum = 0.5f;
  SparseTables t(o);
  CHECK(t.create("emb", 10, 2, 0.0f, 0) && t.create("emb", 10, 2, 0.5f, 9) && !t.create("emb", 10, 4, 0, 0));
  CHECK(!t.create("", 1, 1, 0, 0) && !t.create("x", 0, 1, 0, 0));
  int32_t got = 0, dim = 0;
  std::string err;
  std::vector<float> rows;
  CHECK(!t.push(0, 0, grad({1, 1, 3, -1, 10}, 1.0f, 2), 2, got, err) && got == 1 && err.empty());
  CHECK(!t.ready(0, 0) && t.ready(-1, 0));
  CHECK(t.push(1, 0, grad({1}, 2.0f, 2), 2, got, err) && got == 2 && t.ready(0, 0));
  CHECK(t.push(1, 0, grad({1}, 2.0f, 2), 2, got, err));  // idempotent after completion
  CHECK(t.rows("emb", {1, 3, 0, -1, 10}, rows, dim) && dim == 2 && rows.size() == 10);
  // row 1: (1 + 1 + 2) / 2 = 2 -> m = 2, p = -1; row 3: 1 / 2 -> m = .5, p = -.25; row 0 untouched
  CHECK(std::fabs(rows[0] + 1.0f) < 1e-6f && std::fabs(rows[2] + 0.25f) < 1e-6f && rows[4] == 0.0f);
  CHECK(rows[6] == 0.0f && rows[8] == 0.0f);
  // a pending iteration completes when the membership shrinks to the worker that did push
  CHECK(!t.push(0, 1, grad({3}, 1.0f, 2), 2, got, err) && !t.ready(1, 0));
  CHECK(t.set_total_workers(1) == 1 && t.ready(1, 0));
  CHECK(t.rows("emb", {3}, rows, dim));
  CHECK(std::fabs(rows[0] - (-0.25f - 0.5f * (0.5f * 0.5f + 1.0f))) < 1e-6f);  // divisor = 1 pusher
  // malformed pushes are rejected whole
  CHECK(!t.push(0, 2, grad({1}, 1.0f, 4), 1, got, err) && !err.empty());
  const std::string path = "/tmp/psb_selftest.tables";
  CHECK(t.save(path));
  SparseTables u(o);
  CHECK(u.load(path));
  std::vector<float> a, b;
  err.clear();
  CHECK(t.push(0, 5, grad({1, 3}, 1.0f, 2), 1, got, err) && u.push(0, 5, grad({1, 3}, 1.0f, 2), 1, got, err));
  CHECK(t.rows("emb", {0, 1, 3}, a, dim) && u.rows("emb", {0, 1, 3}, b, dim) && a == b);  // state too
  std::remove(path.c_str());
  // large pushes update their rows on several threads: same bits as on one
  {
    PsOptions o1, o4;
    o1.optimizer = o4.optimizer = "adam", o1.lr = o4.lr = 0.01f;
    o1.threads = 1, o4.threads = 4;
    SparseTables one(o1), four(o4);
    CHECK(one.create("emb", 20000, 8, 0.1f, 3) && four.create("emb", 20000, 8, 0.1f, 3));
    SparseGrad g;
    g.name = "emb", g.dim = 8;
    for (int i = 0; i < 6000; ++i) {
      g.indices.push_back((i * 7919) % 20000);
      for (int k = 0; k < 8; ++k) g.values.push_back(0.001f * static_cast<float>((i * 31 + k) % 97) - 0.04f);
    }
    for (int it = 0; it < 2; ++it) {
      CHECK(one.push(0, it, std::vector<SparseGrad>{g}, 1, got, err));
      CHECK(four.push(0, it, std::vector<SparseGrad>{g}, 1, got, err));
    }
    std::vector<int64_t> all(20000);
    for (int i = 0; i < 20000; ++i) all[i] = i;
    CHECK(one.rows("emb", all, a, dim) && four.rows("emb", all, b, dim) && a == b);
  }
  // async: a share of 1 / N at once, each clock once
  PsOptions as;
  as.mode = PsMode::kAsync, as.lr = 1.0f;
  SparseTables v(as);
  CHECK(v.create("emb", 4, 2, 0.0f, 0));
  CHECK(v.push(0, 0, grad({2}, 1.0f, 2), 2, got, err) && v.push(0, 0, grad({2}, 1.0f, 2), 2, got, err));
  CHECK(v.rows("emb", {2}, rows, dim) && std::fabs(rows[0] + 0.5f) < 1e-6f);
  return 0;
}

int main() {
  using namespace psb;
  if (sparse_tables() != 0) return 1;
  if (hpack_rfc_vectors() != 0) return 1;
  if (threads_do_not_change_results() != 0) return 1;
  if (sharded_parameter_servers() != 0) return 1;
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
