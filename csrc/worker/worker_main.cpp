// worker_main — argv-compatible with the reference binary
// (/root/reference/src/worker_main.cpp:5-45):
//   worker_main [coordinator=localhost:50052] [worker_id=0] [iterations=1] [worker_addr=""]
//               [worker_port=0] [checkpoint_path=""]
// Extensions: --mode=sync|async|ssp  --device=N  --heartbeat-ms=N  --sync-timeout-ms=N
// (or PSB_MODE / PSB_DEVICE / PSB_HEARTBEAT_MS / PSB_SYNC_TIMEOUT_MS), and for the straggler
// benchmark --compute-ms=F --straggle-prob=P --straggle-factor=X --seed=N --quiet (simulated
// compute time per iteration; no per-iteration log line, one summary line at the end), and
// --trace (pull / compute / push / wait milliseconds of every iteration); --join (start at the
// iteration the job has reached instead of 0: a worker that was restarted);
// --shards=N: the model's tensors are split over the N parameter servers that registered with
// --shard-id (tensor k on shard k % N); PSB_SHARDS.
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "../common/memtune.h"
#include "worker.h"

static bool flag(const std::string& arg, const char* key, std::string& val) {
  const std::string pre = std::string("--") + key + "=";
  if (arg.compare(0, pre.size(), pre) != 0) return false;
  val = arg.substr(pre.size());
  return true;
}

static const char kUsage[] =
    "usage: worker_main [coordinator=localhost:50052] [worker_id=0] [iterations=1] [worker_addr]\n"
    "                   [worker_port=0] [checkpoint_path] [flags]\n"
    "  --mode=sync|async|ssp   wait for the barrier | never wait | wait at the staleness gate\n"
    "  --shards=N              the model's tensors are split over N parameter servers\n"
    "  --device=N              synthetic gradients are filled on GPU N\n"
    "  --heartbeat-ms=N (5000) --sync-timeout-ms=N (10000)\n"
    "  --join                  start at the iteration the job has reached (a restarted worker)\n"
    "  --trace                 pull / compute / push / wait milliseconds of every iteration\n"
    "  --quiet                 no per-iteration line, one summary line at the end\n"
    "  --sparse-rows=N [--sparse-table=ROWSxDIM (100000x16)]  also train a sparse embedding table:\n"
    "                          every iteration pushes the gradient 0.01 for N rows and pulls them back\n"
    "  --compute-ms=F --straggle-prob=P --straggle-factor=X --seed=N   simulated compute time\n";

int main(int argc, char** argv) {
  for (int i = 1; i < argc; ++i)
    if (!std::strcmp(argv[i], "--help") || !std::strcmp(argv[i], "-h")) {
      std::cout << kUsage;
      return 0;
    }
  psb::retain_freed_memory();
  std::string coordinator_addr = "localhost:50052";
  int worker_id = 0, iterations = 1;
  std::string worker_addr, checkpoint_path;
  int32_t worker_port = 0;
  psb::WorkerOptions opt;
  bool quiet = false, join = false;
  int sparse_rows = 0;  // > 0: the synthetic sparse workload next to the dense one
  long long table_rows = 100000;
  int table_dim = 16;
  if (const char* e = std::getenv("PSB_MODE")) opt.mode = e;
  if (const char* e = std::getenv("PSB_DEVICE")) opt.device = std::atoi(e);
  if (const char* e = std::getenv("PSB_HEARTBEAT_MS")) opt.heartbeat_ms = std::atoi(e);
  if (const char* e = std::getenv("PSB_SYNC_TIMEOUT_MS")) opt.sync_timeout_ms = std::atoi(e);
  if (const char* e = std::getenv("PSB_SHARDS")) opt.shards = std::atoi(e);
  std::vector<std::string> pos;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i], v;
    if (flag(a, "mode", v)) opt.mode = v;
    else if (flag(a, "device", v)) opt.device = std::atoi(v.c_str());
    else if (flag(a, "heartbeat-ms", v)) opt.heartbeat_ms = std::atoi(v.c_str());
    else if (flag(a, "sync-timeout-ms", v)) opt.sync_timeout_ms = std::atoi(v.c_str());
    else if (flag(a, "compute-ms", v)) opt.compute_ms = std::atof(v.c_str());
    else if (flag(a, "straggle-prob", v)) opt.straggle_prob = std::atof(v.c_str());
    else if (flag(a, "straggle-factor", v)) opt.straggle_factor = std::atof(v.c_str());
    else if (flag(a, "seed", v)) opt.sim_seed = std::strtoull(v.c_str(), nullptr, 10);
    else if (a == "--quiet") quiet = true;
    else if (a == "--trace") opt.trace = true;
    else if (a == "--join") join = true;
    else if (flag(a, "shards", v)) opt.shards = std::atoi(v.c_str());
    else if (flag(a, "sparse-rows", v)) sparse_rows = std::atoi(v.c_str());
    else if (flag(a, "sparse-table", v)) {
      const size_t x = v.find('x');
      if (x == std::string::npos || (table_rows = std::atoll(v.c_str())) <= 0 ||
          (table_dim = std::atoi(v.c_str() + x + 1)) <= 0) {
        std::cerr << "--sparse-table wants ROWSxDIM" << std::endl;
        return 2;
      }
    }
    else if (a.rfind("--", 0) == 0 && a.size() > 2) {
      std::cerr << "unknown flag " << a << std::endl;
      return 2;
    } else pos.push_back(a);
  }
  if (pos.size() > 0) coordinator_addr = pos[0];
  if (pos.size() > 1) worker_id = std::stoi(pos[1]);
  if (pos.size() > 2) iterations = std::stoi(pos[2]);
  if (pos.size() > 3) worker_addr = pos[3];
  if (pos.size() > 4) worker_port = std::stoi(pos[4]);
  if (pos.size() > 5) checkpoint_path = pos[5];

  psb::Worker w(worker_id, coordinator_addr, worker_addr, worker_port, opt);
  if (!w.initialize()) {
    std::cerr << "worker " << worker_id << " failed to initialize" << std::endl;
    return 1;
  }
  if (!checkpoint_path.empty()) {
    int32_t epoch = 0;
    if (w.load_checkpoint_from_server(checkpoint_path, epoch)) {
      std::cout << "worker " << worker_id << " loaded checkpoint from " << checkpoint_path
                << " (epoch " << epoch << ")" << std::endl;
    } else {
      std::cerr << "worker " << worker_id << " failed to load checkpoint from " << checkpoint_path
                << std::endl;  // continue from scratch, like the reference
    }
  }
  int first = 0;
  if (join) {  // continue where the job is, not at iteration 0
    first = std::max(0, w.server_iteration());
    std::cout << "worker " << worker_id << " joining at iteration " << first << std::endl;
  }
  // The sparse twin of the reference worker's fabricated dense gradient (src/worker.cpp:316-320):
  // iteration i touches rows [i*N, (i+1)*N) mod ROWS with 0.01 in every element and reads them back.
  if (sparse_rows > 0 && !w.create_table("embedding", table_rows, table_dim)) {
    std::cerr << "worker " << worker_id << " could not create the sparse table" << std::endl;
    return 1;
  }
  const auto t0 = std::chrono::steady_clock::now();
  int completed = 0;
  for (int it = first; it < iterations; ++it) {
    bool done = w.run_iteration(it);
    if (sparse_rows > 0) {
      std::vector<int64_t> ids(static_cast<size_t>(sparse_rows));
      for (int i = 0; i < sparse_rows; ++i)
        ids[i] = (static_cast<long long>(it) * sparse_rows + i) % table_rows;
      std::vector<float> g(ids.size() * table_dim, opt.synthetic_grad), rows;
      const bool ok = w.push_sparse(it, "embedding", table_dim, ids, g) &&
                      (opt.mode != "sync" ? w.pull_rows("embedding", ids, -1, rows)
                                          : w.pull_rows("embedding", ids, it, rows, opt.sync_timeout_ms));
      done = done && ok && rows.size() == g.size();
      if (!quiet && ok)
        std::cout << "worker " << worker_id << " iter " << it << " sparse rows=" << sparse_rows
                  << " row0[0]=" << rows[0] << std::endl;
    }
    completed += done ? 1 : 0;
    if (!quiet)
      std::cout << "worker " << worker_id << " iter " << it << " done=" << (done ? "true" : "false")
                << std::endl;
  }
  if (quiet || opt.compute_ms > 0.0) {
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    std::cout << "worker " << worker_id << " summary iterations=" << iterations << " done=" << completed
              << " elapsed_ms=" << ms << " mode=" << opt.mode << std::endl;
  }
  return 0;
}
