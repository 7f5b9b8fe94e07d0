// parameter_server — argv-compatible with the reference binary
// (/root/reference/src/parameter_main.cpp:5-22):
//   parameter_server [listen=0.0.0.0:50051] [total_workers=2] [checkpoint_interval=10]
// Extensions are trailing --key=value flags or PSB_* environment variables:
//   --mode=sync|async|ssp  --staleness=N  --optimizer=sgd|momentum|adam|adamw  --lr=F
//   --momentum=F --beta1=F --beta2=F --eps=F --weight-decay=F  --device=N (HBM-resident shard)
//   --coordinator=host:port  --membership-poll-ms=N   follow the coordinator's worker registry:
//       an evicted worker no longer blocks the sync barrier (PSB_COORDINATOR, PSB_MEMBERSHIP_POLL_MS)
//   --shard-id=I [--advertise=host:port]   one of several parameter servers: registers in the
//       coordinator's shard map (needs --coordinator); workers run with --shards=N and split the
//       model's tensors over the shards (PSB_SHARD_ID, PSB_ADVERTISE)
//   --resume=PATH|latest   load this checkpoint (latest: the newest checkpoint_epoch_*.ckpt in the
//       working directory) before serving, so a restarted server carries on (PSB_RESUME)
//   --threads=N   host threads for aggregation and update of large tensors (default: all cores,
//       at most 16; results do not depend on it) (PSB_THREADS)
//   --long-poll-ms=N   hold a "not ready" status answer up to N ms for the barrier to open
//       (default 25; 0 = answer at once like the reference) (PSB_LONG_POLL_MS)
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "../common/memtune.h"
#include "ps_service.h"

static bool flag(const std::string& arg, const char* key, std::string& val) {
  const std::string pre = std::string("--") + key + "=";
  if (arg.compare(0, pre.size(), pre) != 0) return false;
  val = arg.substr(pre.size());
  return true;
}

static const char kUsage[] =
    "usage: parameter_server [listen=0.0.0.0:50051] [total_workers=2] [checkpoint_interval=10] [flags]\n"
    "  --mode=sync|async|ssp      barrier per iteration | apply every push | bounded staleness\n"
    "  --staleness=N              SSP bound (clocks a worker may run ahead)\n"
    "  --optimizer=sgd|momentum|adam|adamw  --lr=F --momentum=F --beta1=F --beta2=F --eps=F\n"
    "  --weight-decay=F           server-side update (default: sgd, lr 1 = the reference's p -= mean)\n"
    "  --device=N                 keep master weights and optimizer state in the HBM of GPU N\n"
    "  --threads=N                host threads for aggregation / update of large tensors\n"
    "  --coordinator=host:port    follow the worker registry: an evicted worker stops blocking\n"
    "  --membership-poll-ms=N     ... polled this often (1000)\n"
    "  --shard-id=I [--advertise=host:port]  one of several servers; registers in the shard map\n"
    "  --resume=PATH|latest       load a checkpoint before serving\n"
    "  --long-poll-ms=N           hold a 'not ready' status answer up to N ms (25)\n"
    "  --stats-every-s=N          one JSON line of counters on stdout every N seconds\n"
    "  Besides the reference's ParameterServer service it answers psb.ParameterServerExt\n"
    "  (CreateTable / PushSparse / PullRows: sparse embedding tables, same optimizer and mode).\n"
    "  Every flag has a PSB_* environment twin (PSB_MODE, PSB_OPTIMIZER, ...).\n";

int main(int argc, char** argv) {
  for (int i = 1; i < argc; ++i)
    if (!std::strcmp(argv[i], "--help") || !std::strcmp(argv[i], "-h")) {
      std::cout << kUsage;
      return 0;
    }
  psb::retain_freed_memory();
  std::string listen = "0.0.0.0:50051";
  int total_workers = 2, checkpoint_interval = 10;
  psb::PsOptions opt;
  auto env = [](const char* k) -> const char* { return std::getenv(k); };
  if (const char* e = env("PSB_MODE")) {
    std::string m = e;
    opt.mode = m == "async" ? psb::PsMode::kAsync : m == "ssp" ? psb::PsMode::kSsp
                                                               : psb::PsMode::kSync;
  }
  if (const char* e = env("PSB_STALENESS")) opt.staleness = std::atoi(e);
  if (const char* e = env("PSB_OPTIMIZER")) opt.optimizer = e;
  if (const char* e = env("PSB_LR")) opt.lr = std::strtof(e, nullptr);
  if (const char* e = env("PSB_DEVICE")) opt.device = std::atoi(e);
  if (const char* e = env("PSB_LONG_POLL_MS")) opt.long_poll_ms = std::atoi(e);
  if (const char* e = env("PSB_THREADS")) opt.threads = std::atoi(e);
  std::string coordinator, resume, advertise;
  int shard_id = -1;
  if (const char* e = env("PSB_SHARD_ID")) shard_id = std::atoi(e);
  if (const char* e = env("PSB_ADVERTISE")) advertise = e;
  int membership_poll_ms = 1000;
  if (const char* e = env("PSB_RESUME")) resume = e;
  if (const char* e = env("PSB_COORDINATOR")) coordinator = e;
  if (const char* e = env("PSB_MEMBERSHIP_POLL_MS")) membership_poll_ms = std::atoi(e);
  std::vector<std::string> pos;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i], v;
    if (flag(a, "mode", v)) {
      opt.mode = v == "async" ? psb::PsMode::kAsync : v == "ssp" ? psb::PsMode::kSsp
                                                                 : psb::PsMode::kSync;
    } else if (flag(a, "staleness", v)) opt.staleness = std::atoi(v.c_str());
    else if (flag(a, "optimizer", v)) opt.optimizer = v;
    else if (flag(a, "lr", v)) opt.lr = std::strtof(v.c_str(), nullptr);
    else if (flag(a, "momentum", v)) opt.momentum = std::strtof(v.c_str(), nullptr);
    else if (flag(a, "beta1", v)) opt.beta1 = std::strtof(v.c_str(), nullptr);
    else if (flag(a, "beta2", v)) opt.beta2 = std::strtof(v.c_str(), nullptr);
    else if (flag(a, "eps", v)) opt.eps = std::strtof(v.c_str(), nullptr);
    else if (flag(a, "weight-decay", v)) opt.weight_decay = std::strtof(v.c_str(), nullptr);
    else if (flag(a, "device", v)) opt.device = std::atoi(v.c_str());
    else if (flag(a, "long-poll-ms", v)) opt.long_poll_ms = std::atoi(v.c_str());
    else if (flag(a, "threads", v)) opt.threads = std::atoi(v.c_str());
    else if (flag(a, "stats-every-s", v)) setenv("PSB_STATS_EVERY_S", v.c_str(), 1);
    else if (flag(a, "coordinator", v)) coordinator = v;
    else if (flag(a, "membership-poll-ms", v)) membership_poll_ms = std::atoi(v.c_str());
    else if (flag(a, "resume", v)) resume = v;
    else if (flag(a, "shard-id", v)) shard_id = std::atoi(v.c_str());
    else if (flag(a, "advertise", v)) advertise = v;
    else if (a.rfind("--", 0) == 0) {
      std::cerr << "unknown flag " << a << std::endl;
      return 2;
    } else pos.push_back(a);
  }
  if (pos.size() > 0) listen = pos[0];
  if (pos.size() > 1) total_workers = std::stoi(pos[1]);
  if (pos.size() > 2) checkpoint_interval = std::stoi(pos[2]);
  psb::run_server(listen, total_workers, checkpoint_interval, opt, coordinator, membership_poll_ms,
                  resume, shard_id, advertise);
  return 0;
}
