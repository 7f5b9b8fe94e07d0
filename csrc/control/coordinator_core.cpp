// coordinator_core.cpp — see coordinator_core.h.
#include "coordinator_core.h"

#include <algorithm>

namespace psb {

CoordinatorCore::CoordinatorCore(const std::string& ps_host, int32_t ps_port)
    : ps_host_(ps_host), ps_port_(ps_port) {}

bool CoordinatorCore::register_worker(const WorkerEntry& w, std::string& ps_address,
                                      int32_t& total_workers) {
  std::lock_guard<std::mutex> lk(mu_);
  WorkerEntry e = w;
  e.last_heartbeat = std::chrono::steady_clock::now();
  workers_[w.worker_id] = e;
  ps_address = ps_host_.find(':') == std::string::npos
                   ? ps_host_ + ":" + std::to_string(ps_port_)
                   : ps_host_;
  total_workers = static_cast<int32_t>(workers_.size());
  cv_.notify_all();
  return true;
}

bool CoordinatorCore::update_heartbeat(int32_t worker_id, int32_t status) {
  std::lock_guard<std::mutex> lk(mu_);
  auto it = workers_.find(worker_id);
  if (it == workers_.end()) return false;
  it->second.last_heartbeat = std::chrono::steady_clock::now();
  it->second.status = status;
  return true;
}

std::vector<WorkerEntry> CoordinatorCore::list_workers() {
  std::lock_guard<std::mutex> lk(mu_);
  std::vector<WorkerEntry> out;
  out.reserve(workers_.size());
  for (const auto& kv : workers_) out.push_back(kv.second);
  return out;
}

void CoordinatorCore::get_parameter_server_address(std::string& host, int32_t& port) {
  std::lock_guard<std::mutex> lk(mu_);
  host = ps_host_;
  port = ps_port_;
}

size_t CoordinatorCore::num_workers() {
  std::lock_guard<std::mutex> lk(mu_);
  return workers_.size();
}

std::vector<int32_t> CoordinatorCore::remove_stale_workers(int64_t timeout_seconds) {
  std::lock_guard<std::mutex> lk(mu_);
  std::vector<int32_t> evicted;
  const auto now = std::chrono::steady_clock::now();
  for (auto it = workers_.begin(); it != workers_.end();) {
    auto silent = std::chrono::duration_cast<std::chrono::milliseconds>(now - it->second.last_heartbeat);
    if (silent > std::chrono::seconds(timeout_seconds)) {
      evicted.push_back(it->first);
      clocks_.erase(it->first);
      it = workers_.erase(it);
    } else {
      ++it;
    }
  }
  if (!evicted.empty()) cv_.notify_all();  // barriers re-evaluate their membership
  return evicted;
}

void CoordinatorCore::register_shard(const ShardEntry& s) {
  std::lock_guard<std::mutex> lk(mu_);
  shards_[s.shard_id] = s;
}

std::vector<ShardEntry> CoordinatorCore::shard_map() {
  std::lock_guard<std::mutex> lk(mu_);
  std::vector<ShardEntry> out;
  for (const auto& kv : shards_) out.push_back(kv.second);
  return out;
}

int32_t CoordinatorCore::barrier(const std::string& name, int32_t worker_id, int32_t generation,
                                 int32_t expected, int64_t timeout_ms, bool& released) {
  std::unique_lock<std::mutex> lk(mu_);
  BarrierState& b = barriers_[name];
  if (generation > b.generation) {  // first arrival of a new generation resets the barrier
    b.generation = generation;
    b.arrived.clear();
    b.open = false;
  }
  if (generation < b.generation) {  // late arrival for a generation that already passed
    released = true;
    return static_cast<int32_t>(b.arrived.size());
  }
  b.arrived[worker_id] = true;
  auto need = [&]() -> size_t {
    return expected > 0 ? static_cast<size_t>(expected) : std::max<size_t>(workers_.size(), 1);
  };
  auto live_arrivals = [&]() -> size_t {
    if (expected > 0) return b.arrived.size();
    size_t n = 0;
    for (const auto& kv : b.arrived) n += workers_.count(kv.first);
    return n;
  };
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  while (true) {
    if (b.generation != generation || b.open || live_arrivals() >= need()) {
      if (b.generation == generation) b.open = true;
      cv_.notify_all();
      released = true;
      return static_cast<int32_t>(b.arrived.size());
    }
    if (cv_.wait_until(lk, deadline) == std::cv_status::timeout) {
      released = false;
      return static_cast<int32_t>(b.arrived.size());
    }
  }
}

int32_t CoordinatorCore::update_clock(int32_t worker_id, int32_t clock) {
  std::lock_guard<std::mutex> lk(mu_);
  clocks_[worker_id] = std::max(clocks_[worker_id], clock);
  int32_t lo = clock;
  for (const auto& kv : clocks_) lo = std::min(lo, kv.second);
  return lo;
}

void CoordinatorCore::put_handle(const std::string& key, int32_t rank, const std::string& blob) {
  std::lock_guard<std::mutex> lk(mu_);
  if (blob.empty()) {  // an empty blob withdraws the entry (keeps repeated exchanges bounded)
    auto it = handles_.find(key);
    if (it != handles_.end()) {
      it->second.erase(rank);
      if (it->second.empty()) handles_.erase(it);
    }
    return;
  }
  handles_[key][rank] = blob;
}

std::vector<std::pair<int32_t, std::string>> CoordinatorCore::get_handles(const std::string& key) {
  std::lock_guard<std::mutex> lk(mu_);
  std::vector<std::pair<int32_t, std::string>> out;
  auto it = handles_.find(key);
  if (it != handles_.end())
    for (const auto& kv : it->second) out.emplace_back(kv.first, kv.second);
  return out;
}

}  // namespace psb
