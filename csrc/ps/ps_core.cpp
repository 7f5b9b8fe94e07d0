// ps_core.cpp — see ps_core.h.
#include "ps_core.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <thread>

#include "ps_device.h"

namespace psb {

namespace {
// Element-wise loops over large tensors run on a few threads.  Every element is still computed
// by the same expression in the same order (the sum over workers is ascending in worker id per
// element), so results do not depend on the thread count.  Threads are created per call: the
// loops this is used for run for milliseconds.
constexpr size_t kParallelGrain = 1 << 16;

template <class F>
void parallel_for(size_t n, int max_threads, F&& body) {  // body(lo, hi)
  size_t t = n / kParallelGrain;
  if (t > static_cast<size_t>(max_threads)) t = static_cast<size_t>(max_threads);
  if (t <= 1) {
    body(size_t{0}, n);
    return;
  }
  std::vector<std::thread> th;
  th.reserve(t - 1);
  const size_t step = (n + t - 1) / t;
  for (size_t k = 1; k < t; ++k) {
    const size_t lo = k * step, hi = std::min(n, lo + step);
    if (lo >= hi) break;
    th.emplace_back([&body, lo, hi] { body(lo, hi); });
  }
  body(size_t{0}, std::min(n, step));
  for (auto& x : th) x.join();
}

int resolve_threads(int requested) {
  int t = requested > 0 ? requested : static_cast<int>(std::thread::hardware_concurrency());
  if (t < 1) t = 1;
  return t > 16 ? 16 : t;
}
}  // namespace


ParameterServerCore::ParameterServerCore(int total_workers, const PsOptions& opt)
    : total_workers_(total_workers), opt_(opt) {
  opt_.threads = resolve_threads(opt_.threads);
  if (opt_.device >= 0) dev_ = DeviceShard::create(opt_);  // nullptr if CUDA is unavailable
}

ParameterServerCore::~ParameterServerCore() { delete dev_; }

void ParameterServerCore::initialize_parameters(const std::vector<HostTensor>& initial) {
  std::lock_guard<std::mutex> lk(mu_);
  params_ = initial;
  m_.assign(params_.size(), {});
  v_.assign(params_.size(), {});
  if (dev_) dev_->upload(params_);
  version_++;
}

bool ParameterServerCore::receive_gradients(int32_t worker_id, int32_t iteration,
                                            const std::vector<HostTensor>& grads) {
  return receive_gradients(worker_id, iteration, std::vector<HostTensor>(grads));
}

bool ParameterServerCore::receive_gradients(int32_t worker_id, int32_t iteration,
                                            std::vector<HostTensor>&& grads) {
  std::lock_guard<std::mutex> lk(mu_);
  if (iteration > current_iteration_.load()) current_iteration_.store(iteration);

  if (opt_.mode != PsMode::kSync) {
    // async / SSP: apply this worker's contribution at once, scaled like one share of the mean.
    // A clock is applied once: the transport re-sends an RPC whose connection died after the
    // request may already have been delivered, and such a duplicate must not count twice.
    auto ck = clocks_.find(worker_id);
    if (ck != clocks_.end() && iteration <= ck->second) return true;
    if (params_.empty()) {
      apply_locked(grads);  // bootstrap from the unscaled tensor, as in sync mode
    } else {
      const float share = 1.0f / static_cast<float>(std::max(1, total_workers_.load()));
      for (auto& t : grads) {  // ours to scale in place
        float* x = t.data.data();
        parallel_for(t.data.size(), opt_.threads, [x, share](size_t lo, size_t hi) {
          for (size_t j = lo; j < hi; ++j) x[j] *= share;
        });
      }
      apply_locked(grads);
    }
    clocks_[worker_id] = iteration;
    IterState& st = iters_[iteration];
    st.seen[worker_id] = true;
    st.aggregated = true;
    prune_locked();
    changed_.notify_all();
    return true;
  }

  IterState& st = iters_[iteration];
  if (st.aggregated) return true;  // idempotent retry after completion
  st.staged[worker_id] = std::move(grads);  // a re-push from the same worker overwrites its slot
  st.seen[worker_id] = true;
  if (st.staged.size() < static_cast<size_t>(std::max(1, total_workers_.load()))) return false;
  // names / shapes / dtypes come from the completing push
  aggregate_locked(st, worker_id);
  prune_locked();
  return true;
}

// Mean over the staged contributions in ascending worker id (deterministic summation order),
// shaped like `tmpl`; shorter contributions are truncated element-wise (reference semantics,
// src/parameter_server.cpp:40-63).
void ParameterServerCore::aggregate_locked(IterState& st, int32_t shape_from) {
  const float div = static_cast<float>(st.staged.size());
  // The mean is written over the staged gradient of worker `shape_from` (we own that memory):
  // no third model-sized buffer, no page faults on the completing push.  Element j of the
  // output is stored after every read of element j, so summing in place is safe.
  std::vector<HostTensor> mean = std::move(st.staged[shape_from]);
  std::vector<std::pair<const float*, size_t>> src;  // per tensor: every worker's (data, size)
  for (size_t i = 0; i < mean.size(); ++i) {
    const size_t n = mean[i].data.size();
    float* out = mean[i].data.data();
    src.clear();
    for (const auto& kv : st.staged) {  // ascending worker id
      if (kv.first == shape_from) {
        src.emplace_back(out, n);
      } else if (i < kv.second.size()) {
        const std::vector<float>& g = kv.second[i].data;
        src.emplace_back(g.data(), std::min(n, g.size()));
      }
    }
    parallel_for(n, opt_.threads, [&src, out, div](size_t lo, size_t hi) {
      constexpr size_t kBlock = 2048;  // sum block-wise: streams every source, keeps `acc` in L1
      float acc[kBlock];
      for (size_t b = lo; b < hi; b += kBlock) {
        const size_t e = std::min(hi, b + kBlock);
        for (size_t j = b; j < e; ++j) acc[j - b] = 0.0f;  // the sum starts from 0.0f (reference)
        for (const auto& s : src) {
          const size_t lim = std::min(e, s.second);
          const float* g = s.first;
          for (size_t j = b; j < lim; ++j) acc[j - b] += g[j];
        }
        for (size_t j = b; j < e; ++j) out[j] = acc[j - b] / div;
      }
    });
  }
  st.aggregated = true;
  st.staged.clear();  // the reference keeps every copy forever
  apply_locked(mean);
  changed_.notify_all();
}

void ParameterServerCore::prune_locked() {
  while (static_cast<int>(iters_.size()) > opt_.keep_iterations) iters_.erase(iters_.begin());
}

// Membership change (a worker was evicted by the coordinator, or came back).  Iterations that
// were waiting for the missing worker are re-evaluated at once and complete with the mean of
// the live workers' gradients — the reference keeps total_workers fixed for the life of the
// process and stalls for ever (src/coordinator.cpp:52-67 evicts,
// src/parameter_main.cpp:13-15 never hears of it).
int ParameterServerCore::set_total_workers(int n) {
  std::lock_guard<std::mutex> lk(mu_);
  total_workers_.store(n);
  changed_.notify_all();  // the SSP gate depends on the count; sync completions notify below
  int completed = 0;
  if (opt_.mode != PsMode::kSync) return 0;
  const size_t need = static_cast<size_t>(std::max(1, n));
  for (auto& kv : iters_) {  // ascending iteration
    IterState& st = kv.second;
    if (st.aggregated || st.staged.empty() || st.staged.size() < need) continue;
    aggregate_locked(st, st.staged.rbegin()->first);
    ++completed;
  }
  return completed;
}

void ParameterServerCore::apply_locked(const std::vector<HostTensor>& g) {
  if (params_.empty()) {  // bootstrap: the first aggregate becomes the model
    params_ = g;
    m_.assign(params_.size(), {});
    v_.assign(params_.size(), {});
    if (dev_) dev_->upload(params_);
    version_++;
    return;
  }
  ++opt_step_;
  for (size_t i = 0; i < g.size() && i < params_.size(); ++i) {
    if (g[i].name != params_[i].name || g[i].shape != params_[i].shape) continue;  // positional
    update_tensor(i, g[i].data);
  }
  if (dev_ && dev_->active() && !dev_->sync()) {
    // the device lost this step: bring back what it still has and continue on the host
    dev_->download_state(m_, v_);
    dev_->download(params_);
  }
  version_++;
}

void ParameterServerCore::update_tensor(size_t idx, const std::vector<float>& g) {
  std::vector<float>& p = params_[idx].data;
  const size_t n = std::min(p.size(), g.size());
  if (dev_ && dev_->active()) {
    if (dev_->update(idx, g.data(), n, opt_step_)) return;
    // the launch failed: fetch the newest moments and weights (the download retires the GPU
    // copy), then this and every later update run on the host
    dev_->download_state(m_, v_);
    dev_->download(params_);
  }
  const std::string& kind = opt_.optimizer;
  const float lr = opt_.lr, wd = opt_.weight_decay;
  float* pp = p.data();
  const float* gg = g.data();
  const int nt = opt_.threads;
  if (kind == "sgd") {
    if (lr == 1.0f && wd == 0.0f) {
      parallel_for(n, nt, [pp, gg](size_t lo, size_t hi) {
        for (size_t k = lo; k < hi; ++k) pp[k] -= gg[k];  // the reference's exact expression
      });
    } else {
      parallel_for(n, nt, [pp, gg, lr, wd](size_t lo, size_t hi) {
        for (size_t k = lo; k < hi; ++k) pp[k] -= lr * (gg[k] + wd * pp[k]);
      });
    }
    return;
  }
  if (m_[idx].size() != p.size()) m_[idx].assign(p.size(), 0.0f);
  float* m = m_[idx].data();
  if (kind == "momentum") {
    const float mu = opt_.momentum;
    parallel_for(n, nt, [pp, gg, m, lr, wd, mu](size_t lo, size_t hi) {
      for (size_t k = lo; k < hi; ++k) {
        m[k] = mu * m[k] + (gg[k] + wd * pp[k]);
        pp[k] -= lr * m[k];
      }
    });
    return;
  }
  if (v_[idx].size() != p.size()) v_[idx].assign(p.size(), 0.0f);
  float* v = v_[idx].data();
  const bool decoupled = kind == "adamw";
  const float b1 = opt_.beta1, b2 = opt_.beta2, eps = opt_.eps;
  const float bias1 = 1.0f - std::pow(b1, static_cast<float>(opt_step_));
  const float bias2 = 1.0f - std::pow(b2, static_cast<float>(opt_step_));
  parallel_for(n, nt, [=](size_t lo, size_t hi) {
    for (size_t k = lo; k < hi; ++k) {
      float gj = gg[k];
      if (decoupled) pp[k] *= 1.0f - lr * wd;
      else gj += wd * pp[k];
      m[k] = b1 * m[k] + (1.0f - b1) * gj;
      v[k] = b2 * v[k] + (1.0f - b2) * gj * gj;
      const float denom = std::sqrt(v[k]) / std::sqrt(bias2) + eps;
      pp[k] -= (lr / bias1) * (m[k] / denom);
    }
  });
}

std::vector<HostTensor> ParameterServerCore::serve_parameters(int32_t /*iteration*/) {
  std::lock_guard<std::mutex> lk(mu_);
  if (dev_) dev_->download(params_);
  return params_;
}

void ParameterServerCore::with_parameters(
    const std::function<void(const std::vector<HostTensor>&, int64_t)>& fn) {
  std::lock_guard<std::mutex> lk(mu_);
  if (dev_) dev_->download(params_);
  fn(params_, version_.load());
}

bool ParameterServerCore::sync_status_locked(int32_t iteration, int32_t& workers_received) {
  auto it = iters_.find(iteration);
  if (it == iters_.end()) {
    workers_received = 0;
    return false;
  }
  workers_received = static_cast<int32_t>(it->second.seen.size());
  return it->second.aggregated;
}

bool ParameterServerCore::check_sync_status(int32_t iteration, int32_t& workers_received) {
  std::lock_guard<std::mutex> lk(mu_);
  return sync_status_locked(iteration, workers_received);
}

bool ParameterServerCore::wait_sync_status(int32_t iteration, int32_t& workers_received) {
  std::unique_lock<std::mutex> lk(mu_);
  if (sync_status_locked(iteration, workers_received) || opt_.long_poll_ms <= 0) {
    return sync_status_locked(iteration, workers_received);
  }
  changed_.wait_for(lk, std::chrono::milliseconds(opt_.long_poll_ms), [&] {
    int32_t r = 0;
    return sync_status_locked(iteration, r);
  });
  return sync_status_locked(iteration, workers_received);
}

bool ParameterServerCore::staleness_ok_locked(int32_t iteration) {
  if (opt_.mode == PsMode::kAsync) return true;
  if (opt_.mode == PsMode::kSync) {
    auto it = iters_.find(iteration);
    return it != iters_.end() && it->second.aggregated;
  }
  int32_t slowest = -1;  // a worker that never pushed has completed nothing
  if (static_cast<int>(clocks_.size()) >= total_workers_.load()) {
    slowest = INT32_MAX;
    for (const auto& kv : clocks_) slowest = std::min(slowest, kv.second);
  }
  return iteration - (slowest + 1) <= opt_.staleness;
}

bool ParameterServerCore::staleness_ok(int32_t /*worker_id*/, int32_t iteration) {
  std::lock_guard<std::mutex> lk(mu_);
  return staleness_ok_locked(iteration);
}

bool ParameterServerCore::wait_staleness_ok(int32_t /*worker_id*/, int32_t iteration) {
  std::unique_lock<std::mutex> lk(mu_);
  if (opt_.mode == PsMode::kSsp && opt_.long_poll_ms > 0 && !staleness_ok_locked(iteration)) {
    changed_.wait_for(lk, std::chrono::milliseconds(opt_.long_poll_ms),
                      [&] { return staleness_ok_locked(iteration); });
  }
  return staleness_ok_locked(iteration);
}

// ---- checkpoints: SURVEY App. B (raw little-endian structs, size_t = 8 bytes, no padding) ----
bool ParameterServerCore::write_file(int32_t epoch, const std::string& path) {
  const std::string tmp = path + ".tmp";
  {
    std::ofstream f(tmp, std::ios::binary | std::ios::trunc);
    if (!f.is_open()) return false;
    const int32_t iter = current_iteration_.load();
    f.write(reinterpret_cast<const char*>(&epoch), 4);
    f.write(reinterpret_cast<const char*>(&iter), 4);
    const uint64_t nt = params_.size();
    f.write(reinterpret_cast<const char*>(&nt), 8);
    for (const auto& t : params_) {
      const uint64_t nl = t.name.size(), nd = t.shape.size(), ne = t.data.size();
      f.write(reinterpret_cast<const char*>(&nl), 8);
      f.write(t.name.data(), static_cast<std::streamsize>(nl));
      f.write(reinterpret_cast<const char*>(&nd), 8);
      f.write(reinterpret_cast<const char*>(t.shape.data()), static_cast<std::streamsize>(nd * 4));
      f.write(reinterpret_cast<const char*>(&t.dtype), 4);
      f.write(reinterpret_cast<const char*>(&ne), 8);
      f.write(reinterpret_cast<const char*>(t.data.data()), static_cast<std::streamsize>(ne * 4));
    }
    f.flush();
    if (!f.good()) return false;
  }
  return std::rename(tmp.c_str(), path.c_str()) == 0;  // atomic publish (reference writes in place)
}

bool ParameterServerCore::save_checkpoint(int32_t epoch, const std::string& path) {
  std::lock_guard<std::mutex> lk(mu_);
  if (dev_) {
    // moments of an HBM-resident shard live only on the device: without this the sidecar would
    // carry opt_step_ with empty m / v and a resumed Adam would run late-step bias correction
    // on zero moments
    if (!dev_->download(params_)) return false;
    if (!dev_->download_state(m_, v_)) return false;
  }
  if (!write_file(epoch, path)) return false;
  // sidecar: optimizer state and mode, so the base file stays loadable by the reference
  bool has_state = false;
  for (const auto& m : m_) has_state |= !m.empty();
  const std::string side = path + ".psb";
  if (!has_state && opt_step_ == 0) {
    std::remove(side.c_str());
    return true;
  }
  std::ofstream f(side + ".tmp", std::ios::binary | std::ios::trunc);
  if (!f.is_open()) return true;  // the parameters are safe; the sidecar is best effort
  const char magic[4] = {'P', 'S', 'B', 'S'};
  const uint32_t ver = 1;
  f.write(magic, 4);
  f.write(reinterpret_cast<const char*>(&ver), 4);
  f.write(reinterpret_cast<const char*>(&opt_step_), 8);
  const uint64_t nk = opt_.optimizer.size(), nt = params_.size();
  f.write(reinterpret_cast<const char*>(&nk), 8);
  f.write(opt_.optimizer.data(), static_cast<std::streamsize>(nk));
  f.write(reinterpret_cast<const char*>(&nt), 8);
  for (size_t i = 0; i < params_.size(); ++i) {
    for (const std::vector<float>* s : {&m_[i], &v_[i]}) {
      const uint64_t n = s->size();
      f.write(reinterpret_cast<const char*>(&n), 8);
      f.write(reinterpret_cast<const char*>(s->data()), static_cast<std::streamsize>(n * 4));
    }
  }
  f.close();
  std::rename((side + ".tmp").c_str(), side.c_str());
  return true;
}

bool ParameterServerCore::load_checkpoint(const std::string& path, int32_t& epoch) {
  std::lock_guard<std::mutex> lk(mu_);
  std::ifstream f(path, std::ios::binary);
  if (!f.is_open()) return false;
  int32_t iter = 0;
  uint64_t nt = 0;
  f.read(reinterpret_cast<char*>(&epoch), 4);
  f.read(reinterpret_cast<char*>(&iter), 4);
  f.read(reinterpret_cast<char*>(&nt), 8);
  if (!f.good() || nt > (1u << 24)) return false;
  std::vector<HostTensor> loaded;
  loaded.reserve(nt);
  for (uint64_t i = 0; i < nt; ++i) {
    HostTensor t;
    uint64_t nl = 0, nd = 0, ne = 0;
    f.read(reinterpret_cast<char*>(&nl), 8);
    if (!f.good() || nl > (1u << 20)) return false;
    t.name.resize(nl);
    f.read(&t.name[0], static_cast<std::streamsize>(nl));
    f.read(reinterpret_cast<char*>(&nd), 8);
    if (!f.good() || nd > 64) return false;
    t.shape.resize(nd);
    f.read(reinterpret_cast<char*>(t.shape.data()), static_cast<std::streamsize>(nd * 4));
    f.read(reinterpret_cast<char*>(&t.dtype), 4);
    f.read(reinterpret_cast<char*>(&ne), 8);
    if (!f.good() || ne > (1ull << 36)) return false;
    t.data.resize(ne);
    f.read(reinterpret_cast<char*>(t.data.data()), static_cast<std::streamsize>(ne * 4));
    if (!f.good()) return false;  // truncated file (the reference would load garbage)
    loaded.push_back(std::move(t));
  }
  params_ = std::move(loaded);
  current_iteration_.store(iter);
  m_.assign(params_.size(), {});
  v_.assign(params_.size(), {});
  opt_step_ = 0;
  std::ifstream s(path + ".psb", std::ios::binary);
  if (s.is_open()) {
    char magic[4];
    uint32_t ver = 0;
    uint64_t nk = 0, n2 = 0;
    int64_t step = 0;
    s.read(magic, 4);
    s.read(reinterpret_cast<char*>(&ver), 4);
    s.read(reinterpret_cast<char*>(&step), 8);
    s.read(reinterpret_cast<char*>(&nk), 8);
    if (s.good() && memcmp(magic, "PSBS", 4) == 0 && ver == 1 && nk < 64) {
      std::string kind(nk, '\0');
      s.read(&kind[0], static_cast<std::streamsize>(nk));
      s.read(reinterpret_cast<char*>(&n2), 8);
      if (s.good() && n2 == params_.size() && kind == opt_.optimizer) {
        opt_step_ = step;
        for (size_t i = 0; i < params_.size() && s.good(); ++i) {
          for (std::vector<float>* st : {&m_[i], &v_[i]}) {
            uint64_t n = 0;
            s.read(reinterpret_cast<char*>(&n), 8);
            if (!s.good() || n > (1ull << 36)) break;
            st->resize(n);
            s.read(reinterpret_cast<char*>(st->data()), static_cast<std::streamsize>(n * 4));
          }
        }
      }
    }
  }
  if (dev_ && dev_->upload(params_)) dev_->upload_state(m_, v_);
  version_++;
  return true;
}

}  // namespace psb
