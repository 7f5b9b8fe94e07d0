// ps_sparse.cpp — see ps_sparse.h.
#include "ps_sparse.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <fstream>
#include <random>
#include <thread>
#include <unordered_map>

namespace psb {

namespace {
// body(lo, hi) over [0, n) on up to `threads` threads (the caller's included)
template <class F>
void parallel_rows(size_t n, int threads, const F& body) {
  const size_t t = std::min<size_t>(static_cast<size_t>(std::max(1, threads)), n / 1024 + 1);
  if (t <= 1) {
    body(size_t{0}, n);
    return;
  }
  std::vector<std::thread> th;
  const size_t step = (n + t - 1) / t;
  for (size_t k = 1; k < t; ++k) {
    const size_t lo = k * step, hi = std::min(n, lo + step);
    if (lo >= hi) break;
    th.emplace_back([&body, lo, hi] { body(lo, hi); });
  }
  body(size_t{0}, std::min(n, step));
  for (auto& x : th) x.join();
}
}  // namespace

SparseTables::SparseTables(const PsOptions& opt) : opt_(opt) {
  int t = opt_.threads > 0 ? opt_.threads : static_cast<int>(std::thread::hardware_concurrency());
  opt_.threads = t < 1 ? 1 : (t > 16 ? 16 : t);
}

bool SparseTables::create(const std::string& name, int64_t rows, int32_t dim, float init_std,
                          int64_t seed) {
  if (name.empty() || rows <= 0 || dim <= 0 || rows > (int64_t{1} << 40) / dim) return false;
  std::lock_guard<std::mutex> lk(mu_);
  auto it = tables_.find(name);
  if (it != tables_.end()) return it->second.rows == rows && it->second.dim == dim;
  Table t;
  t.rows = rows, t.dim = dim;
  t.data.assign(static_cast<size_t>(rows) * dim, 0.0f);
  if (init_std > 0.0f) {
    std::mt19937_64 rng(static_cast<uint64_t>(seed));
    std::normal_distribution<float> normal(0.0f, init_std);
    for (float& x : t.data) x = normal(rng);
  }
  tables_.emplace(name, std::move(t));
  return true;
}

bool SparseTables::empty() {
  std::lock_guard<std::mutex> lk(mu_);
  return tables_.empty();
}

bool SparseTables::valid_locked(const std::vector<SparseGrad>& grads, std::string& err) const {
  for (const auto& g : grads) {
    auto it = tables_.find(g.name);
    if (it == tables_.end()) {
      err = "unknown table '" + g.name + "'";
      return false;
    }
    if (g.dim != it->second.dim ||
        g.values.size() != g.indices.size() * static_cast<size_t>(g.dim)) {
      err = "table '" + g.name + "': width / value count mismatch";
      return false;
    }
  }
  return true;
}

// One row through the configured optimizer; g is scaled by `scale` first (1 / pushers).
void SparseTables::update_row(Table& t, int64_t row, const float* g, float scale) {
  const size_t n = static_cast<size_t>(t.dim), at = static_cast<size_t>(row) * n;
  float* p = t.data.data() + at;
  const std::string& kind = opt_.optimizer;
  const float lr = opt_.lr, wd = opt_.weight_decay;
  if (kind == "sgd") {
    for (size_t k = 0; k < n; ++k) p[k] -= lr * (g[k] * scale + wd * p[k]);
    return;
  }
  float* m = t.m.data() + at;  // state arrays are sized by apply_locked before the row loop
  if (kind == "momentum") {
    for (size_t k = 0; k < n; ++k) {
      m[k] = opt_.momentum * m[k] + (g[k] * scale + wd * p[k]);
      p[k] -= lr * m[k];
    }
    return;
  }
  float* v = t.v.data() + at;
  const bool decoupled = kind == "adamw";
  const float b1 = opt_.beta1, b2 = opt_.beta2;
  const float bias1 = 1.0f - std::pow(b1, static_cast<float>(step_));
  const float bias2 = 1.0f - std::pow(b2, static_cast<float>(step_));
  for (size_t k = 0; k < n; ++k) {
    float gj = g[k] * scale;
    if (decoupled) p[k] *= 1.0f - lr * wd;
    else gj += wd * p[k];
    m[k] = b1 * m[k] + (1.0f - b1) * gj;
    v[k] = b2 * v[k] + (1.0f - b2) * gj * gj;
    const float denom = std::sqrt(v[k]) / std::sqrt(bias2) + opt_.eps;
    p[k] -= (lr / bias1) * (m[k] / denom);
  }
}

// Sum the staged rows per table in ascending worker id (rows in arrival order inside a push),
// then apply every touched row once.  Rows outside the table are ignored (padding ids).
void SparseTables::apply_locked(Iter& it) {
  ++step_;
  const float scale = 1.0f / static_cast<float>(std::max<size_t>(1, it.staged.size()));
  for (auto& tk : tables_) {
    Table& t = tk.second;
    const size_t dim = static_cast<size_t>(t.dim);
    std::unordered_map<int64_t, size_t> slot;
    std::vector<int64_t> touched;
    std::vector<float> acc;
    for (const auto& wk : it.staged)
      for (const auto& g : wk.second) {
        if (g.name != tk.first) continue;
        for (size_t i = 0; i < g.indices.size(); ++i) {
          const int64_t row = g.indices[i];
          if (row < 0 || row >= t.rows) continue;
          auto ins = slot.emplace(row, touched.size());
          if (ins.second) {
            touched.push_back(row);
            acc.resize(acc.size() + dim, 0.0f);
          }
          float* a = acc.data() + ins.first->second * dim;
          const float* src = g.values.data() + i * dim;
          for (size_t k = 0; k < dim; ++k) a[k] += src[k];
        }
      }
    if (touched.empty()) continue;
    // optimizer state appears with the first update (before the threads start)
    if (opt_.optimizer != "sgd" && t.m.size() != t.data.size()) t.m.assign(t.data.size(), 0.0f);
    if ((opt_.optimizer == "adam" || opt_.optimizer == "adamw") && t.v.size() != t.data.size())
      t.v.assign(t.data.size(), 0.0f);
    // touched rows are distinct: their updates are independent (a random walk over the table,
    // bound by memory latency - worth spreading over cores for large batches)
    parallel_rows(touched.size(), opt_.threads, [&](size_t lo, size_t hi) {
      for (size_t s = lo; s < hi; ++s) update_row(t, touched[s], acc.data() + s * dim, scale);
    });
  }
  it.staged.clear();  // staged gradients are not kept after aggregation
  it.applied = true;
  changed_.notify_all();
}

bool SparseTables::push(int32_t worker_id, int32_t iteration, std::vector<SparseGrad>&& grads,
                        int total_workers, int32_t& workers_received, std::string& err) {
  std::lock_guard<std::mutex> lk(mu_);
  workers_received = 0;
  if (!valid_locked(grads, err)) return false;
  const size_t need = static_cast<size_t>(std::max(1, total_workers));
  if (opt_.mode != PsMode::kSync) {
    auto ck = clocks_.find(worker_id);
    if (ck != clocks_.end() && iteration <= ck->second) return true;  // a re-sent RPC
    Iter once;
    once.staged[worker_id] = std::move(grads);
    // one worker's share of the mean: apply_locked divides by the pusher count (1), so scale here
    for (auto& g : once.staged[worker_id])
      for (float& x : g.values) x /= static_cast<float>(need);
    apply_locked(once);
    clocks_[worker_id] = iteration;
    iters_[iteration].applied = true;
    workers_received = 1;
  } else {
    Iter& it = iters_[iteration];
    if (it.applied) return true;  // idempotent retry after completion
    it.staged[worker_id] = std::move(grads);
    workers_received = static_cast<int32_t>(it.staged.size());
    if (it.staged.size() < need) return false;
    apply_locked(it);
  }
  while (iters_.size() > static_cast<size_t>(std::max(8, opt_.keep_iterations)))
    iters_.erase(iters_.begin());
  return true;
}

bool SparseTables::ready(int32_t iteration, int wait_ms) {
  std::unique_lock<std::mutex> lk(mu_);
  auto done = [&] {
    if (iteration < 0) return true;
    auto it = iters_.find(iteration);
    if (it != iters_.end()) return it->second.applied;
    // pruned long ago (everything older than the oldest kept entry was applied)
    return !iters_.empty() && iteration < iters_.begin()->first;
  };
  if (done()) return true;
  if (wait_ms > 0) changed_.wait_for(lk, std::chrono::milliseconds(wait_ms), done);
  return done();
}

bool SparseTables::rows(const std::string& name, const std::vector<int64_t>& indices,
                        std::vector<float>& out, int32_t& dim) {
  std::lock_guard<std::mutex> lk(mu_);
  auto it = tables_.find(name);
  if (it == tables_.end()) return false;
  const Table& t = it->second;
  dim = t.dim;
  const size_t n = static_cast<size_t>(t.dim);
  // an 8-byte id asks for dim * 4 bytes: bound what one request can make the server allocate
  if (indices.size() > (size_t{1} << 30) / n) return false;
  out.assign(indices.size() * n, 0.0f);
  for (size_t i = 0; i < indices.size(); ++i) {
    const int64_t row = indices[i];
    if (row < 0 || row >= t.rows) continue;
    std::copy_n(t.data.data() + static_cast<size_t>(row) * n, n, out.data() + i * n);
  }
  return true;
}

int SparseTables::set_total_workers(int n) {
  std::lock_guard<std::mutex> lk(mu_);
  if (opt_.mode != PsMode::kSync) return 0;
  const size_t need = static_cast<size_t>(std::max(1, n));
  int completed = 0;
  for (auto& kv : iters_) {
    Iter& it = kv.second;
    if (it.applied || it.staged.empty() || it.staged.size() < need) continue;
    apply_locked(it);
    ++completed;
  }
  return completed;
}

// File: "PSBT" u32 version | i64 step | u64 count | per table: u64 name_len, name, i64 rows,
// i32 dim, u8 has_m, u8 has_v, data[, m][, v] (float32, little-endian host).
bool SparseTables::save(const std::string& path) {
  std::lock_guard<std::mutex> lk(mu_);
  const std::string tmp = path + ".tmp";
  {
    std::ofstream f(tmp, std::ios::binary | std::ios::trunc);
    if (!f) return false;
    const uint32_t version = 1;
    const uint64_t count = tables_.size();
    f.write("PSBT", 4);
    f.write(reinterpret_cast<const char*>(&version), 4);
    f.write(reinterpret_cast<const char*>(&step_), 8);
    f.write(reinterpret_cast<const char*>(&count), 8);
    for (const auto& kv : tables_) {
      const Table& t = kv.second;
      const uint64_t len = kv.first.size();
      const uint8_t has_m = t.m.size() == t.data.size(), has_v = t.v.size() == t.data.size();
      f.write(reinterpret_cast<const char*>(&len), 8);
      f.write(kv.first.data(), static_cast<std::streamsize>(len));
      f.write(reinterpret_cast<const char*>(&t.rows), 8);
      f.write(reinterpret_cast<const char*>(&t.dim), 4);
      f.write(reinterpret_cast<const char*>(&has_m), 1);
      f.write(reinterpret_cast<const char*>(&has_v), 1);
      const std::streamsize bytes = static_cast<std::streamsize>(t.data.size() * sizeof(float));
      f.write(reinterpret_cast<const char*>(t.data.data()), bytes);
      if (has_m) f.write(reinterpret_cast<const char*>(t.m.data()), bytes);
      if (has_v) f.write(reinterpret_cast<const char*>(t.v.data()), bytes);
    }
    if (!f) return false;
  }
  return std::rename(tmp.c_str(), path.c_str()) == 0;
}

bool SparseTables::load(const std::string& path) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  char magic[4];
  uint32_t version = 0;
  int64_t step = 0;
  uint64_t count = 0;
  f.read(magic, 4);
  f.read(reinterpret_cast<char*>(&version), 4);
  f.read(reinterpret_cast<char*>(&step), 8);
  f.read(reinterpret_cast<char*>(&count), 8);
  if (!f || std::string(magic, 4) != "PSBT" || version != 1 || count > (1u << 20)) return false;
  std::map<std::string, Table> loaded;
  for (uint64_t i = 0; i < count; ++i) {
    uint64_t len = 0;
    f.read(reinterpret_cast<char*>(&len), 8);
    if (!f || len > 4096) return false;
    std::string name(len, '\0');
    f.read(&name[0], static_cast<std::streamsize>(len));
    Table t;
    uint8_t has_m = 0, has_v = 0;
    f.read(reinterpret_cast<char*>(&t.rows), 8);
    f.read(reinterpret_cast<char*>(&t.dim), 4);
    f.read(reinterpret_cast<char*>(&has_m), 1);
    f.read(reinterpret_cast<char*>(&has_v), 1);
    if (!f || t.rows <= 0 || t.dim <= 0 || t.rows > (int64_t{1} << 40) / t.dim) return false;
    const size_t n = static_cast<size_t>(t.rows) * t.dim;
    const std::streamsize bytes = static_cast<std::streamsize>(n * sizeof(float));
    t.data.resize(n);
    f.read(reinterpret_cast<char*>(t.data.data()), bytes);
    if (has_m) {
      t.m.resize(n);
      f.read(reinterpret_cast<char*>(t.m.data()), bytes);
    }
    if (has_v) {
      t.v.resize(n);
      f.read(reinterpret_cast<char*>(t.v.data()), bytes);
    }
    if (!f) return false;
    loaded.emplace(std::move(name), std::move(t));
  }
  std::lock_guard<std::mutex> lk(mu_);
  tables_ = std::move(loaded);
  step_ = step;
  iters_.clear();
  clocks_.clear();
  changed_.notify_all();
  return true;
}

}  // namespace psb
