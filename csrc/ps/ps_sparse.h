// ps_sparse.h — sparse embedding tables of the host-plane parameter server.
//
// No counterpart in the reference: its Tensor is dense (proto/parameter_server.proto:19-24) and
// a push or pull always carries the whole model (src/worker.cpp:240-312,
// src/parameter_server.cpp:93-97).  A table lives outside that model (ServeParameters never
// ships it): workers pull the rows a batch touches (PullRows) and push only those rows'
// gradients (PushSparse).  Sync mode stages the pushes of an iteration per worker, completes
// when every live worker has pushed — the rule of ParameterServerCore::receive_gradients,
// src/parameter_server.cpp:35-37 — sums duplicates in ascending worker id, divides by the number
// of pushers and applies the configured optimizer *lazily*: rows without a gradient keep value
// and state.  Async / SSP apply a push at once, scaled by 1 / total_workers, once per clock.
// The GPU plane's counterpart is csrc/kernels/sparse.cu (K5).
#pragma once

#include <condition_variable>
#include <cstdint>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "ps_core.h"

namespace psb {

struct SparseGrad {
  std::string name;
  int32_t dim = 0;
  std::vector<int64_t> indices;
  std::vector<float> values;  // indices.size() * dim
};

class SparseTables {
 public:
  explicit SparseTables(const PsOptions& opt);

  // Idempotent for an identical (rows, dim) spec; false on a conflicting or invalid one.
  bool create(const std::string& name, int64_t rows, int32_t dim, float init_std, int64_t seed);
  // Returns true when `iteration` is (now or already) applied.  Malformed entries (unknown
  // table, wrong width, values.size() != indices.size() * dim) are rejected as a whole: `err`.
  bool push(int32_t worker_id, int32_t iteration, std::vector<SparseGrad>&& grads,
            int total_workers, int32_t& workers_received, std::string& err);
  // Is `iteration` applied?  A negative answer is held up to `wait_ms` for it to change.
  bool ready(int32_t iteration, int wait_ms);
  // Rows of `name` (zero rows for ids outside the table); false if the table does not exist or
  // the answer would exceed 2^30 floats (4 GiB) — split such a pull.
  bool rows(const std::string& name, const std::vector<int64_t>& indices, std::vector<float>& out,
            int32_t& dim);
  // Membership follows the coordinator: pending iterations complete for the new count apply.
  int set_total_workers(int n);
  bool empty();
  // Tables, optimizer state and step count next to a checkpoint (`<path>.tables`), atomically.
  bool save(const std::string& path);
  bool load(const std::string& path);

 private:
  struct Table {
    int64_t rows = 0;
    int32_t dim = 0;
    std::vector<float> data, m, v;
  };
  struct Iter {
    std::map<int32_t, std::vector<SparseGrad>> staged;  // ascending worker id: fixed sum order
    bool applied = false;
  };
  void apply_locked(Iter& it);
  void update_row(Table& t, int64_t row, const float* g, float scale);
  bool valid_locked(const std::vector<SparseGrad>& grads, std::string& err) const;

  PsOptions opt_;
  std::mutex mu_;
  std::condition_variable changed_;
  std::map<std::string, Table> tables_;
  std::map<int32_t, Iter> iters_;
  std::map<int32_t, int32_t> clocks_;  // async / ssp: last clock applied per worker
  int64_t step_ = 0;                   // applied updates (Adam bias correction)
};

}  // namespace psb
