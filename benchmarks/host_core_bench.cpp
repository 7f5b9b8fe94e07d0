// host_core_bench — server-side aggregation + update on the CPU: this repo's
// psb::ParameterServerCore against the reference's ParameterServerCore, same workload, same box.
//
// The reference core compiles stand-alone from its unmodified sources
// (/root/reference/src/parameter_server.cpp, include/parameter_server.h — BASELINE.md §3), so this
// is a like-for-like timing of what a ReceiveGradients RPC does after decoding: stage the
// worker's gradient, and on the completing push compute the mean over workers and apply it
// (src/parameter_server.cpp:18-91).  W workers push T tensors of n floats each for K iterations;
// reported per iteration: total time inside receive_gradients (all W calls) and the time of the
// completing call alone.  The final parameters of the two cores are compared bit for bit.
//
//   g++ -std=c++17 -O2 -pthread -I baseline/_ref/include benchmarks/host_core_bench.cpp \
//       baseline/_ref/src/parameter_server.cpp csrc/ps/ps_core.cpp csrc/ps/ps_device.cpp -ldl \
//       -o build/bin/host_core_bench          (benchmarks/host_core_bench.sh does this)
//   build/bin/host_core_bench [workers=4] [iters=5] [scale=1.0] [optimizer=sgd]
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../csrc/ps/ps_core.h"
#include "parameter_server.h"  // the reference's (global namespace)

namespace {

using clk = std::chrono::steady_clock;
double ms_since(clk::time_point t0) {
  return std::chrono::duration<double, std::milli>(clk::now() - t0).count();
}

struct Shape {
  const char* name;
  int rows, cols;
};
// BASELINE config 2: MLP 4096-8192-8192-1024 (weights + biases, 109 M parameters at scale 1)
const Shape kModel[] = {{"layers.0.weight", 8192, 4096}, {"layers.0.bias", 8192, 1},
                        {"layers.1.weight", 8192, 8192}, {"layers.1.bias", 8192, 1},
                        {"layers.2.weight", 1024, 8192}, {"layers.2.bias", 1024, 1}};

template <class T>
std::vector<T> make(int worker, int iteration, double scale, bool params) {
  std::vector<T> out;
  uint32_t s = 0x9E3779B9u * static_cast<uint32_t>(worker * 131 + iteration * 7 + (params ? 1 : 2));
  for (const Shape& sh : kModel) {
    T t;
    t.name = sh.name;
    const int rows = std::max(1, static_cast<int>(sh.rows * scale));
    t.shape = sh.cols == 1 ? std::vector<int32_t>{rows} : std::vector<int32_t>{rows, sh.cols};
    t.dtype = 0;
    const size_t n = static_cast<size_t>(rows) * sh.cols;
    t.data.resize(n);
    for (size_t j = 0; j < n; ++j) {
      s = s * 1664525u + 1013904223u;
      t.data[j] = (static_cast<int32_t>(s >> 8) - (1 << 23)) * (1.0f / (1 << 23)) * 0.01f;
    }
    out.push_back(std::move(t));
  }
  return out;
}

struct Timing {
  double total_ms = 0, last_ms = 0;
};

// What the service does with a decoded push: the reference's API takes a const reference and
// copies into its staging map; ours can take ownership of the decoded tensors.
bool push(ParameterServerCore& core, int w, int it, std::vector<tensor>& decoded) {
  return core.receive_gradients(w, it, decoded);
}
bool push(psb::ParameterServerCore& core, int w, int it, std::vector<psb::HostTensor>& decoded) {
  return core.receive_gradients(w, it, std::move(decoded));
}

template <class Core, class T>
Timing run(Core& core, int workers, int iters, double scale) {
  Timing tm;
  // one gradient set per worker; every push gets its own copy made OUTSIDE the timed region —
  // that copy stands for the tensors the RPC layer has just decoded
  std::vector<std::vector<T>> grads;
  for (int w = 0; w < workers; ++w) grads.push_back(make<T>(w, 0, scale, false));
  for (int it = 0; it < iters; ++it) {
    for (int w = 0; w < workers; ++w) {
      std::vector<T> decoded = grads[w];
      const auto t0 = clk::now();
      const bool done = push(core, w, it, decoded);
      const double ms = ms_since(t0);
      if (it > 0) {  // iteration 0 is the bootstrap (the first aggregate becomes the model)
        tm.total_ms += ms;
        if (w == workers - 1) tm.last_ms += ms;
      }
      if (done != (w == workers - 1)) {
        std::fprintf(stderr, "unexpected completion state at it %d w %d\n", it, w);
        std::exit(1);
      }
    }
  }
  tm.total_ms /= std::max(1, iters - 1);
  tm.last_ms /= std::max(1, iters - 1);
  return tm;
}

}  // namespace

int main(int argc, char** argv) {
  const int workers = argc > 1 ? std::atoi(argv[1]) : 4;
  const int iters = argc > 2 ? std::atoi(argv[2]) : 5;
  const double scale = argc > 3 ? std::atof(argv[3]) : 1.0;
  const std::string optimizer = argc > 4 ? argv[4] : "sgd";
  size_t n = 0;
  for (const auto& t : make<tensor>(0, 0, scale, false)) n += t.data.size();

  Timing ref_t{}, our_t{};
  bool identical = true;
  double max_diff = 0.0;
  const bool compat = optimizer == "sgd";
  std::vector<tensor> ref_params;
  if (compat) {
    ParameterServerCore ref(workers);
    ref_t = run<ParameterServerCore, tensor>(ref, workers, iters, scale);
    ref_params = ref.serve_parameters(0);
  }
  {
    psb::PsOptions opt;
    opt.optimizer = optimizer;
    if (!compat) opt.lr = 0.001f;
    psb::ParameterServerCore ours(workers, opt);
    our_t = run<psb::ParameterServerCore, psb::HostTensor>(ours, workers, iters, scale);
    if (compat) {
      const auto p = ours.serve_parameters(0);
      // The reference sums the workers in the iteration order of a std::unordered_map
      // (src/parameter_server.cpp:49-55), this repo in ascending worker id: bit-identical for
      // <= 2 workers, equal up to fp32 summation order beyond that.
      bool same_shape = p.size() == ref_params.size();
      for (size_t i = 0; same_shape && i < p.size(); ++i) {
        same_shape = p[i].data.size() == ref_params[i].data.size();
        if (!same_shape) break;
        if (std::memcmp(p[i].data.data(), ref_params[i].data.data(),
                        p[i].data.size() * sizeof(float)) != 0)
          identical = false;
        for (size_t j = 0; j < p[i].data.size(); ++j) {
          const double d = std::fabs(static_cast<double>(p[i].data[j]) - ref_params[i].data[j]);
          if (d > max_diff) max_diff = d;
        }
      }
      if (!same_shape) {
        std::fprintf(stderr, "parameter shapes differ\n");
        return 1;
      }
    }
  }
  std::printf("{\"bench\": \"host_core\", \"workers\": %d, \"iters\": %d, \"params\": %zu, "
              "\"optimizer\": \"%s\", \"ours_ms_per_iter\": %.2f, \"ours_completing_push_ms\": %.2f",
              workers, iters, n, optimizer.c_str(), our_t.total_ms, our_t.last_ms);
  if (compat)
    std::printf(", \"reference_ms_per_iter\": %.2f, \"reference_completing_push_ms\": %.2f, "
                "\"speedup\": %.2f, \"bit_identical\": %s, \"max_abs_diff\": %.3g",
                ref_t.total_ms, ref_t.last_ms, ref_t.total_ms / our_t.total_ms,
                identical ? "true" : "false", max_diff);
  std::printf("}\n");
  return (workers <= 2 && !identical) || max_diff > 1e-6 ? 1 : 0;
}
