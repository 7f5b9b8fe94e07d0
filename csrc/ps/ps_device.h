// ps_device.h — optional HBM residency of a host-path PS shard (parameter_server --device=N).
// The master weights live on the GPU; the aggregated gradient is staged host -> device and the
// update runs as kernel K3 (psb_optim_step) from libpsb_kernels.so, loaded with dlopen so the
// binary still builds and runs on a box without CUDA.  No counterpart in the reference.
#pragma once

#include <cstddef>
#include <cstdint>
#include <vector>

#include "ps_core.h"

namespace psb {

class DeviceShard {
 public:
  static DeviceShard* create(const PsOptions& opt);  // nullptr when CUDA / the library is missing
  ~DeviceShard();
  void upload(const std::vector<HostTensor>& params);
  bool update(size_t idx, const float* grad, size_t n, int64_t opt_step);
  void download(std::vector<HostTensor>& params);

 private:
  DeviceShard() = default;
  struct Impl;
  Impl* impl_ = nullptr;
};

}  // namespace psb
