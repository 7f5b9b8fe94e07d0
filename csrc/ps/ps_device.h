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
  // Allocates the slots and copies the parameters in.  On any allocation or copy failure the
  // slots are released and the shard disables itself: update() then answers false and the
  // caller's host path takes over.  Returns whether the shard is usable.
  bool upload(const std::vector<HostTensor>& params);
  // Queues copy + K3 for one tensor; sync() once per optimizer step waits for all of them
  // (one stream synchronisation per push instead of one per tensor).
  bool update(size_t idx, const float* grad, size_t n, int64_t opt_step);
  bool sync();
  // false: a copy failed and `params` may be stale (the caller keeps its host copy)
  bool download(std::vector<HostTensor>& params);
  // Optimizer moments for the checkpoint sidecar: device -> host before a save, host -> device
  // after a load (sizes that do not match a slot are left alone).
  bool download_state(std::vector<std::vector<float>>& m, std::vector<std::vector<float>>& v);
  bool upload_state(const std::vector<std::vector<float>>& m,
                    const std::vector<std::vector<float>>& v);
  bool active() const;

 private:
  DeviceShard() = default;
  struct Impl;
  Impl* impl_ = nullptr;
};

}  // namespace psb
