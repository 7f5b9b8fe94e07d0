// ps_device.cpp — see ps_device.h.
#include "ps_device.h"

#include <dlfcn.h>
#include <unistd.h>

#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>

#include "../../include/psb_kernels.h"

namespace psb {

struct DeviceShard::Impl {
  void* lib = nullptr;
  PsOptions opt;
  decltype(&psb_set_device) set_device = nullptr;
  decltype(&psb_heap_alloc) heap_alloc = nullptr;
  decltype(&psb_heap_free) heap_free = nullptr;
  decltype(&psb_memcpy_h2d) h2d = nullptr;
  decltype(&psb_memcpy_d2h) d2h = nullptr;
  decltype(&psb_optim_step) optim_step = nullptr;
  decltype(&psb_stream_sync) stream_sync = nullptr;
  decltype(&psb_device_count) device_count = nullptr;
  decltype(&psb_last_error) last_error = nullptr;
  struct Slot {
    float* master = nullptr;
    float* m = nullptr;
    float* v = nullptr;
    float* grad = nullptr;
    uint32_t* step = nullptr;
    size_t n = 0;
  };
  std::vector<Slot> slots;
  bool dirty = false;  // device copy is newer than the host copy
  bool disabled = false;       // allocation / copy failure: the host path owns the shard
  bool failed_update = false;  // a launch failed: retire the GPU copy at the next download

  void free_slots() {
    for (auto& s : slots) {
      for (void* p : {(void*)s.master, (void*)s.m, (void*)s.v, (void*)s.grad, (void*)s.step})
        if (p) heap_free(p);
    }
    slots.clear();
  }
};

static std::string default_lib_path() {
  if (const char* e = std::getenv("PSB_KERNELS_LIB")) return e;
  char exe[4096];
  ssize_t n = readlink("/proc/self/exe", exe, sizeof(exe) - 1);
  if (n <= 0) return "libpsb_kernels.so";
  exe[n] = 0;
  std::string p(exe);
  size_t cut = p.rfind("/build/bin/");
  if (cut == std::string::npos) return "libpsb_kernels.so";
  return p.substr(0, cut) + "/parameter_server_distributed_b200/_native/libpsb_kernels.so";
}

DeviceShard* DeviceShard::create(const PsOptions& opt) {
  void* lib = dlopen(default_lib_path().c_str(), RTLD_NOW | RTLD_LOCAL);
  if (!lib) {
    std::cerr << "device shard disabled: " << dlerror() << std::endl;
    return nullptr;
  }
  auto* im = new Impl;
  im->lib = lib;
  im->opt = opt;
#define PSB_SYM(field, name) im->field = reinterpret_cast<decltype(im->field)>(dlsym(lib, #name))
  PSB_SYM(set_device, psb_set_device);
  PSB_SYM(heap_alloc, psb_heap_alloc);
  PSB_SYM(heap_free, psb_heap_free);
  PSB_SYM(h2d, psb_memcpy_h2d);
  PSB_SYM(d2h, psb_memcpy_d2h);
  PSB_SYM(optim_step, psb_optim_step);
  PSB_SYM(stream_sync, psb_stream_sync);
  PSB_SYM(device_count, psb_device_count);
  PSB_SYM(last_error, psb_last_error);
#undef PSB_SYM
  int count = 0;
  if (!im->device_count || !im->optim_step || im->device_count(&count) != 0 ||
      opt.device >= count || im->set_device(opt.device) != 0) {
    std::cerr << "device shard disabled: GPU " << opt.device << " is not available" << std::endl;
    dlclose(lib);
    delete im;
    return nullptr;
  }
  auto* d = new DeviceShard;
  d->impl_ = im;
  return d;
}

DeviceShard::~DeviceShard() {
  if (!impl_) return;
  impl_->set_device(impl_->opt.device);
  impl_->free_slots();
  dlclose(impl_->lib);
  delete impl_;
}

bool DeviceShard::active() const { return impl_ && !impl_->disabled; }

bool DeviceShard::upload(const std::vector<HostTensor>& params) {
  Impl& im = *impl_;
  im.set_device(im.opt.device);
  im.free_slots();
  im.disabled = false;
  im.slots.resize(params.size());
  const bool need_m = im.opt.optimizer != "sgd";
  const bool need_v = im.opt.optimizer == "adam" || im.opt.optimizer == "adamw";
  bool ok = true;
  // every allocation is checked: psb_heap_alloc leaves *ptr untouched on failure, so a reused
  // pointer would silently alias two buffers
  auto alloc = [&](size_t bytes) -> void* {
    void* p = nullptr;
    if (!ok || im.heap_alloc(bytes, &p) != 0 || !p) {
      ok = false;
      return nullptr;
    }
    return p;
  };
  for (size_t i = 0; i < params.size() && ok; ++i) {
    Impl::Slot& s = im.slots[i];
    s.n = params[i].data.size();
    if (s.n == 0) continue;
    const size_t bytes = s.n * sizeof(float);
    s.master = static_cast<float*>(alloc(bytes));
    s.grad = static_cast<float*>(alloc(bytes));
    if (need_m) s.m = static_cast<float*>(alloc(bytes));
    if (need_v) s.v = static_cast<float*>(alloc(bytes));
    s.step = static_cast<uint32_t*>(alloc(256));
    if (ok && im.h2d(s.master, params[i].data.data(), bytes, nullptr) != 0) ok = false;
  }
  if (ok && im.stream_sync(nullptr) != 0) ok = false;
  if (!ok) {
    std::cerr << "device shard disabled (falling back to the host path): " << im.last_error()
              << std::endl;
    im.free_slots();
    im.disabled = true;
  }
  im.dirty = false;
  return ok;
}

bool DeviceShard::update(size_t idx, const float* grad, size_t n, int64_t opt_step) {
  Impl& im = *impl_;
  if (im.disabled) return false;
  if (idx >= im.slots.size() || im.slots[idx].n == 0 || n != im.slots[idx].n) return false;
  Impl::Slot& s = im.slots[idx];
  im.set_device(im.opt.device);
  PsbOptim o{};
  const std::string& k = im.opt.optimizer;
  o.kind = k == "sgd" ? PSB_OPT_SGD : k == "momentum" ? PSB_OPT_MOMENTUM
           : k == "adam" ? PSB_OPT_ADAM : PSB_OPT_ADAMW;
  o.lr = im.opt.lr, o.momentum = im.opt.momentum, o.beta1 = im.opt.beta1, o.beta2 = im.opt.beta2;
  o.eps = im.opt.eps, o.weight_decay = im.opt.weight_decay, o.grad_scale = 1.0f;
  const uint32_t completed = static_cast<uint32_t>(opt_step > 0 ? opt_step - 1 : 0);
  int rc = im.h2d(s.step, &completed, sizeof(completed), nullptr);
  if (rc == 0) rc = im.h2d(s.grad, grad, n * sizeof(float), nullptr);
  if (rc == 0)
    rc = im.optim_step(&o, s.master, s.m, s.v, s.grad, PSB_F32, nullptr, n, nullptr, s.step,
                       nullptr);
  // No synchronisation here: a copy from pageable host memory has left the caller's buffer
  // when the call returns, and the update kernels of one step queue up behind each other;
  // ParameterServerCore::apply_locked waits once per step (sync()).
  if (rc != 0) {
    // The device copy of this tensor can no longer be trusted as the newest one.  Bring back
    // what is still readable, then stop using the GPU: the host path continues from there.
    std::cerr << "device update failed (" << im.last_error() << "): shard falls back to the host"
              << std::endl;
    im.failed_update = true;
    return false;
  }
  im.dirty = true;
  return true;
}

bool DeviceShard::sync() {
  Impl& im = *impl_;
  if (im.disabled) return true;
  im.set_device(im.opt.device);
  if (im.stream_sync(nullptr) == 0) return true;
  std::cerr << "device shard: update step failed (" << im.last_error()
            << "): falling back to the host copy" << std::endl;
  im.failed_update = true;
  return false;
}

bool DeviceShard::download(std::vector<HostTensor>& params) {
  Impl& im = *impl_;
  if (im.disabled || !im.dirty) return true;
  im.set_device(im.opt.device);
  bool ok = true;
  for (size_t i = 0; i < params.size() && i < im.slots.size(); ++i) {
    if (im.slots[i].n == 0 || im.slots[i].n != params[i].data.size()) continue;
    ok = ok && im.d2h(params[i].data.data(), im.slots[i].master, im.slots[i].n * sizeof(float),
                      nullptr) == 0;
  }
  ok = ok && im.stream_sync(nullptr) == 0;
  if (!ok) {
    std::cerr << "device download failed: " << im.last_error() << std::endl;
    return false;
  }
  im.dirty = false;
  if (im.failed_update) {  // host copy is current again: retire the GPU copy for good
    im.free_slots();
    im.disabled = true;
  }
  return true;
}

bool DeviceShard::download_state(std::vector<std::vector<float>>& m,
                                 std::vector<std::vector<float>>& v) {
  Impl& im = *impl_;
  if (im.disabled) return true;
  im.set_device(im.opt.device);
  bool ok = true;
  for (size_t i = 0; i < im.slots.size(); ++i) {
    const Impl::Slot& s = im.slots[i];
    if (s.n == 0) continue;
    if (s.m && i < m.size()) {
      m[i].resize(s.n);
      ok = ok && im.d2h(m[i].data(), s.m, s.n * sizeof(float), nullptr) == 0;
    }
    if (s.v && i < v.size()) {
      v[i].resize(s.n);
      ok = ok && im.d2h(v[i].data(), s.v, s.n * sizeof(float), nullptr) == 0;
    }
  }
  return ok && im.stream_sync(nullptr) == 0;
}

bool DeviceShard::upload_state(const std::vector<std::vector<float>>& m,
                               const std::vector<std::vector<float>>& v) {
  Impl& im = *impl_;
  if (im.disabled) return true;
  im.set_device(im.opt.device);
  bool ok = true;
  for (size_t i = 0; i < im.slots.size(); ++i) {
    const Impl::Slot& s = im.slots[i];
    if (s.n == 0) continue;
    if (s.m && i < m.size() && m[i].size() == s.n)
      ok = ok && im.h2d(s.m, m[i].data(), s.n * sizeof(float), nullptr) == 0;
    if (s.v && i < v.size() && v[i].size() == s.n)
      ok = ok && im.h2d(s.v, v[i].data(), s.n * sizeof(float), nullptr) == 0;
  }
  return ok && im.stream_sync(nullptr) == 0;
}

}  // namespace psb
