// sparse_emu_lib.cpp — libpsb_kernels' sparse entry points (same C ABI) on the host: the device
// code of csrc/kernels/sparse.cu run thread by thread (cuda_shim.h) behind psb_sparse_init /
// _stage / _pull / _apply.  tests/test_sparse_device_path_emulated.py loads it in place of the
// real library, so that the Python device path — ctypes struct, pointer tables, argument order —
// runs end to end without a GPU.  Launch shapes mirror the host code of sparse.cu.
#include "cuda_shim.h"

#include <algorithm>
#include <cstring>

namespace psb {
#include "optim_math.inc"
namespace {
#include "sparse_kernels.inc"
}  // namespace
}  // namespace psb
using namespace psb;

static int lanes(int dim) { const int v = dim / 8; return v <= 8 ? 8 : (v <= 16 ? 16 : 32); }

extern "C" {

const char* psb_last_error(void) { return "emulated library"; }

int psb_counter_add(uint32_t* counter, uint32_t delta, void*) {
  *counter += delta;
  return 0;
}

int psb_sparse_init(int32_t* slot, int64_t rows, void*) {
  std::memset(slot, 0xFF, (size_t)rows * sizeof(int32_t));
  return 0;
}

int psb_sparse_stage(const PsbSparse* s, const int64_t* idx, const void* grad, int grad_dtype,
                     int64_t n, void*) {
  if (n < 0 || n > s->cap) return PSB_ERANGE;
  const int r = s->fabric.rank;
  int64_t* sidx = const_cast<int64_t*>(reinterpret_cast<const int64_t*>(s->stage_idx[r]));
  void* sgrad = const_cast<void*>(s->stage_grad[r]);
  uint32_t* scnt = const_cast<uint32_t*>(reinterpret_cast<const uint32_t*>(s->stage_cnt[r]));
  const int64_t work = std::max<int64_t>(n * s->dim / 8, n);
  const int grid = (int)std::min<int64_t>(std::max<int64_t>(1, (work + kThreads - 1) / kThreads),
                                          (int64_t)kNumSMs * 8);
  const bool in_bf = grad_dtype == PSB_BF16, out_bf = s->grad_dtype == PSB_BF16;
  if (in_bf && out_bf) emu_launch(sparse_stage_kernel<PSB_BF16, PSB_BF16>, grid, kThreads, idx, grad, n, s->dim, sidx, sgrad, scnt);
  else if (in_bf) emu_launch(sparse_stage_kernel<PSB_BF16, PSB_F32>, grid, kThreads, idx, grad, n, s->dim, sidx, sgrad, scnt);
  else if (out_bf) emu_launch(sparse_stage_kernel<PSB_F32, PSB_BF16>, grid, kThreads, idx, grad, n, s->dim, sidx, sgrad, scnt);
  else emu_launch(sparse_stage_kernel<PSB_F32, PSB_F32>, grid, kThreads, idx, grad, n, s->dim, sidx, sgrad, scnt);
  return 0;
}

int psb_sparse_pull(const PsbSparse* s, const int64_t* idx, int64_t n, void* out, void*) {
  if (n == 0) return 0;
  emu_launch(sparse_pull_kernel, grid_for(n, lanes(s->dim)), kThreads, *s, idx, n,
             reinterpret_cast<__nv_bfloat16*>(out));
  return 0;
}

int psb_sparse_apply(const PsbSparse* s, void*) {
  const int64_t total = (int64_t)s->fabric.world * s->cap;
  const int g = grid_for(total, lanes(s->dim));
  emu_launch(sparse_mark_kernel, (int)((total + kThreads - 1) / kThreads), kThreads, *s);
  if (s->grad_dtype == PSB_BF16) emu_launch(sparse_accumulate_kernel<PSB_BF16>, g, kThreads, *s);
  else emu_launch(sparse_accumulate_kernel<PSB_F32>, g, kThreads, *s);
  switch (s->opt.kind) {
    case PSB_OPT_SGD: emu_launch(sparse_apply_kernel<PSB_OPT_SGD>, g, kThreads, *s); break;
    case PSB_OPT_MOMENTUM: emu_launch(sparse_apply_kernel<PSB_OPT_MOMENTUM>, g, kThreads, *s); break;
    case PSB_OPT_ADAM: emu_launch(sparse_apply_kernel<PSB_OPT_ADAM>, g, kThreads, *s); break;
    default: emu_launch(sparse_apply_kernel<PSB_OPT_ADAMW>, g, kThreads, *s); break;
  }
  *s->n_touched = 0;  // cudaMemsetAsync in the real entry point
  return 0;
}

}  // extern "C"
