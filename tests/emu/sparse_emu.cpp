// sparse_emu.cpp — the device code of csrc/kernels/sparse.cu, compiled for the host and run
// thread by thread (cuda_shim.h), against a dense restatement.  Checks what a GPU is not needed
// for: index arithmetic (owner / local row / sub-warp groups / grid-stride loops), the slot
// protocol (mark -> accumulate -> apply leaves slot / acc / n_touched at rest), lazy updates
// (untouched rows bit-identical), padding, duplicates within and across workers.
// Built and run by tests/test_sparse_kernel_emulation.py, which generates the two .inc files
// from the real sources so that this file never holds a copy of the kernels.
#include "cuda_shim.h"

#include <algorithm>
#include <cstdio>
#include <map>
#include <random>
#include <vector>

namespace psb {
#include "optim_math.inc"
namespace {
#include "sparse_kernels.inc"
}  // namespace
}  // namespace psb
using namespace psb;

namespace {

struct Rank {
  std::vector<__nv_bfloat16> table;
  std::vector<int64_t> stage_idx;
  std::vector<unsigned char> stage_grad;
  uint32_t cnt[64] = {0};
  std::vector<float> master, m, v, acc;
  std::vector<int32_t> slot, touched;
  uint32_t n_touched = 0, step = 0;
};

template <int KIND>
void ref_update(const PsbOptim& o, uint32_t step, float gs, float* p, float* m, float* v,
                const double* g, int dim) {
  const OptimCoef c = make_coef(o, step);
  for (int d = 0; d < dim; ++d) {
    float mm = m ? m[d] : 0.f, vv = v ? v[d] : 0.f;
    update1<KIND>(c, p[d], mm, vv, (float)g[d] * gs);
    if (m) m[d] = mm;
    if (v) v[d] = vv;
  }
}

bool close(float a, float b) { return std::fabs(a - b) <= 1e-5f + 1e-5f * std::fabs(b); }

int fail(const char* what, long a = 0, long b = 0) {
  std::printf("FAIL: %s (%ld, %ld)\n", what, a, b);
  return 1;
}

int run(int world, int64_t vocab, int dim, int cap, int kind, int gdt, int steps, unsigned seed,
        bool tiny_grid) {
  std::mt19937 rng(seed);
  std::normal_distribution<float> normal(0.f, 1.f);
  const int64_t R = (vocab + world - 1) / world;
  const int gsz = gdt == PSB_BF16 ? 2 : 4;
  std::vector<Rank> rk(world);
  std::vector<float> ref_p((size_t)vocab * dim), ref_m(ref_p.size(), 0.f), ref_v(ref_p.size(), 0.f);
  for (auto& x : ref_p) x = 0.02f * normal(rng);
  const bool has_m = kind != PSB_OPT_SGD, has_v = kind == PSB_OPT_ADAM || kind == PSB_OPT_ADAMW;
  for (int r = 0; r < world; ++r) {
    Rank& k = rk[r];
    k.table.assign((size_t)R * dim, __nv_bfloat16{0});
    k.master.assign((size_t)R * dim, 0.f);
    if (has_m) k.m.assign((size_t)R * dim, 0.f);
    if (has_v) k.v.assign((size_t)R * dim, 0.f);
    for (int64_t row = 0; row < R && r * R + row < vocab; ++row)
      for (int d = 0; d < dim; ++d) {
        k.master[row * dim + d] = ref_p[(r * R + row) * dim + d];
        k.table[row * dim + d] = float_to_bf16_rn(k.master[row * dim + d]);
      }
    k.stage_idx.assign(cap, 0);
    k.stage_grad.assign((size_t)cap * dim * gsz, 0);
    k.acc.assign((size_t)world * cap * dim, 0.f);
    k.slot.assign(R, -1);
    k.touched.assign((size_t)world * cap, 0);
  }
  PsbOptim opt = {kind, 0.05f, 0.9f, 0.9f, 0.999f, 1e-8f, 0.01f, 1.0f / world};
  auto make = [&](int r) {
    PsbSparse s;
    std::memset(&s, 0, sizeof(s));
    s.fabric.rank = r, s.fabric.world = world;
    for (int w = 0; w < world; ++w) {
      s.table[w] = rk[w].table.data(), s.stage_idx[w] = rk[w].stage_idx.data();
      s.stage_grad[w] = rk[w].stage_grad.data(), s.stage_cnt[w] = rk[w].cnt;
    }
    s.vocab = vocab, s.rows_per_rank = R, s.dim = dim, s.cap = cap, s.grad_dtype = gdt;
    Rank& k = rk[r];
    s.master = k.master.data(), s.m = has_m ? k.m.data() : nullptr, s.v = has_v ? k.v.data() : nullptr;
    s.slot = k.slot.data(), s.touched = k.touched.data(), s.n_touched = &k.n_touched;
    s.acc = k.acc.data(), s.step = &k.step, s.opt = opt;
    return s;
  };
  const int vecs = dim / 8, lpr = vecs <= 8 ? 8 : (vecs <= 16 ? 16 : 32);
  const int64_t total = (int64_t)world * cap;
  for (int it = 0; it < steps; ++it) {
    // ---- workers stage: random rows, duplicates inside a worker, rows shared by all workers,
    //      padding (-1) and out-of-range (vocab) entries; worker 1 stages fewer than cap
    std::map<int64_t, std::vector<double>> sums;
    for (int w = 0; w < world; ++w) {
      Rank& k = rk[w];
      const int n = w == 1 ? cap / 2 + 1 : cap;
      k.cnt[0] = n;
      for (int i = 0; i < cap; ++i) {  // entries beyond n hold stale garbage on purpose
        int64_t id = (int64_t)(rng() % (uint64_t)vocab);
        if (i < 8) id = (int64_t)((it * 7 + i) % vocab);          // pushed by every worker
        if (i == cap - 1) id = -1;
        if (i == cap - 2) id = vocab;
        k.stage_idx[i] = id;
      }
      for (int i = 8; i < 16 && i + 8 < cap; ++i) k.stage_idx[i] = k.stage_idx[i + 8];  // duplicates
      for (int i = 0; i < cap; ++i)
        for (int d = 0; d < dim; ++d) {
          float g = normal(rng);
          if (gdt == PSB_BF16) {
            __nv_bfloat16 h = float_to_bf16_rn(g);
            std::memcpy(&k.stage_grad[((size_t)i * dim + d) * 2], &h, 2);
            g = bf16_to_float(h);
          } else {
            std::memcpy(&k.stage_grad[((size_t)i * dim + d) * 4], &g, 4);
          }
          const int64_t id = k.stage_idx[i];
          if (i < n && id >= 0 && id < vocab) {
            auto& s = sums[id];
            if (s.empty()) s.assign(dim, 0.0);
            s[d] += g;
          }
        }
    }
    // ---- the staging kernel must reproduce exactly what was written above: hand it the first
    //      n entries as user buffers (fp32 or bf16 gradients), wipe the staged copy, restage
    for (int w = 0; w < world; ++w) {
      Rank& k = rk[w];
      const int64_t n = k.cnt[0];
      const bool in_bf = ((seed + w + it) & 1) != 0;
      const std::vector<int64_t> want_idx(k.stage_idx.begin(), k.stage_idx.begin() + n);
      const std::vector<unsigned char> want_grad(k.stage_grad.begin(),
                                                 k.stage_grad.begin() + (size_t)n * dim * gsz);
      std::vector<unsigned char> user((size_t)n * dim * (in_bf ? 2 : 4) + 16);
      for (int64_t e = 0; e < n * dim; ++e) {
        float g;
        if (gdt == PSB_BF16) {
          __nv_bfloat16 h;
          std::memcpy(&h, &want_grad[e * 2], 2);
          g = bf16_to_float(h);
        } else {
          std::memcpy(&g, &want_grad[e * 4], 4);
          if (in_bf) g = bf16_to_float(float_to_bf16_rn(g));  // what a bf16 caller would hold
        }
        if (in_bf) {
          __nv_bfloat16 h = float_to_bf16_rn(g);
          std::memcpy(&user[e * 2], &h, 2);
        } else {
          std::memcpy(&user[e * 4], &g, 4);
        }
      }
      std::fill(k.stage_idx.begin(), k.stage_idx.begin() + n, (int64_t)-77);
      std::fill(k.stage_grad.begin(), k.stage_grad.begin() + (size_t)n * dim * gsz, (unsigned char)0xAB);
      k.cnt[0] = 0xdead;
      const int64_t work = std::max<int64_t>(n * dim / 8, n);
      const int grid = tiny_grid ? 1 : (int)std::max<int64_t>(1, (work + kThreads - 1) / kThreads);
      if (in_bf && gdt == PSB_BF16)
        emu_launch(sparse_stage_kernel<PSB_BF16, PSB_BF16>, grid, kThreads, want_idx.data(), (const void*)user.data(), n, dim, k.stage_idx.data(), (void*)k.stage_grad.data(), k.cnt);
      else if (in_bf)
        emu_launch(sparse_stage_kernel<PSB_BF16, PSB_F32>, grid, kThreads, want_idx.data(), (const void*)user.data(), n, dim, k.stage_idx.data(), (void*)k.stage_grad.data(), k.cnt);
      else if (gdt == PSB_BF16)
        emu_launch(sparse_stage_kernel<PSB_F32, PSB_BF16>, grid, kThreads, want_idx.data(), (const void*)user.data(), n, dim, k.stage_idx.data(), (void*)k.stage_grad.data(), k.cnt);
      else
        emu_launch(sparse_stage_kernel<PSB_F32, PSB_F32>, grid, kThreads, want_idx.data(), (const void*)user.data(), n, dim, k.stage_idx.data(), (void*)k.stage_grad.data(), k.cnt);
      if (k.cnt[0] != (uint32_t)n) return fail("staged count", k.cnt[0], n);
      if (!std::equal(want_idx.begin(), want_idx.end(), k.stage_idx.begin())) return fail("staged ids", w);
      if (gdt == PSB_F32 && in_bf) {
        // fp32 staging of bf16 inputs holds the bf16-rounded values: restate the reference sums
        for (int64_t i = 0; i < n; ++i) {
          const int64_t id = want_idx[i];
          if (id < 0 || id >= vocab) continue;
          for (int d = 0; d < dim; ++d) {
            float a, b;
            std::memcpy(&a, &want_grad[((size_t)i * dim + d) * 4], 4);
            std::memcpy(&b, &k.stage_grad[((size_t)i * dim + d) * 4], 4);
            if (b != bf16_to_float(float_to_bf16_rn(a))) return fail("staged bf16 -> fp32", w, i);
            sums[id][d] += (double)b - (double)a;
          }
        }
      } else if (!std::equal(want_grad.begin(), want_grad.end(), k.stage_grad.begin())) {
        return fail("staged gradients", w);
      }
    }
    std::vector<Rank> before = rk;  // untouched rows must come out bit-identical
    // ---- owners apply (the launches of psb_sparse_apply)
    for (int r = 0; r < world; ++r) {
      PsbSparse s = make(r);
      emu_launch(sparse_mark_kernel, (int)((total + kThreads - 1) / kThreads), kThreads, s);
      long owned = 0;
      for (auto& kv : sums) owned += kv.first / R == r;
      if ((long)rk[r].n_touched != owned) return fail("touched rows", rk[r].n_touched, owned);
      const int g = tiny_grid ? 1 : grid_for(total, lpr);
      if (gdt == PSB_BF16) emu_launch(sparse_accumulate_kernel<PSB_BF16>, g, kThreads, s);
      else emu_launch(sparse_accumulate_kernel<PSB_F32>, g, kThreads, s);
      switch (kind) {
        case PSB_OPT_SGD: emu_launch(sparse_apply_kernel<PSB_OPT_SGD>, g, kThreads, s); break;
        case PSB_OPT_MOMENTUM: emu_launch(sparse_apply_kernel<PSB_OPT_MOMENTUM>, g, kThreads, s); break;
        case PSB_OPT_ADAM: emu_launch(sparse_apply_kernel<PSB_OPT_ADAM>, g, kThreads, s); break;
        default: emu_launch(sparse_apply_kernel<PSB_OPT_ADAMW>, g, kThreads, s); break;
      }
      rk[r].n_touched = 0;  // cudaMemsetAsync in psb_sparse_apply
    }
    // ---- dense restatement
    for (auto& kv : sums) {
      float* p = &ref_p[kv.first * dim];
      float* m = has_m ? &ref_m[kv.first * dim] : nullptr;
      float* v = has_v ? &ref_v[kv.first * dim] : nullptr;
      switch (kind) {
        case PSB_OPT_SGD: ref_update<PSB_OPT_SGD>(opt, it, opt.grad_scale, p, m, v, kv.second.data(), dim); break;
        case PSB_OPT_MOMENTUM: ref_update<PSB_OPT_MOMENTUM>(opt, it, opt.grad_scale, p, m, v, kv.second.data(), dim); break;
        case PSB_OPT_ADAM: ref_update<PSB_OPT_ADAM>(opt, it, opt.grad_scale, p, m, v, kv.second.data(), dim); break;
        default: ref_update<PSB_OPT_ADAMW>(opt, it, opt.grad_scale, p, m, v, kv.second.data(), dim); break;
      }
    }
    for (int r = 0; r < world; ++r) {
      Rank& k = rk[r];
      k.step += 1;  // psb_counter_add
      for (int64_t row = 0; row < R && r * R + row < vocab; ++row) {
        const int64_t id = r * R + row;
        const bool hit = sums.count(id) != 0;
        for (int d = 0; d < dim; ++d) {
          const float a = k.master[row * dim + d];
          const float b = hit ? ref_p[id * dim + d] : before[r].master[row * dim + d];
          if (hit ? !close(a, b) : a != b) return fail("master", id, d);
          if (has_m) {
            const float x = k.m[row * dim + d];
            const float y = hit ? ref_m[id * dim + d] : before[r].m[row * dim + d];
            if (hit ? !close(x, y) : x != y) return fail("m", id, d);
          }
          if (has_v) {
            const float x = k.v[row * dim + d];
            const float y = hit ? ref_v[id * dim + d] : before[r].v[row * dim + d];
            if (hit ? !close(x, y) : x != y) return fail("v", id, d);
          }
          if (k.table[row * dim + d].bits != float_to_bf16_rn(a).bits) return fail("bf16 row", id, d);
        }
      }
      for (int32_t x : k.slot) if (x != -1) return fail("slot not released", r, x);
      for (float x : k.acc) if (x != 0.f) return fail("accumulator not cleared", r);
    }
  }
  // ---- pull from every rank
  for (int r = 0; r < world; ++r) {
    const int64_t n = 301;
    std::vector<int64_t> idx(n);
    for (auto& x : idx) x = (int64_t)(rng() % (uint64_t)vocab);
    idx[0] = -1, idx[1] = vocab, idx[2] = 0, idx[3] = vocab - 1;
    std::vector<__nv_bfloat16> out((size_t)n * dim, __nv_bfloat16{0x7fc0});
    emu_launch(sparse_pull_kernel, tiny_grid ? 1 : grid_for(n, lpr), kThreads, make(r),
               (const int64_t*)idx.data(), n, out.data());
    for (int64_t i = 0; i < n; ++i)
      for (int d = 0; d < dim; ++d) {
        const int64_t id = idx[i];
        const uint16_t want = id < 0 || id >= vocab ? 0 : rk[id / R].table[(id % R) * dim + d].bits;
        if (out[i * dim + d].bits != want) return fail("pull", i, d);
      }
  }
  return 0;
}

}  // namespace

int main() {
  int bad = 0;
  // world, vocab, dim, cap, kind, grad dtype, steps, seed, one-CTA grid
  bad += run(1, 1003, 16, 64, PSB_OPT_SGD, PSB_F32, 3, 1, false);
  bad += run(3, 1003, 16, 64, PSB_OPT_ADAM, PSB_BF16, 3, 2, false);     // 8 lanes per row
  bad += run(4, 2050, 128, 40, PSB_OPT_MOMENTUM, PSB_BF16, 3, 3, false);  // 16 lanes per row
  bad += run(2, 777, 264, 48, PSB_OPT_ADAMW, PSB_F32, 3, 4, false);     // 33 vectors: 32 lanes, 2 passes
  bad += run(8, 4099, 64, 300, PSB_OPT_ADAM, PSB_BF16, 2, 5, true);     // grid-stride loops
  if (bad) return 1;
  std::printf("sparse kernels (host emulation): ok\n");
  return 0;
}
