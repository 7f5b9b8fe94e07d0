#!/usr/bin/env python
"""Single-GPU kernel micro-benchmarks with roofline fractions (MEASURED_PEAKS.json).

Device-timed with CUDA events, 3 warm-ups, L2 flushed between iterations.  Prints one JSON
line per kernel; `frac` is achieved / measured peak (HBM copy bandwidth or cuBLAS bf16)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

from parameter_server_distributed_b200 import OptimConfig, _lib
from parameter_server_distributed_b200.ops import native
from parameter_server_distributed_b200.utils import cuda_time_ms

DEV = torch.device("cuda:0")
try:
    PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    HBM, TF, SRC = PEAKS["hbm_gbs"], PEAKS["bf16_tflops"], "measured"
except Exception:
    HBM, TF, SRC = 6650.0, 1590.0, "fallback"


def report(name, ms, bytes_=None, flops=None, **kw):
    ms = sorted(ms)
    med = ms[len(ms) // 2]
    out = {"kernel": name, "ms_median": round(med, 4), "ms_min": round(ms[0], 4), **kw}
    if bytes_:
        out["GBps"] = round(bytes_ / med / 1e6, 1)
        out["frac_of_%s_hbm" % SRC] = round(bytes_ / med / 1e6 / HBM, 3)
    if flops:
        out["TFLOPs"] = round(flops / med / 1e9, 1)
        out["frac_of_%s_bf16" % SRC] = round(flops / med / 1e9 / TF, 3)
    print(json.dumps(out), flush=True)


def bench_optim(n=256 << 20):
    for kind, state_bytes in [("sgd", 4 + 4), ("momentum", 16), ("adamw", 24)]:
        cfg = OptimConfig(kind=kind, lr=1e-3)
        master = torch.randn(n, device=DEV)
        m = torch.zeros(n, device=DEV)
        v = torch.zeros(n, device=DEV)
        g = torch.randn(n, device=DEV).to(torch.bfloat16)
        out = torch.empty(n, dtype=torch.bfloat16, device=DEV)
        ms = cuda_time_ms(lambda: native.optim_step(cfg, master, g, m, v, out), device=DEV)
        report(f"K3 optim {kind} n={n}", ms, bytes_=n * (state_bytes + 2 + 2))
        if kind == "adamw":
            p = torch.nn.Parameter(master.clone())
            p.grad = g.float()
            opt = torch.optim.AdamW([p], lr=1e-3, fused=True)
            ms = cuda_time_ms(opt.step, device=DEV)
            report(f"torch fused AdamW (fp32 grad, no bf16 emit) n={n}", ms, bytes_=n * (24 + 4))
        del master, m, v, g, out


def bench_gemm():
    for M, N, K in [(4096, 8192, 4096), (4096, 8192, 8192), (8192, 8192, 8192), (1024, 8192, 4096),
                    (4096, 4096, 1024), (16384, 4096, 4096)]:
        x = (torch.randn(M, K, device=DEV) * 0.1).to(torch.bfloat16)
        w = (torch.randn(N, K, device=DEV) * 0.1).to(torch.bfloat16)
        y = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
        fl = 2.0 * M * N * K
        for variant, label in ((1, "cta_group::2"), (0, "cta_group::1")):
            _lib.check(_lib.load().psb_gemm_select(variant))
            ms = cuda_time_ms(lambda: native.gemm_bf16(x, w, out=y), device=DEV)
            report(f"K2f tcgen05 gemm {label} {M}x{N}x{K}", ms, flops=fl)
        _lib.check(_lib.load().psb_gemm_select(1))
        ms = cuda_time_ms(lambda: torch.matmul(x, w.t(), out=y), device=DEV)
        report(f"cuBLAS torch.matmul {M}x{N}x{K}", ms, flops=fl)


def bench_misc(n=512 << 20):
    x = torch.randn(n, device=DEV).to(torch.bfloat16)
    ms = cuda_time_ms(lambda: native.sumsq(x), device=DEV)
    report(f"K4 sumsq bf16 n={n}", ms, bytes_=n * 2)
    f = torch.empty(n // 2, device=DEV)
    ms = cuda_time_ms(lambda: native.fill_(f, 0.01), device=DEV)
    report(f"K7 fill f32 n={n // 2}", ms, bytes_=n * 2)
    dst = torch.empty(n // 2, dtype=torch.bfloat16, device=DEV)
    ms = cuda_time_ms(lambda: native.cast_scale(f, dst, 0.5), device=DEV)
    report(f"K6 cast f32->bf16 n={n // 2}", ms, bytes_=(n // 2) * 6)


if __name__ == "__main__":
    which = sys.argv[1:] or ["optim", "gemm", "misc"]
    if "optim" in which:
        bench_optim()
    if "gemm" in which:
        bench_gemm()
    if "misc" in which:
        bench_misc()
