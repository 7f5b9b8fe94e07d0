#!/usr/bin/env python
"""Training GEMMs of the flagship MLP: the repo's CTA-pair tcgen05 kernel (K-major and MN-major
operands, fused GELU / dGELU epilogues) against the library path of PyTorch (cuBLAS GEMM +
separate elementwise kernels).  One JSON line per measurement -> stdout.

    python benchmarks/linear_bench.py [--batch 4096] [--iters 20]

Timing: CUDA events on the launching stream, 3 warm-up runs, operands rotate over enough
copies to exceed the 126 MB L2.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.nn.functional as F

from parameter_server_distributed_b200.models import MLP, MLPConfig, mse_loss
from parameter_server_distributed_b200.ops import linear, native

DEV = torch.device("cuda:0")


def timed(fn, iters, warmup=3):
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def rot(shape, n):
    return [(torch.randn(*shape, device=DEV) * 0.05).to(torch.bfloat16) for _ in range(n)]


def copies(*shapes):
    byts = sum(2 * s[0] * s[1] for s in shapes)
    return max(2, int(200e6 // byts) + 1)


def bench_layer(batch, k_in, n_out, iters):
    out = []
    flops = 2.0 * batch * k_in * n_out
    n = copies((batch, k_in), (n_out, k_in), (batch, n_out))
    x, w, gy, z = rot((batch, k_in), n), rot((n_out, k_in), n), rot((batch, n_out), n), rot((batch, n_out), n)
    zin = rot((batch, k_in), n)
    bias = torch.randn(n_out, device=DEV)
    bias16 = bias.to(torch.bfloat16)

    def rec(name, ours_ms, lib_ms):
        out.append({"bench": "linear", "op": name, "batch": batch, "k_in": k_in, "n_out": n_out,
                    "ours_ms": ours_ms, "library_ms": lib_ms, "ours_tflops": flops / ours_ms / 1e9,
                    "library_tflops": flops / lib_ms / 1e9, "speedup": lib_ms / ours_ms})

    rec("fwd", timed(lambda i: native.gemm(x[i % n], w[i % n], bias=bias), iters),
        timed(lambda i: F.linear(x[i % n], w[i % n], bias16), iters))
    rec("fwd+gelu", timed(lambda i: native.gemm(x[i % n], w[i % n], bias=bias, epilogue="gelu", aux=z[i % n]), iters),
        timed(lambda i: F.gelu(F.linear(x[i % n], w[i % n], bias16), approximate="tanh"), iters))
    rec("dX", timed(lambda i: native.gemm(gy[i % n], w[i % n], trans_b=True), iters),
        timed(lambda i: gy[i % n] @ w[i % n], iters))

    def lib_dx_dgelu(i):
        zf = zin[i % n].detach().requires_grad_()
        a = F.gelu(zf, approximate="tanh")
        a.backward(gy[i % n] @ w[i % n])

    rec("dX*dgelu", timed(lambda i: native.gemm(gy[i % n], w[i % n], trans_b=True, epilogue="dgelu", aux=zin[i % n]), iters),
        timed(lib_dx_dgelu, iters))
    rec("dW", timed(lambda i: native.gemm(gy[i % n], x[i % n], trans_a=True, trans_b=True), iters),
        timed(lambda i: gy[i % n].t() @ x[i % n], iters))
    return out


def bench_model(batch, iters):
    cfg = MLPConfig()
    model = MLP(cfg).to(DEV, torch.bfloat16)
    xs, ys = rot((batch, cfg.d_in), 4), rot((batch, cfg.d_out), 4)

    def step(i):
        model.zero_grad(set_to_none=True)
        mse_loss(model(xs[i % 4]), ys[i % 4]).backward()

    res = {}
    for name, flag in (("ours", True), ("library", False)):
        linear.ENABLED = flag
        res[name] = timed(step, iters)
    linear.ENABLED = True
    flops = cfg.flops_per_sample() * batch
    return {"bench": "mlp_fwd_bwd", "batch": batch, "dims": cfg.dims, "ours_ms": res["ours"],
            "library_ms": res["library"], "ours_tflops": flops / res["ours"] / 1e9,
            "library_tflops": flops / res["library"] / 1e9, "speedup": res["library"] / res["ours"]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    torch.manual_seed(0)
    d = MLPConfig().dims
    for i in range(len(d) - 1):
        for r in bench_layer(args.batch, d[i], d[i + 1], args.iters):
            print(json.dumps(r), flush=True)
    print(json.dumps(bench_model(args.batch, args.iters)), flush=True)


if __name__ == "__main__":
    main()
