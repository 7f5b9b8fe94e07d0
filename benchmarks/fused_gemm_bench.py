#!/usr/bin/env python
"""K2f: parameter pull fused with the first GEMM, against its baseline and its roofline.

    python -m torch.distributed.run --nproc-per-node N benchmarks/fused_gemm_bench.py

W [Nout, K] bf16 is row-sharded over the N ranks (rank r owns rows [r*Nout/N, (r+1)*Nout/N) of
its replica).  Per shape, device-timed between device barriers, max over ranks:
  fused      psb_gemm_allgather_bf16: ONE kernel gathers the remote slices over NVLink
             (cp.async.bulk peer->smem->local) while tcgen05 computes Y = X W^T tile by tile
  baseline   NCCL all_gather_into_tensor of W, then cuBLAS torch.matmul  ("B1" of BASELINE.md)
  local      the same GEMM with W already local (this repo's kernel and cuBLAS): compute floor
Roofline = max(2*M*Nout*K / measured bf16 peak, (N-1)/N * |W| / 770 GB/s)."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

import parameter_server_distributed_b200 as psd
from parameter_server_distributed_b200 import _lib
from parameter_server_distributed_b200.ops import native
from parameter_server_distributed_b200.utils import flush_l2

try:
    PEAK_TF = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]
except Exception:
    PEAK_TF = 1590.0
NVLINK_GBS = 770.0


def timed(ctx, fn, iters=6, warmup=2):
    for _ in range(warmup):
        fn()
    ts = []
    for _ in range(iters):
        flush_l2(ctx.device)
        ctx.fabric.barrier(2)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    t = torch.tensor(ts, dtype=torch.float64, device=ctx.device)
    if ctx.world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return sorted(t.tolist())


def main():
    shapes = [(1024, 8192, 4096), (4096, 8192, 4096), (16384, 8192, 4096), (4096, 16384, 8192)]
    max_w = max(n * k for _, n, k in shapes) * 2
    ctx = psd.init(heap_bytes=max_w + (64 << 20))
    world, rank = ctx.world, ctx.rank
    if world > 1 and not dist.is_initialized():
        # the product path runs on the coordinator; NCCL exists here only as the baseline
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=ctx.device)
    lib = _lib.load()
    for M, N, K in shapes:
        if N % (256 * world):
            continue
        ctx.fabric.reset_allocations()
        buf = ctx.fabric.alloc(N * K * 2, align=4096)
        wt = buf.view(torch.bfloat16).view(N, K)
        torch.manual_seed(7)
        w_full = (torch.randn(N, K, device=ctx.device) * 0.05).to(torch.bfloat16)
        rows = N // world
        wt.zero_()
        wt[rank * rows:(rank + 1) * rows].copy_(w_full[rank * rows:(rank + 1) * rows])
        x = (torch.randn(M, K, device=ctx.device) * 0.1).to(torch.bfloat16)
        y = torch.empty(M, N, dtype=torch.bfloat16, device=ctx.device)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        a = _lib.PsbGemmAG()
        a.fabric = ctx.fabric.c_fabric()
        for q in range(world):
            a.w[q] = buf.peer_ptr(q)
        a.x, a.bias, a.y = x.data_ptr(), None, y.data_ptr()
        a.M, a.N, a.K = M, N, K
        a.out_dtype, a.bucket, a.epoch, a.skip_wait = _lib.BF16, 0, None, 1
        stream = torch.cuda.current_stream().cuda_stream

        def fused():
            _lib.check(lib.psb_gemm_allgather_bf16(C.byref(a), stream))

        t_fused = timed(ctx, fused)
        torch.cuda.synchronize()
        ok = bool(torch.equal(wt, w_full))
        exp = x.float() @ w_full.float().t()
        err = ((y.float() - exp).abs().max() / exp.abs().max()).item()

        shard = w_full[rank * rows:(rank + 1) * rows].contiguous()
        gathered = torch.empty(N, K, dtype=torch.bfloat16, device=ctx.device)

        def baseline():
            if world > 1:
                dist.all_gather_into_tensor(gathered, shard)
                torch.matmul(x, gathered.t(), out=y)
            else:
                torch.matmul(x, w_full.t(), out=y)

        t_base = timed(ctx, baseline)
        t_local = timed(ctx, lambda: native.gemm_bf16(x, w_full, out=y))
        t_cublas = timed(ctx, lambda: torch.matmul(x, w_full.t(), out=y))
        if rank == 0:
            flops = 2.0 * M * N * K
            link_ms = (world - 1) / world * N * K * 2 / (NVLINK_GBS * 1e6)
            roof_ms = max(flops / (PEAK_TF * 1e9), link_ms)
            med = lambda t: t[len(t) // 2]
            print(json.dumps({
                "M": M, "N": N, "K": K, "n_gpus": world, "weight_mb": N * K * 2 / 2**20,
                "fused_ms": round(med(t_fused), 4), "nccl_allgather_cublas_ms": round(med(t_base), 4),
                "local_gemm_ms": round(med(t_local), 4), "local_cublas_ms": round(med(t_cublas), 4),
                "roofline_ms": round(roof_ms, 4), "bound": "nvlink" if link_ms > flops / (PEAK_TF * 1e9) else "compute",
                "fused_frac_of_roofline": round(roof_ms / med(t_fused), 3),
                "speedup_vs_baseline": round(med(t_base) / med(t_fused), 3),
                "gathered_weight_exact": ok, "rel_err": round(err, 5)}), flush=True)
        del w_full, x, y, gathered
        torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
    ctx.close()


if __name__ == "__main__":
    main()
