#!/usr/bin/env python
"""BASELINE.json config 5: push/pull bandwidth sweep, 1 MB - 4 GB parameter vectors at 2/4/8
GPUs, fused kernels (K1 / K2) vs the strong NCCL baseline B1 (reduce_scatter + fused
torch optimizer step on the shard + all_gather).

    python -m torch.distributed.run --nproc-per-node N benchmarks/bw_sweep.py [--max-mb 4096]

Device-timed (CUDA events between device-side barriers), max over ranks, L2 flushed between
iterations.  One JSON line per (size, variant) on rank 0.  Roofline: the bytes that must
enter each GPU over NVLink, (N-1)/N * S per direction, at the measured 770 GB/s peer-copy
bandwidth (B200_PROFILING.md)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

import parameter_server_distributed_b200 as psd
from parameter_server_distributed_b200 import OptimConfig
from parameter_server_distributed_b200.parallel.buckets import plan_buckets
from parameter_server_distributed_b200.parallel.device_backend import DeviceBackend
from parameter_server_distributed_b200.utils import flush_l2

NVLINK_GBS = 770.0


def timed(ctx, fn, iters, warmup=2):
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
        dist.all_reduce(t, op=dist.ReduceOp.MAX)  # max over ranks, per iteration
    return sorted(t.tolist())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-mb", type=float, default=1)
    ap.add_argument("--max-mb", type=float, default=4096)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--optimizer", default="sgd")
    ap.add_argument("--grids", default="0")
    ap.add_argument("--tma-grids", default="16,24,32,48",
                    help="CTA counts of the bulk-copy K1 variant (empty = skip)")
    args = ap.parse_args()
    max_elems = int(args.max_mb * (1 << 20) / 2)
    ctx = psd.init(heap_bytes=max_elems * 4 + (64 << 20))
    world, rank = ctx.world, ctx.rank
    if world > 1 and not dist.is_initialized():
        # the product path runs on the coordinator; NCCL exists here only as the baseline B1
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=ctx.device)
    cfg = OptimConfig(kind=args.optimizer, lr=1e-3, grad_scale=1.0 / world)
    mb = args.min_mb
    sizes = []
    while mb <= args.max_mb:
        sizes.append(mb)
        mb *= 4
    if sizes[-1] != args.max_mb:
        sizes.append(args.max_mb)
    # one plan sized for the largest vector; smaller sizes reuse a prefix via fresh plans
    for mb in sizes:
        elems = int(mb * (1 << 20) / 2)
        plan = plan_buckets([("vec", (elems,))], world, bucket_elems=1 << 62)
        # re-use the heap: reset the bump allocator past the signal pad
        ctx.fabric.reset_allocations()
        for grid in [int(g) for g in args.grids.split(",")]:
            be = DeviceBackend(ctx.fabric, plan, cfg, broadcast=True, push_grid=grid)
            ctx.fabric.reset_allocations()
            be.grad_flat.normal_()
            S = plan.total_elems * 2
            link_bytes = (world - 1) / world * S
            floor_ms = link_bytes / (NVLINK_GBS * 1e6) if world > 1 else 0.0
            hbm_bytes = plan.shard_elems * (2 * world + 8 + 2)
            be.epoch.zero_()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()

            def emit(variant, ts, nv_bytes):
                if rank != 0:
                    return
                med = ts[len(ts) // 2]
                out = {"size_mb": mb, "n_gpus": world, "variant": variant, "grid": grid,
                       "ms_median": round(med, 4), "ms_min": round(ts[0], 4),
                       "optimizer": args.optimizer}
                if world > 1:
                    out["nvlink_in_GBps_per_gpu"] = round(nv_bytes / med / 1e6, 1)
                    out["frac_of_770GBps"] = round(nv_bytes / med / 1e6 / NVLINK_GBS, 3)
                else:
                    out["hbm_GBps"] = round(hbm_bytes / med / 1e6, 1)
                print(json.dumps(out), flush=True)

            # fused push with push-mode all-gather: grads in + params in per GPU
            be.broadcast = True
            ts = timed(ctx, lambda: be.push(0), args.iters)
            emit("K1 push+update+broadcast", ts, 2 * link_bytes)
            # pull mode: K1 (reduce-scatter + update) then K2 (all-gather)
            if grid == 0 and args.tma_grids:
                be.push_tma = True
                for tg in [int(g) for g in args.tma_grids.split(",")]:
                    be._push_args[0].grid = tg
                    ts = timed(ctx, lambda: be.push(0), args.iters)
                    emit(f"K1-TMA push+update+broadcast grid={tg}", ts, 2 * link_bytes)
                be.push_tma = False
                be._push_args[0].grid = grid
            if grid == 0 and world > 1 and be.nvls_available:
                # the switch reduces and replicates: S/N + (N-1)/N*S = S bytes in per GPU
                for ng in (0, 16, 32):
                    be._push_args[0].grid = ng
                    ts = timed(ctx, lambda: be.push(0, nvls=True), args.iters)
                    emit(f"K1-NVLS push+update+broadcast grid={ng or 'auto'}", ts, S)
                be._push_args[0].grid = grid
                ts = timed(ctx, lambda: be.bcast_slice_nvls(0), args.iters)
                emit("K2-NVLS all-gather (multimem.st)", ts, link_bytes)
            be.broadcast = False
            ts = timed(ctx, lambda: be.push(0), args.iters)
            emit("K1 push+update (no broadcast)", ts, link_bytes)
            ts = timed(ctx, lambda: be.pull(0), args.iters)
            emit("K2 pull all-gather", ts, link_bytes)
            del be
        # ---- B1: NCCL reduce_scatter + torch fused optimizer + all_gather
        if world > 1:
            g = torch.randn(plan.total_elems, device=ctx.device).to(torch.bfloat16)
            shard_g = torch.empty(plan.shard_elems, dtype=torch.bfloat16, device=ctx.device)
            master = torch.nn.Parameter(torch.zeros(plan.shard_elems, device=ctx.device))
            full_p = torch.empty(plan.total_elems, dtype=torch.bfloat16, device=ctx.device)
            opt = (torch.optim.SGD([master], lr=1e-3) if args.optimizer == "sgd"
                   else torch.optim.AdamW([master], lr=1e-3, fused=True))

            def nccl_step():
                dist.reduce_scatter_tensor(shard_g, g, op=dist.ReduceOp.AVG)
                master.grad = shard_g.float()
                opt.step()
                dist.all_gather_into_tensor(full_p, master.detach().to(torch.bfloat16))

            ts = timed(ctx, nccl_step, args.iters)
            if rank == 0:
                med = ts[len(ts) // 2]
                print(json.dumps({"size_mb": mb, "n_gpus": world,
                                  "variant": "B1 NCCL reduce_scatter + torch optim + all_gather",
                                  "ms_median": round(med, 4), "ms_min": round(ts[0], 4),
                                  "nvlink_in_GBps_per_gpu": round(2 * (world - 1) / world * plan.total_elems * 2 / med / 1e6, 1),
                                  "optimizer": args.optimizer}), flush=True)
            del g, shard_g, master, full_p, opt
        torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
    ctx.close()


if __name__ == "__main__":
    main()
