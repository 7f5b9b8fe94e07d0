#!/usr/bin/env python
"""Kernel timeline of a few training steps (CUPTI through torch.profiler): which kernel ran on
which stream, when, for how long.  A diagnostic for overlap questions, never a benchmark
number (profiling perturbs timing).  1 GPU:

    python benchmarks/timeline.py [--config mlp] [--steps 2] > gpurun_out/timeline.tsv
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
from torch.profiler import ProfilerActivity, profile

import parameter_server_distributed_b200 as psd
from bench import Workload


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="mlp")
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=4)
    args = ap.parse_args()
    wl = Workload(args.config, 0)
    ctx = psd.init(heap_bytes=wl.nparams * 6 + (512 << 20))
    torch.manual_seed(1234)
    model = wl.build().to(ctx.device, torch.bfloat16)
    trainer = psd.Trainer(model, wl.loss, ctx, psd.OptimConfig(kind="sgd", lr=1e-2))
    gen = torch.Generator().manual_seed(100 + ctx.rank)
    x, y = [t.to(ctx.device) for t in wl.make(gen)]
    for _ in range(args.warmup):
        trainer.step_device(x, y)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(args.steps):
            trainer.step_device(x, y)
        torch.cuda.synchronize()
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "trace.json")
        prof.export_chrome_trace(path)
        events = json.load(open(path))["traceEvents"]
    ker = [e for e in events if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
    ker.sort(key=lambda e: e["ts"])
    if ctx.rank == 0 and ker:
        t0 = ker[0]["ts"]
        print("start_us\tdur_us\tstream\tname")
        for e in ker:
            print(f"{e['ts'] - t0:.1f}\t{e['dur']:.1f}\t{e['args'].get('stream')}\t{e['name'][:90]}")
    ctx.close()


if __name__ == "__main__":
    main()
