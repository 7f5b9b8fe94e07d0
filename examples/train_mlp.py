#!/usr/bin/env python
"""Minimal end-to-end use of the public API: synthetic-data MLP training through the sharded
parameter server, with a checkpoint in the reference's file format at the end.

    python examples/train_mlp.py --steps 20                       # one process (CPU or one GPU)
    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 8 examples/train_mlp.py

On GPUs every rank is a worker and a parameter-server shard (HBM data plane, no NCCL on the hot
path); without a GPU the same engine runs on the host backend (gloo when WORLD_SIZE > 1).
The reference's equivalent is its `worker_main` loop, which pushes a fabricated gradient
(/root/reference/src/worker.cpp:316-320, 331-406).
"""
from __future__ import annotations

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

import parameter_server_distributed_b200 as psd
from parameter_server_distributed_b200.models import MLP, MLPConfig, mse_loss
from parameter_server_distributed_b200.utils.checkpoint import load_checkpoint, save_checkpoint


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--dims", default="256,512,512,64", help="d_in,hidden...,d_out")
    ap.add_argument("--optimizer", default="adamw", choices=["sgd", "momentum", "adam", "adamw"])
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--clip-norm", type=float, default=None)
    ap.add_argument("--mode", default="sync", choices=["sync", "async", "ssp"])
    ap.add_argument("--checkpoint", default="", help="write (and read back) this file at the end")
    args = ap.parse_args()

    ctx = psd.init(heap_bytes=256 << 20)
    dims = [int(d) for d in args.dims.split(",")]
    torch.manual_seed(0)  # same initial weights on every rank
    model = MLP(MLPConfig(d_in=dims[0], hidden=tuple(dims[1:-1]), d_out=dims[-1]))
    dtype = torch.bfloat16 if ctx.device.type == "cuda" else torch.float32
    model = model.to(ctx.device, dtype)
    # (without a GPU the async / ssp engines run their host emulation: collectives are blocking
    # there, so every clock sees every worker's gradient — the arithmetic of the device path when
    # nobody is ahead)
    trainer = psd.Trainer(model, mse_loss, ctx, psd.OptimConfig(kind=args.optimizer, lr=args.lr),
                          mode=args.mode, staleness=1, clip_norm=args.clip_norm, bucket_mb=4)

    # a fixed random teacher so that the loss has somewhere to go
    gen = torch.Generator().manual_seed(1)
    teacher = torch.randn(dims[0], dims[-1], generator=gen) / dims[0] ** 0.5
    data = torch.Generator().manual_seed(100 + ctx.rank)  # every rank sees its own batches
    first = last = None
    for step in range(args.steps):
        x = torch.randn(args.batch, dims[0], generator=data)
        y = x @ teacher
        x, y = x.to(dtype), y.to(dtype)
        if ctx.device.type == "cuda":
            x, y = x.pin_memory(), y.pin_memory()
        loss = trainer.step(x, y).item()
        first = loss if first is None else first
        last = loss
        if ctx.rank == 0 and (step % 5 == 0 or step == args.steps - 1):
            print(f"step {step:4d}  loss {loss:.5f}", flush=True)
    if args.mode != "sync":
        trainer.ps.drain()
    if args.checkpoint:
        save_checkpoint(trainer.ps, args.checkpoint, epoch=1, iteration=args.steps)
        epoch, iteration = load_checkpoint(trainer.ps, args.checkpoint)
        if ctx.rank == 0:
            print(f"checkpoint {args.checkpoint}: epoch {epoch}, iteration {iteration}", flush=True)
    if ctx.rank == 0:
        print(f"loss {first:.5f} -> {last:.5f}", flush=True)
    ctx.close()
    return 0 if last == last and abs(last) != float("inf") else 1  # a finite loss


if __name__ == "__main__":
    sys.exit(main())
