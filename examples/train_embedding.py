#!/usr/bin/env python
"""Sparse embedding push / pull (K5) end to end: a categorical-feature classifier whose table is
sharded over the parameter-server ranks, its dense tower trained through the dense engine.

    python examples/train_embedding.py --steps 60                  # one process (CPU or one GPU)
    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 examples/train_embedding.py

Per step a worker pulls only the rows its batch touches and pushes only their gradients; the
owners apply the mean lazily (rows nobody touched keep value and optimizer state).  The reference
has no counterpart: its pushes and pulls always carry the whole model
(/root/reference/src/worker.cpp:240-312).
"""
from __future__ import annotations

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.nn.functional as F

import parameter_server_distributed_b200 as psd
from parameter_server_distributed_b200.models import EmbeddingMLP, EmbeddingMLPConfig
from parameter_server_distributed_b200.parallel import ShardedParameterServer


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--vocab", type=int, default=20000)
    ap.add_argument("--dim", type=int, default=32)
    ap.add_argument("--fields", type=int, default=4)
    ap.add_argument("--lr", type=float, default=0.01)
    args = ap.parse_args()

    ctx = psd.init(heap_bytes=256 << 20)
    cfg = EmbeddingMLPConfig(vocab=args.vocab, dim=args.dim, fields=args.fields, hidden=(64,))
    optim = psd.OptimConfig(kind="adam", lr=args.lr)
    table = psd.ShardedEmbedding.from_context(ctx, cfg.vocab, cfg.dim, optim,
                                              capacity=args.batch * cfg.fields, init_std=0.1)
    torch.manual_seed(0)  # same dense weights on every rank
    dtype = torch.bfloat16 if ctx.device.type == "cuda" else torch.float32
    model = EmbeddingMLP(cfg, table).to(ctx.device, dtype)
    dense = ShardedParameterServer(list(model.named_parameters()), optim, ctx.fabric,
                                   world=ctx.world, rank=ctx.rank, bucket_mb=1)

    # synthetic task with structure: every id carries a fixed hidden score, the label is the
    # sign of the scores' sum, so the table has something to learn row by row
    score = torch.randn(cfg.vocab, generator=torch.Generator().manual_seed(1))
    data = torch.Generator().manual_seed(100 + ctx.rank)  # every rank sees its own batches
    hot = max(64, cfg.vocab // 50)                        # a skewed id distribution, as in practice
    first = last = None
    for step in range(args.steps):
        ids = torch.randint(0, hot, (args.batch, cfg.fields), generator=data)
        cold = torch.rand(args.batch, cfg.fields, generator=data) < 0.2
        ids[cold] = torch.randint(0, cfg.vocab, (int(cold.sum()),), generator=data)
        y = (score[ids].sum(1) > 0).long()
        loss = F.cross_entropy(model(ids.to(ctx.device)).float(), y.to(ctx.device))
        loss.backward()
        dense.step()               # dense tower: bucketed push / pull
        table.apply_gradients()    # table: sparse push of the touched rows
        first = loss.item() if first is None else first
        last = loss.item()
        if ctx.rank == 0 and (step % 10 == 0 or step == args.steps - 1):
            print(f"step {step:4d}  loss {last:.4f}", flush=True)
    if ctx.rank == 0:
        print(f"loss {first:.4f} -> {last:.4f}", flush=True)
    table.close()
    ctx.close()
    return 0 if last < first else 1


if __name__ == "__main__":
    sys.exit(main())
