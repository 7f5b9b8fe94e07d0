#!/usr/bin/env python
"""Train a torch model through the C++ coordinator + parameter server (host plane) with
``psd.RemoteTrainer`` — the reference's worker loop (/root/reference/src/worker.cpp:331-406) with
real gradients.  Start the two servers, then one process per worker:

    build/bin/coordinator 0.0.0.0:50052 127.0.0.1:50051 &
    build/bin/parameter_server 0.0.0.0:50051 2 0 --optimizer=adam --lr=0.01 &
    python examples/remote_trainer.py --worker-id 0 &
    python examples/remote_trainer.py --worker-id 1

``--mode ssp|async`` needs the server started with the same ``--mode`` (and ``--staleness``).
``--embedding`` trains a categorical-feature classifier instead: its table lives on the server as
a sparse table (``psd.RemoteEmbedding``: only the rows a batch touches are pulled and pushed).
``examples/python_worker.py`` shows the same protocol at the level of single RPCs.
"""
from __future__ import annotations

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

import parameter_server_distributed_b200 as psd


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--coordinator", default="127.0.0.1:50052")
    ap.add_argument("--worker-id", type=int, default=0)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--mode", default="sync", choices=["sync", "async", "ssp"])
    ap.add_argument("--checkpoint", default="", help="ask the server to save here at the end")
    ap.add_argument("--embedding", action="store_true", help="sparse-table model (see above)")
    a = ap.parse_args()
    if a.embedding:
        return embedding_main(a)

    torch.manual_seed(0)  # every worker starts from the same weights
    model = torch.nn.Sequential(torch.nn.Linear(32, 64), torch.nn.GELU(), torch.nn.Linear(64, 8))
    w_true = torch.randn(32, 8, generator=torch.Generator().manual_seed(1)) * 0.3
    data = torch.Generator().manual_seed(100 + a.worker_id)  # every worker sees its own batches

    def loss_fn(m, x, y):
        return torch.nn.functional.mse_loss(m(x), y)

    with psd.RemoteParameterServer(a.coordinator, a.worker_id) as ps:
        print(f"worker {a.worker_id}: parameter server at {ps.ps_address}, "
              f"{ps.total_workers} worker(s) registered", flush=True)
        trainer = psd.RemoteTrainer(model, loss_fn, ps, mode=a.mode)
        for step in range(a.steps):
            x = torch.randn(128, 32, generator=data)
            loss = trainer.step(x, torch.tanh(x @ w_true))
            if step % 10 == 0 or step == a.steps - 1:
                print(f"worker {a.worker_id} iter {trainer.iteration - 1} loss {float(loss):.5f}",
                      flush=True)
        if a.checkpoint and a.worker_id == 0:
            print("checkpoint:", ps.save_checkpoint(trainer.iteration, a.checkpoint), flush=True)
        if trainer.skipped:
            print(f"worker {a.worker_id}: {trainer.skipped} barrier(s) did not open in time")
    return 0


def embedding_main(a) -> int:
    vocab, dim, fields = 5000, 16, 4
    torch.manual_seed(0)
    tower = torch.nn.Sequential(torch.nn.Linear(fields * dim, 32), torch.nn.GELU(),
                                torch.nn.Linear(32, 2))
    score = torch.randn(vocab, generator=torch.Generator().manual_seed(1))  # hidden per-id signal
    data = torch.Generator().manual_seed(100 + a.worker_id)
    with psd.RemoteParameterServer(a.coordinator, a.worker_id) as ps:
        table = psd.RemoteEmbedding(ps, "ids", vocab, dim, init_std=0.1)

        def loss_fn(m, ids, y):
            return torch.nn.functional.cross_entropy(m(table.lookup(ids).reshape(len(ids), -1)), y)

        trainer = psd.RemoteTrainer(tower, loss_fn, ps, mode=a.mode, tables=[table])
        for step in range(a.steps):
            ids = torch.randint(0, vocab // 20, (256, fields), generator=data)  # a hot id range
            loss = trainer.step(ids, (score[ids].sum(1) > 0).long())
            if step % 10 == 0 or step == a.steps - 1:
                print(f"worker {a.worker_id} iter {trainer.iteration - 1} loss {float(loss):.5f}",
                      flush=True)
        if a.checkpoint and a.worker_id == 0:  # writes <path> and <path>.tables
            print("checkpoint:", ps.save_checkpoint(trainer.iteration, a.checkpoint), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
