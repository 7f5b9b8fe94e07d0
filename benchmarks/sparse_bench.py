#!/usr/bin/env python
"""K5 micro-benchmark: sparse embedding pull and push against the library formulation.

    python benchmarks/sparse_bench.py                                  # one GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 \
        benchmarks/sparse_bench.py                                     # rows spread over 8 GPUs

Device-timed with CUDA events (3 warm-ups, L2 flushed between iterations, max over ranks), one
JSON line per measurement.  `GBps` counts the bytes that have to move: pulled rows x dim x 2
for a pull; for a push the staged rows (read once) plus master / state / bf16 row of every
touched row (read and written).  The library arm (one GPU only) is what a user writes without
this kernel: `torch.unique` + `index_add_` into a dense scratch + a masked Adam on the touched
rows (the arithmetic of `torch.optim.SparseAdam`), and `F.embedding` for the pull.
Without a GPU it runs the host path on a small table (smoke test of the script itself).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

import parameter_server_distributed_b200 as psd


def timed(fn, device, iters=10, warmup=3):
    if device.type == "cuda":
        from parameter_server_distributed_b200.utils import cuda_time_ms
        return sorted(cuda_time_ms(fn, iters=iters, warmup=warmup, device=device))
    out = []
    for i in range(warmup + iters):
        t0 = time.perf_counter()
        fn()
        if i >= warmup:
            out.append((time.perf_counter() - t0) * 1e3)
    return sorted(out)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--vocab", type=int, default=0, help="rows (default 8 Mi on a GPU, 50 k on the host)")
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--rows", type=int, default=0, help="rows per worker and step (default 256 Ki / 4 Ki)")
    ap.add_argument("--optimizer", default="adam")
    args = ap.parse_args()
    ctx = psd.init(heap_bytes=8 << 30)
    cuda = ctx.device.type == "cuda"
    vocab = args.vocab or ((8 << 20) if cuda else 50_000)
    n = args.rows or ((256 << 10) if cuda else 4096)
    dim, dev = args.dim, ctx.device
    cfg = psd.OptimConfig(kind=args.optimizer, lr=1e-3)
    table = psd.ShardedEmbedding.from_context(ctx, vocab, dim, cfg, capacity=n)
    gen = torch.Generator().manual_seed(ctx.rank)
    # a skewed id distribution (half of the draws from 1 % of the rows), as in recommender data
    ids = torch.randint(0, vocab, (n,), generator=gen)
    hot = torch.rand(n, generator=gen) < 0.5
    ids[hot] = torch.randint(0, max(1, vocab // 100), (int(hot.sum()),), generator=gen)
    ids = ids.to(dev)
    grads = torch.randn(n, dim, generator=gen).to(dev, table.grad_dtype)
    state_b = {"sgd": 8, "momentum": 16}.get(args.optimizer, 24)
    uniq = int(torch.unique(ids).numel())

    def line(name, ms, nbytes, **kw):
        med = ms[len(ms) // 2]
        med = ctx.reduce_max(med, key=name)
        if ctx.rank == 0:
            print(json.dumps({"kernel": name, "n_gpus": ctx.world, "vocab": vocab, "dim": dim,
                              "rows_per_worker": n, "unique_rows": uniq, "ms_median": round(med, 4),
                              "GBps_per_gpu": round(nbytes / med / 1e6, 1), **kw}), flush=True)

    line("K5 pull", timed(lambda: table.pull(ids), dev), n * dim * 2)
    gsz = 2 if table.grad_dtype == torch.bfloat16 else 4
    line("K5 push (stage + barriers + mark / accumulate / apply)",
         timed(lambda: table.push(ids, grads), dev), n * dim * gsz + uniq * dim * (state_b + 2 + 8),
         optimizer=args.optimizer)

    if cuda and ctx.world == 1:   # library arm
        w = table.master.clone()
        wb = w.to(torch.bfloat16)
        m, v = torch.zeros_like(w), torch.zeros_like(w)
        line("library pull (F.embedding on the bf16 table)",
             timed(lambda: torch.nn.functional.embedding(ids, wb), dev), n * dim * 2)

        def lib_push():
            u, inv = torch.unique(ids, return_inverse=True)
            g = torch.zeros(u.numel(), dim, device=dev).index_add_(0, inv, grads.float())
            mm = m[u].mul_(0.9).add_(g, alpha=0.1)
            vv = v[u].mul_(0.999).addcmul_(g, g, value=0.001)
            m[u], v[u] = mm, vv
            p = w[u].addcdiv_(mm, vv.sqrt().add_(1e-8), value=-1e-3)
            w[u] = p
            wb[u] = p.to(torch.bfloat16)
        line("library push (unique + index_add_ + lazy Adam on torch ops)", timed(lib_push, dev),
             n * dim * gsz + uniq * dim * (24 + 2 + 8))
    table.close()
    ctx.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
