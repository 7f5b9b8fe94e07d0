"""Multi-rank worker for the sparse embedding path (one rank per GPU under torch.distributed.run;
also runs on CPU ranks over gloo, which is how its own logic is tested without GPUs).
Every rank pushes its own rows for a few steps — rows shared by all ranks, duplicates, padding —
then compares every owner's rows with a one-worker host oracle fed with the mean gradient.
Prints 'MP_OK <rank>' on success."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

import parameter_server_distributed_b200 as psd
from parameter_server_distributed_b200 import OptimConfig


def main():
    ctx = psd.init(heap_bytes=256 << 20)
    vocab, dim, n = 20011, 64, 1024
    for kind in ("sgd", "adam"):
        cfg = OptimConfig(kind=kind, lr=0.05)
        table = psd.ShardedEmbedding.from_context(ctx, vocab, dim, cfg, capacity=n, seed=3)
        oracle = psd.ShardedEmbedding(vocab, dim, cfg, capacity=n * ctx.world, seed=3,
                                       grad_dtype=torch.float32)  # takes the mean as it is
        gen = torch.Generator().manual_seed(17)
        for _ in range(3):
            idx = torch.randint(0, vocab, (ctx.world, n), generator=gen)
            idx[:, :64] = idx[0, :64].clone()
            idx[:, 64:128] = idx[:, 128:192].clone()
            idx[:, -1] = -1
            g = torch.randn(ctx.world, n, dim, generator=gen)
            table.push(idx[ctx.rank].to(ctx.device), g[ctx.rank].to(ctx.device))
            gb = g.to(torch.bfloat16).float()  # what the workers staged (bf16 gradients)
            oracle.push(idx.reshape(-1), (gb / ctx.world).reshape(-1, dim))
        own = (table.master[:table.hi - table.lo].cpu() - oracle.master[table.lo:table.hi]).abs()
        bad = (own > 2e-4).float().mean().item()
        assert bad < 1e-4 and own.max().item() < 0.2, (kind, bad, own.max().item())
        full = (table.full_table().cpu().float() - oracle.table.float()).abs()
        assert (full > 1e-3).float().mean().item() < 1e-4 and full.max().item() < 0.2, \
            (kind, (full > 1e-3).float().mean().item(), full.max().item())
        probe = torch.tensor([0, vocab - 1, -1, vocab, table.lo, max(table.lo, table.hi - 1)])
        got = table.pull(probe.to(ctx.device)).cpu().float()
        assert got[2].abs().sum() == 0 and got[3].abs().sum() == 0
        assert (got[0] - oracle.table[0].float()).abs().max() < 1e-3
        table.close()
        ctx.barrier()
    print(f"MP_OK {ctx.rank}", flush=True)
    ctx.close()


if __name__ == "__main__":
    try:
        main()
    except BaseException as e:
        import traceback
        tb = traceback.extract_tb(e.__traceback__)[-1]
        print(f"MP_FAIL rank={os.environ.get('RANK')} {type(e).__name__}: {str(e)[:400]} "
              f"({os.path.basename(tb.filename)}:{tb.lineno})", flush=True)
        os._exit(1)
