"""K5 on the GPU: csrc/kernels/sparse.cu (pull, mark / accumulate / apply) against the host path
of the same class (plain PyTorch fp32, tests/test_sparse_host.py pins it to torch.nn.Embedding +
torch.optim).  One rank without a heap, and four emulated ranks of a ``Fabric.local_group`` on
one GPU (every rank's staging and table sit in a different heap segment, addressed through the
peer-pointer tables exactly as over NVLink)."""
import pytest
import torch

from parameter_server_distributed_b200 import OptimConfig, _lib
from parameter_server_distributed_b200.parallel.fabric import Fabric
from parameter_server_distributed_b200.parallel.sparse import ShardedEmbedding

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)


def _batches(world, n, vocab, dim, steps, seed):
    gen = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(steps):
        idx = torch.randint(0, vocab, (world, n), generator=gen)
        idx[:, :n // 8] = idx[0, :n // 8].clone()        # rows that every worker pushes
        idx[:, n // 8:n // 4] = idx[:, n // 4:n // 4 + n // 8].clone()  # duplicates inside one worker
        idx[:, -1] = -1                           # padding entries are ignored
        g = torch.randn(world, n, dim, generator=gen)
        out.append((idx, g))
    return out


def _assert_close(got, want, tol):
    """Sums of three or more duplicates come out of the atomics in another order than out of
    index_add_; where such a sum cancels to ~0, Adam's m / sqrt(v) turns the last bit into an
    lr-sized difference.  Tolerate a vanishing fraction of those, nothing else."""
    bad = (got - want).abs() > tol
    assert bad.float().mean().item() < 1e-4, (got - want).abs().max().item()
    assert (got - want).abs().max().item() < 0.2


@pytest.mark.parametrize("kind,dim,gdt", [("sgd", 16, torch.float32), ("adam", 128, torch.bfloat16),
                                          ("momentum", 512, torch.bfloat16)])
def test_one_rank_matches_the_host_path(kind, dim, gdt):
    vocab, n = 5003, 512
    cfg = OptimConfig(kind=kind, lr=0.05)
    dev = ShardedEmbedding(vocab, dim, cfg, capacity=n, device=DEV, grad_dtype=gdt, seed=4)
    host = ShardedEmbedding(vocab, dim, cfg, capacity=n, grad_dtype=gdt, seed=4)
    assert torch.equal(dev.master.cpu(), host.master)
    launches = _lib.launch_count()
    for idx, g in _batches(1, n, vocab, dim, 3, seed=11):
        dev.push(idx[0].to(DEV), g[0].to(DEV))
        host.push(idx[0], g[0])
    assert _lib.launch_count() - launches >= 9      # mark + accumulate + apply per step
    _assert_close(dev.master.cpu(), host.master, 1e-5 if gdt == torch.float32 else 2e-4)
    q = torch.cat([torch.randint(0, vocab, (300,)), torch.tensor([-1, vocab, 0, vocab - 1])])
    got, want = dev.pull(q.to(DEV)).cpu(), host.pull(q)
    assert got.dtype == torch.bfloat16 and got.shape == (304, dim)
    _assert_close(got.float(), want.float(), 1e-3)
    assert got[300].abs().sum() == 0 and got[301].abs().sum() == 0
    assert torch.equal(got[302], dev.master[0].to(torch.bfloat16).cpu())
    # scratch is back to its resting state: nothing leaks into the next step
    assert int(dev.n_touched.item()) == 0 and bool((dev.slot == -1).all())
    assert float(dev.acc.abs().max()) == 0.0


def test_four_emulated_ranks_match_one_host_worker_with_the_mean_gradient():
    world, vocab, dim, n = 4, 4001, 64, 256
    cfg = OptimConfig(kind="adamw", lr=0.02, weight_decay=0.05)
    fabs = Fabric.local_group(world, DEV, 64 << 20)
    embs = [ShardedEmbedding(vocab, dim, cfg, capacity=n, fabric=f, seed=7) for f in fabs]
    host = ShardedEmbedding(vocab, dim, cfg, capacity=n * world, seed=7, grad_dtype=torch.float32)
    try:
        for idx, g in _batches(world, n, vocab, dim, 3, seed=12):
            for r, e in enumerate(embs):
                e.stage(idx[r].to(DEV), g[r].to(DEV))
            for e in embs:
                e.apply()
            gb = g.to(torch.bfloat16).float()     # what the workers staged
            host.push(idx.reshape(-1), (gb / world).reshape(-1, dim).to(torch.float32))
        torch.cuda.synchronize()
        for r, e in enumerate(embs):
            _assert_close(e.master[:e.hi - e.lo].cpu(), host.master[e.lo:e.hi], 2e-4)
        want = host.table.float()
        for e in (embs[0], embs[3]):              # every rank reads every owner's rows
            _assert_close(e.full_table().cpu().float(), want, 1e-3)
    finally:
        for e in embs:
            e.close()
        for f in reversed(fabs):
            f.close()
