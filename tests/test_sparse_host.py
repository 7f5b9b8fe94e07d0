"""K5 sparse embedding push / pull, host path (CPU tensors, gloo): the oracle of the kernels in
csrc/kernels/sparse.cu and the proof that the shard / staging / mean logic is right without a GPU.
Oracles are independent of the implementation: torch.nn.Embedding + torch.optim (dense SGD on
the summed gradient, SparseAdam for the lazy Adam)."""
import multiprocessing as mp
import os
import socket

import pytest
import torch
import torch.distributed as dist

from parameter_server_distributed_b200 import OptimConfig
from parameter_server_distributed_b200.parallel.sparse import ShardedEmbedding, _init_rows

V, D = 1000, 16


def test_init_is_independent_of_the_shard_layout():
    whole = _init_rows(0, 9000, 8, 0.02, seed=3)
    parts = torch.cat([_init_rows(lo, hi, 8, 0.02, seed=3)
                       for lo, hi in ((0, 2250), (2250, 4500), (4500, 6750), (6750, 9000))])
    assert torch.equal(whole, parts)
    assert abs(whole.std().item() - 0.02) < 1e-3


def test_pull_returns_rows_and_zero_rows_for_padding():
    e = ShardedEmbedding(V, D, OptimConfig(kind="sgd", lr=0.1), capacity=64)
    idx = torch.tensor([[0, 5, 999], [-1, 1000, 5]])
    out = e.pull(idx)
    assert out.shape == (2, 3, D) and out.dtype == torch.bfloat16
    assert torch.equal(out[0, 1], e.master[5].to(torch.bfloat16))
    assert torch.equal(out[0, 2], e.master[999].to(torch.bfloat16))
    assert out[1, 0].abs().sum() == 0 and out[1, 1].abs().sum() == 0
    assert torch.equal(out[1, 2], out[0, 1])


def test_sgd_push_matches_dense_embedding_with_duplicates():
    e = ShardedEmbedding(V, D, OptimConfig(kind="sgd", lr=0.5), capacity=256,
                         grad_dtype=torch.float32)
    ref = torch.nn.Embedding(V, D)
    ref.weight.data.copy_(e.master)
    opt = torch.optim.SGD(ref.parameters(), lr=0.5)
    gen = torch.Generator().manual_seed(0)
    for _ in range(4):
        idx = torch.randint(0, V, (200,), generator=gen)
        idx[:20] = idx[20:40]  # duplicates are summed
        w = torch.randn(200, D, generator=gen)
        opt.zero_grad()
        (ref(idx) * w).sum().backward()
        opt.step()
        e.push(idx, w)
    assert (e.master - ref.weight.data).abs().max() < 1e-5
    assert torch.equal(e.table, e.master.to(torch.bfloat16))


def test_lazy_adam_matches_torch_sparse_adam():
    e = ShardedEmbedding(V, D, OptimConfig(kind="adam", lr=0.01), capacity=256,
                         grad_dtype=torch.float32)
    ref = torch.nn.Embedding(V, D, sparse=True)
    ref.weight.data.copy_(e.master)
    opt = torch.optim.SparseAdam(ref.parameters(), lr=0.01)
    gen = torch.Generator().manual_seed(1)
    for _ in range(5):
        idx = torch.randint(0, 50, (64,), generator=gen)  # small range: rows come back
        w = torch.randn(64, D, generator=gen)
        opt.zero_grad()
        (ref(idx) * w).sum().backward()
        opt.step()
        e.push(idx, w)
    # SparseAdam puts eps outside the bias correction: a difference of order eps
    assert (e.master - ref.weight.data).abs().max() < 1e-5
    untouched = torch.ones(V, dtype=torch.bool)
    untouched[:50] = False
    assert e.m[untouched].abs().sum() == 0 and e.v[untouched].abs().sum() == 0


def test_lookup_collects_gradients_of_several_lookups_into_one_push():
    e = ShardedEmbedding(V, D, OptimConfig(kind="sgd", lr=1.0), capacity=64,
                         grad_dtype=torch.float32)
    before = e.master.clone()
    a, b = torch.tensor([1, 2, 2]), torch.tensor([[7], [1]])
    loss = e.lookup(a).float().sum() + 2 * e.lookup(b).float().sum()
    loss.backward()
    e.apply_gradients()
    want = before.clone()
    want[1] -= 1 + 2
    want[2] -= 2
    want[7] -= 2
    assert (e.master - want).abs().max() < 1e-6
    assert e.steps == 1
    e.apply_gradients()  # nothing pending: an empty push is still a (collective) step
    assert e.steps == 2 and (e.master - want).abs().max() < 1e-6


def test_capacity_and_shape_errors():
    e = ShardedEmbedding(V, D, OptimConfig(), capacity=4)
    with pytest.raises(ValueError, match="capacity"):
        e.push(torch.arange(5), torch.zeros(5, D))
    with pytest.raises(ValueError, match="one gradient row"):
        e.push(torch.arange(3), torch.zeros(2, D))
    with pytest.raises(ValueError, match="multiple of 8"):
        ShardedEmbedding(V, 12, OptimConfig(), capacity=4)


def test_state_dict_round_trip_continues_identically():
    cfg = OptimConfig(kind="adamw", lr=0.01, weight_decay=0.1)
    a = ShardedEmbedding(V, D, cfg, capacity=64, grad_dtype=torch.float32)
    gen = torch.Generator().manual_seed(2)
    batches = [(torch.randint(0, 30, (32,), generator=gen), torch.randn(32, D, generator=gen))
               for _ in range(4)]
    for i, g in batches[:2]:
        a.push(i, g)
    b = ShardedEmbedding(V, D, cfg, capacity=64, grad_dtype=torch.float32, seed=9)
    b.load_state_dict(a.state_dict())
    for i, g in batches[2:]:
        a.push(i, g)
        b.push(i, g)
    assert torch.equal(a.master, b.master) and torch.equal(a.table, b.table)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, kind, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = OptimConfig(kind=kind, lr=0.05)
    e = ShardedEmbedding(V + 3, D, cfg, capacity=128, rank=rank, world=world,
                         grad_dtype=torch.float32)       # V + 3: the last shard is shorter
    one = ShardedEmbedding(V + 3, D, cfg, capacity=128 * world, grad_dtype=torch.float32)
    gen = torch.Generator().manual_seed(5)
    for _ in range(3):
        idx = torch.randint(0, V + 3, (world, 100), generator=gen)
        idx[:, :10] = idx[0, :10]                          # rows pushed by every worker
        g = torch.randn(world, 100, D, generator=gen)
        e.push(idx[rank], g[rank])
        one.push(idx.reshape(-1), g.reshape(-1, D) / world)  # oracle: one worker, mean gradient
    full = e.full_table()
    err = (full.float() - one.table.float()).abs().max().item()
    own = (e.master[:e.hi - e.lo] - one.master[e.lo:e.hi]).abs().max().item()
    q.put((rank, err, own))
    dist.destroy_process_group()


@pytest.mark.parametrize("kind,world", [("sgd", 2), ("adam", 3)])
def test_gloo_workers_match_one_worker_with_the_mean_gradient(kind, world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, kind, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err, own in res:
        assert own < 1e-5, (rank, own)
        assert err < 1e-2, (rank, err)   # bf16 rows


@pytest.mark.parametrize("nproc", [1, 2])
def test_example_trains_a_sharded_table_with_a_dense_tower(nproc):
    """examples/train_embedding.py: ShardedEmbedding + EmbeddingMLP + the dense engine end to end
    on the host path; exit code 0 means the loss went down."""
    import re
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = os.path.join(root, "examples", "train_embedding.py")
    cmd = [sys.executable, script] if nproc == 1 else [
        sys.executable, "-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1",
        "--nproc-per-node", str(nproc), script]
    r = subprocess.run(cmd + ["--steps", "30"], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
    assert r.returncode == 0, r.stdout[-600:] + r.stderr[-600:]
    first, last = re.search(r"loss ([0-9.]+) -> ([0-9.]+)", r.stdout).groups()
    assert float(last) < 0.8 * float(first)


def _ckpt_worker(rank, world, port, path, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = OptimConfig(kind="adam", lr=0.05)
    e = ShardedEmbedding(V + 3, D, cfg, capacity=64, rank=rank, world=world, grad_dtype=torch.float32)
    gen = torch.Generator().manual_seed(21)
    for _ in range(2):
        idx = torch.randint(0, V + 3, (world, 50), generator=gen)
        g = torch.randn(world, 50, D, generator=gen)
        e.push(idx[rank], g[rank])
    e.save(path, name="emb", epoch=3)
    q.put((rank, e.full_table().float()))
    dist.destroy_process_group()


def test_whole_table_checkpoint_written_by_three_ranks_loads_at_world_one(tmp_path):
    """save(): one tensor in the reference's file format + sidecar; load() at another world size
    restores rows, Adam state and step count — the next push lands where the writers' would."""
    from parameter_server_distributed_b200.utils.checkpoint import read_reference_file
    path = str(tmp_path / "table.ckpt")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_ckpt_worker, args=(r, 3, port, path, q)) for r in range(3)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    epoch, iteration, tensors = read_reference_file(path)
    assert (epoch, iteration) == (3, 2) and tensors[0][0] == "emb" and tensors[0][1] == (V + 3, D)
    cfg = OptimConfig(kind="adam", lr=0.05)
    one = ShardedEmbedding(V + 3, D, cfg, capacity=256, grad_dtype=torch.float32, seed=5)
    assert one.load(path) == 2
    assert torch.equal(one.table.float(), res[0])
    # the state is the writers': replay the same two steps from scratch and take a third one
    ref = ShardedEmbedding(V + 3, D, cfg, capacity=256, grad_dtype=torch.float32)
    gen = torch.Generator().manual_seed(21)
    for _ in range(2):
        idx = torch.randint(0, V + 3, (3, 50), generator=gen)
        g = torch.randn(3, 50, D, generator=gen)
        ref.push(idx.reshape(-1), g.reshape(-1, D) / 3)
    idx, g = torch.arange(0, 40), torch.ones(40, D)
    one.push(idx, g)
    ref.push(idx, g)
    assert (one.master - ref.master).abs().max() < 1e-5
    with pytest.raises(ValueError, match="expected"):
        ShardedEmbedding(V, D, cfg, capacity=8).load(path)


def test_c_abi_rejects_bad_arguments_before_any_launch():
    """psb_sparse_pull / psb_sparse_apply validate on the host (no GPU needed to see it)."""
    import ctypes as C
    from parameter_server_distributed_b200 import _lib
    lib = _lib.load()
    s = _lib.PsbSparse()
    s.fabric.rank, s.fabric.world = 0, 1
    s.vocab, s.rows_per_rank, s.dim, s.cap = 100, 100, 12, 8
    s.table[0] = 4096
    assert lib.psb_sparse_pull(C.byref(s), 4096, 1, 4096, None) != 0
    assert b"multiple of 8" in lib.psb_last_error()
    s.dim, s.rows_per_rank = 16, 10
    assert lib.psb_sparse_pull(C.byref(s), 4096, 1, 4096, None) != 0
    assert b"do not cover" in lib.psb_last_error()
    s.rows_per_rank = 100
    assert lib.psb_sparse_pull(C.byref(s), None, 0, None, None) == 0      # nothing to pull
    assert lib.psb_sparse_apply(C.byref(s), None) != 0
    assert b"staging range" in lib.psb_last_error()
    s.stage_idx[0] = s.stage_grad[0] = s.stage_cnt[0] = 4096
    assert lib.psb_sparse_apply(C.byref(s), None) != 0
    assert b"owner state" in lib.psb_last_error()
    assert lib.psb_sparse_init(None, 5, None) != 0
    s.master = s.slot = s.touched = s.n_touched = s.acc = 4096
    assert lib.psb_sparse_stage(C.byref(s), 4096, 4096, _lib.F32, 9, None) == -2   # PSB_ERANGE
    assert b"staging capacity 8" in lib.psb_last_error()
    assert lib.psb_sparse_stage(C.byref(s), 4096, 4100, _lib.F32, 2, None) != 0
    assert b"16-byte aligned" in lib.psb_last_error()
    assert lib.psb_sparse_stage(C.byref(s), 4096, 4096, 7, 2, None) != 0


def test_trainer_pushes_table_gradients_next_to_the_dense_ones():
    """psd.Trainer(tables=[...]): one call per step trains the dense tower (engine) and the
    embedding rows (sparse push); ids travel as the model input."""
    import parameter_server_distributed_b200 as psd
    from parameter_server_distributed_b200.models import EmbeddingMLP, EmbeddingMLPConfig
    ctx = psd.init()
    cfg = EmbeddingMLPConfig(vocab=300, dim=8, fields=3, hidden=(16,), classes=2)
    optim = OptimConfig(kind="adam", lr=0.02)
    table = ShardedEmbedding.from_context(ctx, cfg.vocab, cfg.dim, optim, capacity=64 * 3,
                                          init_std=0.1, grad_dtype=torch.float32)
    torch.manual_seed(0)
    model = EmbeddingMLP(cfg, table)
    tr = psd.Trainer(model, lambda out, y: torch.nn.functional.cross_entropy(out.float(), y), ctx,
                     optim, bucket_mb=1, tables=[table])
    score = torch.randn(cfg.vocab, generator=torch.Generator().manual_seed(1))
    gen = torch.Generator().manual_seed(2)
    before = table.master.clone()
    losses = []
    for _ in range(60):
        ids = torch.randint(0, cfg.vocab, (64, cfg.fields), generator=gen)
        losses.append(tr.step(ids, (score[ids].sum(1) > 0).long()).item())
    assert table.steps == 60 and tr.steps == 60
    assert (table.master - before).abs().max() > 1e-2
    assert sum(losses[-10:]) < 0.8 * sum(losses[:10]), (losses[:3], losses[-3:])
    ctx.close()


def test_metrics_count_rows_and_bytes():
    e = ShardedEmbedding(V, D, OptimConfig(), capacity=16, grad_dtype=torch.float32)
    e.pull(torch.arange(5))
    e.push(torch.tensor([1, 2, 2]), torch.zeros(3, D))
    snap = e.metrics.snapshot()
    assert snap["rows_pulled"] == 5 and snap["pull_bytes"] == 5 * D * 2
    assert snap["rows_pushed"] == 3 and snap["pushes"] == 1 and snap["push_bytes"] == 3 * (8 + D * 4)
    assert "psb_sparse_rows_pushed 3" in e.metrics.prometheus()


def test_sparse_benchmark_script_runs_on_the_host_path():
    """benchmarks/sparse_bench.py without a GPU: small table, host path — keeps the script that
    will produce the K5 numbers on a GPU from rotting."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "benchmarks", "sparse_bench.py"),
                        "--vocab", "5000", "--dim", "16", "--rows", "512"], capture_output=True,
                       text=True, timeout=300, env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
    assert r.returncode == 0, r.stdout[-400:] + r.stderr[-400:]
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert [l["kernel"].split()[1] for l in lines] == ["pull", "push"]
    assert all(l["ms_median"] > 0 and l["unique_rows"] <= 512 for l in lines)


def test_trainer_checkpoint_carries_the_tables(tmp_path):
    """Trainer.save_checkpoint / load_checkpoint: the model's file plus one `<path>.table<i>` per
    sparse table; a fresh trainer that loads them continues exactly like the one that wrote them."""
    import parameter_server_distributed_b200 as psd
    from parameter_server_distributed_b200.models import EmbeddingMLP, EmbeddingMLPConfig
    cfg = EmbeddingMLPConfig(vocab=120, dim=8, fields=2, hidden=(8,), classes=2)
    optim = OptimConfig(kind="adam", lr=0.05)
    loss = lambda out, y: torch.nn.functional.cross_entropy(out.float(), y)  # noqa: E731
    gen = torch.Generator().manual_seed(4)
    batches = [(torch.randint(0, cfg.vocab, (16, cfg.fields), generator=gen),
                torch.randint(0, 2, (16,), generator=gen)) for _ in range(6)]

    def make(seed):
        ctx = psd.init()
        table = ShardedEmbedding.from_context(ctx, cfg.vocab, cfg.dim, optim, capacity=64,
                                              init_std=0.1, grad_dtype=torch.float32, seed=seed)
        torch.manual_seed(seed)
        return ctx, table, psd.Trainer(EmbeddingMLP(cfg, table), loss, ctx, optim, bucket_mb=1,
                                       tables=[table])

    ctx, table, tr = make(0)
    for ids, y in batches[:3]:
        tr.step(ids, y).item()
    path = str(tmp_path / "job.ckpt")
    tr.save_checkpoint(path, epoch=1)
    assert os.path.exists(path + ".table0") and os.path.exists(path + ".table0.psb")
    ctx2, table2, tr2 = make(7)            # different initial values everywhere
    tr2.load_checkpoint(path)
    assert torch.equal(table2.master, table.master) and table2.steps == 3
    for ids, y in batches[3:]:
        a, b = tr.step(ids, y).item(), tr2.step(ids, y).item()
        assert abs(a - b) < 1e-6
    assert torch.equal(table2.master, table.master)
