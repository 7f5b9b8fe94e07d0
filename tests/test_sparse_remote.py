"""Sparse embedding tables on the host plane: psb.ParameterServerExt (CreateTable / PushSparse /
PullRows) answered by build/bin/parameter_server, driven by RemoteParameterServer.  The oracle
is the host path of ShardedEmbedding (itself pinned to torch.nn.Embedding + torch.optim in
tests/test_sparse_host.py) fed with the mean of the workers' gradients."""
import threading

import numpy as np
import pytest
import torch

import parameter_server_distributed_b200 as psd
from parameter_server_distributed_b200.control import protos
from parameter_server_distributed_b200.control.remote import (decode_rows_response,
                                                               encode_rows_request,
                                                               encode_sparse_push)
from tests._native import BIN, Proc, ensure_binaries, free_port, wait_port

ensure_binaries()
V, D, STEPS = 503, 8, 4


def test_numpy_codec_matches_the_protobuf_runtime():
    ids = np.array([5, 0, 2 ** 40, -1], dtype=np.int64)
    vals = np.arange(8, dtype=np.float32).reshape(4, 2) - 3.5
    mine = encode_sparse_push(3, 7, [("emb", 2, ids, vals), ("empty", 2, ids[:0], vals[:0])])
    m = protos.msg("psb.SparsePush").FromString(mine)
    assert (m.worker_id, m.iteration, len(m.tables)) == (3, 7, 2)
    assert list(m.tables[0].indices) == ids.tolist() and list(m.tables[0].values) == vals.ravel().tolist()
    assert m.tables[1].name == "empty" and len(m.tables[1].indices) == 0
    assert m.SerializeToString() == mine            # byte-identical with the runtime's encoding
    Q = protos.msg("psb.RowsRequest")
    for wid, it, ids_ in ((3, -1, ids), (0, 0, ids[:0]), (1, 7, ids[:1])):
        assert encode_rows_request(wid, it, "emb", ids_) == Q(
            worker_id=wid, iteration=it, name="emb", indices=ids_.tolist()).SerializeToString()
    R = protos.msg("psb.RowsResponse")
    ready, dim, got, msg = decode_rows_response(R(ready=True, dim=2, values=[1.5, -2.0]).SerializeToString())
    assert ready and dim == 2 and got.tolist() == [1.5, -2.0] and msg == ""
    assert decode_rows_response(R(message="no").SerializeToString())[3] == "no"


def _batches(worker, world):
    gen = torch.Generator().manual_seed(7)
    out = []
    for _ in range(STEPS):
        idx = torch.randint(0, V, (world, 40), generator=gen)
        idx[:, :5] = idx[0, :5].clone()      # rows pushed by every worker
        idx[:, 5:8] = idx[:, 8:11].clone()   # duplicates inside one push
        idx[:, -1] = -1                       # padding
        idx[:, -2] = V                        # out of range
        out.append((idx, torch.randn(world, 40, D, generator=gen)))
    return out


def _drive(coord, kind, lr, shards=0, lockstep=True):
    """Two worker threads: create, read the initial table, STEPS x (push, pull after it)."""
    res, errs = {}, []
    pushed = threading.Barrier(2)  # async mode: the final read comes after everybody's last push

    def work(w):
        try:
            with psd.RemoteParameterServer(coord, w, heartbeat_s=0.2, shards=shards) as ps:
                ps.create_table("emb", V, D, init_std=0.1, seed=11)
                init = ps.pull_rows("emb", np.arange(V))
                pushed.wait(timeout=60)  # nobody pushes before everybody has read the start
                last = None
                for it, (idx, g) in enumerate(_batches(w, 2)):
                    ps.push_sparse(it, [("emb", idx[w].numpy(), g[w].numpy())])
                    last = ps.pull_rows("emb", idx[w].numpy(), after_iteration=it)
                pushed.wait(timeout=60)
                res[w] = (init, ps.pull_rows("emb", np.arange(-1, V + 1)), last)
        except BaseException as e:  # noqa: BLE001
            pushed.abort()
            errs.append(repr(e))

    ts = [threading.Thread(target=work, args=(w,)) for w in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=120)
    assert not errs, errs
    init, final, last = res[0]
    assert np.array_equal(init, res[1][0]) and np.array_equal(final, res[1][1])
    assert abs(float(init.std()) - 0.1) < 0.01
    # oracle: one worker that pushes both workers' rows with the mean gradient
    cfg = psd.OptimConfig(kind=kind, lr=lr)
    one = psd.ShardedEmbedding(V, D, cfg, capacity=80, grad_dtype=torch.float32)
    one.master.copy_(torch.from_numpy(init))
    for idx, g in _batches(0, 2):
        one.push(idx.reshape(-1), g.reshape(-1, D) / 2)
    assert np.abs(final[1:-1] - one.master.numpy()).max() < 2e-5
    assert not final[0].any() and not final[-1].any()          # ids -1 and V: zero rows
    idx = _batches(0, 2)[-1][0][0].numpy()
    assert np.array_equal(last[idx == -1], np.zeros((1, D), np.float32))
    ok = (idx >= 0) & (idx < V)
    if lockstep:  # sync: the pull after the last push already is the final table
        assert np.array_equal(last[ok], final[1:-1][idx[ok]])


@pytest.fixture()
def plane(tmp_path, request):
    ps_port, co_port = free_port(), free_port()
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "2", "0", *request.param],
              str(tmp_path), str(tmp_path / "ps.log"))
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"))
    assert wait_port(ps_port) and wait_port(co_port)
    yield f"127.0.0.1:{co_port}", ps, tmp_path
    ps.stop()
    co.stop()


@pytest.mark.parametrize("plane,kind,lr", [(["--optimizer=sgd", "--lr=0.5"], "sgd", 0.5),
                                           (["--optimizer=adam", "--lr=0.01"], "adam", 0.01)],
                         indirect=["plane"])
def test_two_workers_through_the_cpp_server_match_the_oracle(plane, kind, lr):
    _drive(plane[0], kind, lr)


@pytest.mark.parametrize("plane", [["--optimizer=sgd", "--lr=0.5", "--mode=async"]], indirect=True)
def test_async_server_applies_every_sparse_push_once(plane):
    """--mode=async: a push is applied at once with a 1/N share.  SGD updates commute, so the
    final table equals the synchronous mean whatever the interleaving of the two workers."""
    _drive(plane[0], "sgd", 0.5, lockstep=False)


@pytest.mark.parametrize("plane", [["--optimizer=momentum", "--lr=0.05"]], indirect=True)
def test_tables_survive_checkpoint_and_reload_and_reject_bad_pushes(plane):
    coord, ps_proc, tmp = plane
    with psd.RemoteParameterServer(coord, 0, heartbeat_s=0.2) as a, \
            psd.RemoteParameterServer(coord, 1, heartbeat_s=0.2) as b:
        a.create_table("emb", V, D, init_std=0.1)
        b.create_table("emb", V, D)                       # same shape: fine
        with pytest.raises(RuntimeError, match="another shape"):
            b.create_table("emb", V, 2 * D)
        ids, g = np.array([3, 4, 3]), np.ones((3, D), np.float32)
        assert a.push_sparse(0, [("emb", ids, g)]) is False   # waits for worker 1
        assert b.push_sparse(0, [("emb", ids[:0], g[:0])]) is True
        # the dense model needs a pusher before a checkpoint exists (bootstrap)
        for ps in (a, b):
            ps.push(0, [("w", np.ones(4, np.float32))])
        before = a.pull_rows("emb", np.arange(V), after_iteration=0)
        path = a.save_checkpoint(1, str(tmp / "t.ckpt"))
        assert (tmp / "t.ckpt.tables").exists()
        a.push_sparse(1, [("emb", ids, g)])
        b.push_sparse(1, [("emb", ids, g)])
        moved = a.pull_rows("emb", ids, after_iteration=1)
        assert np.abs(moved - before[ids]).max() > 1e-3
        a.load_checkpoint(path)
        assert np.array_equal(a.pull_rows("emb", np.arange(V)), before)
        # momentum state came back too: the same push from the same state lands in the same place
        a.push_sparse(0, [("emb", ids, g)])
        b.push_sparse(0, [("emb", ids, g)])
        assert np.allclose(a.pull_rows("emb", ids, after_iteration=0), moved, atol=1e-6)
        # malformed pushes are refused as a whole
        with pytest.raises(RuntimeError, match="unknown table"):
            a._tables["nope"] = (V, D, V)
            a.push_sparse(5, [("nope", ids, g)])


def test_table_rows_split_over_two_parameter_server_shards(tmp_path):
    """shards=2: rows [0, 252) on shard 0, [252, 503) on shard 1; same result as one server."""
    co_port = free_port()
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", "127.0.0.1:1"], str(tmp_path),
              str(tmp_path / "co.log"))
    assert wait_port(co_port)
    servers = []
    try:
        for sid in range(2):
            d = tmp_path / f"shard{sid}"
            d.mkdir()
            port = free_port()
            servers.append(Proc([f"{BIN}/parameter_server", f"127.0.0.1:{port}", "2", "0",
                                 "--optimizer=adam", "--lr=0.01",
                                 f"--coordinator=127.0.0.1:{co_port}", f"--shard-id={sid}",
                                 "--membership-poll-ms=100"], str(d), str(d / "ps.log")))
            assert wait_port(port)
        _drive(f"127.0.0.1:{co_port}", "adam", 0.01, shards=2)
    finally:
        for s in servers:
            s.stop()
        co.stop()


@pytest.mark.parametrize("plane", [["--optimizer=adam", "--lr=0.02"]], indirect=True)
def test_remote_trainer_with_an_embedding_table_learns(plane):
    """RemoteEmbedding + RemoteTrainer: a classifier over categorical ids, dense tower and table
    both on the C++ server; two workers, the loss falls and both end with the same rows."""
    coord = plane[0]
    vocab, dim = 200, 8
    score = torch.randn(vocab, generator=torch.Generator().manual_seed(1))
    res, errs = {}, []
    towers = []
    for _ in range(2):
        torch.manual_seed(0)
        towers.append(torch.nn.Linear(3 * dim, 2))

    def work(w):
        try:
            with psd.RemoteParameterServer(coord, w, heartbeat_s=0.2) as ps:
                table = psd.RemoteEmbedding(ps, "ids", vocab, dim, init_std=0.1)

                def loss_fn(model, ids, y):
                    x = table.lookup(ids).reshape(ids.shape[0], -1)
                    return torch.nn.functional.cross_entropy(model(x), y)

                tr = psd.RemoteTrainer(towers[w], loss_fn, ps, tables=[table])
                gen = torch.Generator().manual_seed(50 + w)
                losses = []
                for _ in range(40):
                    ids = torch.randint(0, vocab, (64, 3), generator=gen)
                    losses.append(float(tr.step(ids, (score[ids].sum(1) > 0).long())))
                res[w] = (losses, ps.pull_rows("ids", np.arange(vocab), after_iteration=table.after))
        except BaseException as e:  # noqa: BLE001
            errs.append(repr(e))

    ts = [threading.Thread(target=work, args=(w,)) for w in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=180)
    assert not errs, errs
    for w in range(2):
        losses = res[w][0]
        assert sum(losses[-5:]) / 5 < 0.8 * sum(losses[:5]) / 5, losses
    assert np.array_equal(res[0][1], res[1][1])


@pytest.mark.parametrize("plane", [["--optimizer=sgd", "--lr=1.0"]], indirect=True)
def test_cpp_workers_train_a_sparse_table_next_to_the_dense_model(plane):
    """worker_main --sparse-rows: the reference worker's fabricated gradient (0.01 everywhere),
    but for the rows of an embedding table — two C++ workers, then the table read from Python:
    rows touched in iteration i moved by -lr * mean(0.01) exactly once, the others not at all."""
    import subprocess
    coord, ps_proc, tmp = plane
    runs = [subprocess.Popen([f"{BIN}/worker_main", coord, str(w), "4", "--sparse-rows=8",
                              "--sparse-table=100x4"], stdout=subprocess.PIPE,
                             stderr=subprocess.STDOUT, text=True) for w in range(2)]
    outs = [r.communicate(timeout=120)[0] for r in runs]
    for r, o in zip(runs, outs):
        assert r.returncode == 0, o
        assert o.count("done=true") == 4, o
        # the row read back after iteration 3 already holds that iteration's update
        assert "iter 3 sparse rows=8 row0[0]=-0.01" in o, o
    with psd.RemoteParameterServer(coord, 0, heartbeat_s=0.2) as ps:
        ps.create_table("embedding", 100, 4)
        table = ps.pull_rows("embedding", np.arange(100))
    assert np.allclose(table[:32], -0.01, atol=1e-7) and not table[32:].any()


@pytest.mark.parametrize("plane", [["--optimizer=adam", "--lr=0.02"]], indirect=True)
def test_remote_trainer_example_with_embedding(plane):
    """examples/remote_trainer.py --embedding, two worker processes; the server writes the model
    and `<path>.tables`."""
    import os
    import subprocess
    import sys
    coord, _, tmp = plane
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, os.path.join(root, "examples", "remote_trainer.py"), "--coordinator", coord,
           "--embedding", "--steps", "12"]
    ckpt = str(tmp / "e.ckpt")
    ps = [subprocess.Popen(cmd + ["--worker-id", str(w)] + (["--checkpoint", ckpt] if w == 0 else []),
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
          for w in range(2)]
    outs = [p.communicate(timeout=180)[0] for p in ps]
    for p, o in zip(ps, outs):
        assert p.returncode == 0, o[-800:]
        assert "iter 12 loss" in o, o[-400:]
    assert os.path.exists(ckpt) and os.path.exists(ckpt + ".tables")


def test_restarted_server_resumes_its_tables(tmp_path):
    """parameter_server --resume=latest: the newest checkpoint's `.tables` file comes back with
    the model (SIGKILL, so nothing is written on the way out)."""
    ps_port, co_port = free_port(), free_port()
    args = [f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "1", "0", "--optimizer=adam", "--lr=0.05"]
    ps = Proc(args, str(tmp_path), str(tmp_path / "ps.log"))
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"))
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        ids, g = np.arange(0, 40, 3), np.ones((14, D), np.float32)
        with psd.RemoteParameterServer(f"127.0.0.1:{co_port}", 0, heartbeat_s=0.2) as w:
            w.create_table("emb", V, D, init_std=0.1)
            w.push(0, [("w", np.ones(4, np.float32))])          # a dense model to checkpoint
            w.push_sparse(0, [("emb", ids, g)])
            before = w.pull_rows("emb", np.arange(V), after_iteration=0)
            w.save_checkpoint(7, "checkpoint_epoch_7.ckpt")      # relative: the server's directory
            w.push_sparse(1, [("emb", ids, g)])                  # not in the checkpoint
        ps.p.kill()
        ps.p.wait(timeout=10)
        ps.f.close()
        ps = Proc(args + ["--resume=latest"], str(tmp_path), str(tmp_path / "ps2.log"))
        assert wait_port(ps_port)
        with psd.RemoteParameterServer(f"127.0.0.1:{co_port}", 0, heartbeat_s=0.2) as w:
            w.create_table("emb", V, D)                          # exists: same shape, kept as loaded
            assert np.array_equal(w.pull_rows("emb", np.arange(V)), before)
            # Adam's state and step count resumed too: the lost push, repeated, moves the rows as before
            w.push_sparse(0, [("emb", ids, g)])
            again = w.pull_rows("emb", ids, after_iteration=0)
            assert np.abs(again - before[ids]).max() > 1e-3
        assert "resumed from checkpoint_epoch_7.ckpt" in ps.output()
    finally:
        ps.stop()
        co.stop()


def test_evicted_worker_stops_blocking_a_pending_sparse_iteration(tmp_path):
    """Two workers, one goes silent before it pushes: once the coordinator has evicted it the
    server (which follows the registry) completes the sparse iteration with the live worker's
    rows alone — the dense barrier's rule, applied to tables."""
    import time
    fast = {"PSB_STALE_SCAN_MS": "100", "PSB_STALE_TIMEOUT_S": "1"}
    ps_port, co_port = free_port(), free_port()
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"), env=fast)
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "2", "0", "--optimizer=sgd",
               "--lr=1.0", f"--coordinator=127.0.0.1:{co_port}", "--membership-poll-ms=100"],
              str(tmp_path), str(tmp_path / "ps.log"))
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        coord = f"127.0.0.1:{co_port}"
        a = psd.RemoteParameterServer(coord, 0, heartbeat_s=0.1)
        b = psd.RemoteParameterServer(coord, 1, heartbeat_s=0.1)
        a.create_table("emb", 50, D)
        time.sleep(0.5)                      # the server has seen the job complete (2 live)
        b.close()                            # worker 1 goes silent without pushing
        ids, g = np.array([1, 2]), np.ones((2, D), np.float32)
        assert a.push_sparse(0, [("emb", ids, g)]) is False          # 1 of 2 so far
        rows = a.pull_rows("emb", ids, after_iteration=0, timeout_s=20)  # opens after the eviction
        assert np.allclose(rows, -1.0)       # divisor = the one live pusher, lr 1
        assert "evicted stale worker 1" in co.output()
        assert "membership: total_workers 2 -> 1" in ps.output()
        a.close()
    finally:
        ps.stop()
        co.stop()
