"""The Python worker of the host plane (control/remote.py): a torch model trained through the
C++ coordinator + parameter server over the reference's wire protocol
(/root/reference/src/worker.cpp:331-406 is the C++ loop it mirrors)."""
import threading

import numpy as np
import pytest
import torch

import parameter_server_distributed_b200 as psd
from tests._native import BIN, Proc, ensure_binaries, free_port, wait_port

ensure_binaries()


def _model():
    torch.manual_seed(3)
    return torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Tanh(), torch.nn.Linear(16, 4))


def _loss(model, x, y):
    return torch.nn.functional.mse_loss(model(x), y)


def _batches(worker, steps):
    g = torch.Generator().manual_seed(100 + worker)
    w_true = torch.randn(8, 4, generator=torch.Generator().manual_seed(9)) * 0.5
    xs = [torch.randn(32, 8, generator=g) for _ in range(steps)]
    return [(x, torch.tanh(x @ w_true)) for x in xs]  # a learnable target, the same for everybody


@pytest.fixture()
def plane(tmp_path, request):
    flags = getattr(request, "param", ["--optimizer=sgd", "--lr=0.1"])
    ps_port, co_port = free_port(), free_port()
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "2", "0", *flags],
              str(tmp_path), str(tmp_path / "ps.log"))
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"))
    assert wait_port(ps_port) and wait_port(co_port)
    yield f"127.0.0.1:{co_port}", ps, co, tmp_path
    ps.stop()
    co.stop()


def _run_workers(coord, steps, mode="sync"):
    out, errs = {}, []
    models = [_model() for _ in range(2)]  # seeded here: torch's global RNG is not per thread

    def work(w):
        try:
            with psd.RemoteParameterServer(coord, w, heartbeat_s=0.2) as ps:
                tr = psd.RemoteTrainer(models[w], _loss, ps, mode=mode)
                losses = [float(tr.step(x, y)) for x, y in _batches(w, steps)]
                tr.sync_model()
                out[w] = (losses, [p.detach().clone() for p in tr.params], tr.skipped,
                          ps.live_workers(), ps.total_workers)
        except BaseException as e:  # noqa: BLE001
            errs.append(repr(e))

    ts = [threading.Thread(target=work, args=(w,)) for w in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=120)
    assert not errs, errs
    return out


def test_two_python_workers_match_the_single_process_oracle(plane):
    """Sync mode: the server applies p -= lr * mean(g) — exactly what one process computes from
    both workers' batches (fp32 on both sides; the wire carries fp32)."""
    coord, ps, co, _ = plane
    steps = 8
    out = _run_workers(coord, steps)
    oracle = _model()
    params = [p for p in oracle.parameters()]
    b = [_batches(w, steps) for w in range(2)]
    for s in range(steps):
        gs = []
        for w in range(2):
            for p in params:
                p.grad = None
            _loss(oracle, *b[w][s]).backward()
            gs.append([p.grad.clone() for p in params])
        with torch.no_grad():
            for i, p in enumerate(params):
                p -= 0.1 * (gs[0][i] + gs[1][i]) / 2
    for w in range(2):
        losses, got, skipped, live, _ = out[w]
        assert skipped == 0 and live == [0, 1]
        assert losses[-1] < losses[0]
        for a, e in zip(got, params):
            assert torch.allclose(a, e, atol=2e-6), (a - e).abs().max()


@pytest.mark.parametrize("plane", [["--mode=ssp", "--staleness=2", "--optimizer=sgd", "--lr=0.05"]],
                         indirect=True)
def test_ssp_python_workers_finish_and_learn(plane):
    coord, ps, co, _ = plane
    out = _run_workers(coord, 12, mode="ssp")
    for w in range(2):
        losses, got, skipped, _, _ = out[w]
        assert skipped == 0 and np.isfinite(losses).all()
        assert np.mean(losses[-3:]) < np.mean(losses[:3])
    # both replicas were refreshed from the same server state after the last push of each
    for a, e in zip(out[0][1], out[1][1]):
        assert torch.allclose(a, e, atol=1e-2)


def test_checkpoint_round_trip_through_the_python_client(plane):
    coord, ps, co, tmp = plane
    out = _run_workers(coord, 3)
    with psd.RemoteParameterServer(coord, 0, heartbeat_s=0.2) as c:
        path = c.save_checkpoint(7, str(tmp / "m.ckpt"))
        before, _ = c.pull(99)
        assert c.load_checkpoint(path) == 7
        after, _ = c.pull(99)
        assert set(before) == set(after) and len(before) == 4
        for k in before:
            np.testing.assert_array_equal(before[k], after[k])
        assert before["0.weight"].shape == (16, 8)


# ------------------------------------------------------------------ host-plane sharding
@pytest.fixture()
def sharded_plane(tmp_path):
    """coordinator + two parameter servers that register as shards 0 and 1 (own working dirs)."""
    co_port = free_port()
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", "127.0.0.1:1"], str(tmp_path),
              str(tmp_path / "co.log"))
    assert wait_port(co_port)
    servers, ports = [], []
    for sid in range(2):
        d = tmp_path / f"shard{sid}"
        d.mkdir()
        port = free_port()
        ports.append(port)
        servers.append(Proc([f"{BIN}/parameter_server", f"127.0.0.1:{port}", "2", "0",
                             "--optimizer=sgd", "--lr=0.1", f"--coordinator=127.0.0.1:{co_port}",
                             f"--shard-id={sid}", "--membership-poll-ms=100"], str(d),
                            str(d / "ps.log")))
    for p in ports:
        assert wait_port(p)
    yield f"127.0.0.1:{co_port}", ports, servers, tmp_path
    for s in servers:
        s.stop()
    co.stop()


def test_model_split_over_two_parameter_server_shards(sharded_plane):
    """tensor k lives on shard k % 2; two Python workers train through both shards and end
    exactly where the single-process oracle ends; each shard holds (and checkpoints) its half."""
    from parameter_server_distributed_b200.utils.checkpoint import read_reference_file
    coord, ports, servers, tmp = sharded_plane
    steps = 6
    out, errs = {}, []
    models = [_model() for _ in range(2)]

    def work(w):
        try:
            with psd.RemoteParameterServer(coord, w, heartbeat_s=0.2, shards=2) as ps:
                assert ps.shard_addresses == [f"127.0.0.1:{p}" for p in ports]
                tr = psd.RemoteTrainer(models[w], _loss, ps)
                for x, y in _batches(w, steps):
                    tr.step(x, y)
                tr.sync_model()
                if w == 0:
                    ps.save_checkpoint(3, "half.ckpt")
                out[w] = ([p.detach().clone() for p in tr.params], tr.skipped)
        except BaseException as e:  # noqa: BLE001
            errs.append(repr(e))

    ts = [threading.Thread(target=work, args=(w,)) for w in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=120)
    assert not errs, errs
    oracle = _model()
    params = list(oracle.parameters())
    b = [_batches(w, steps) for w in range(2)]
    for s in range(steps):
        gs = []
        for w in range(2):
            for p in params:
                p.grad = None
            _loss(oracle, *b[w][s]).backward()
            gs.append([p.grad.clone() for p in params])
        with torch.no_grad():
            for i, p in enumerate(params):
                p -= 0.1 * (gs[0][i] + gs[1][i]) / 2
    for w in range(2):
        got, skipped = out[w]
        assert skipped == 0
        for a, e in zip(got, params):
            assert torch.allclose(a, e, atol=2e-6), (a - e).abs().max()
    names = [[t[0] for t in read_reference_file(str(tmp / f"shard{s}" / "half.ckpt"))[2]]
             for s in range(2)]
    assert names == [["0.weight", "2.weight"], ["0.bias", "2.bias"]]
    for s in servers:
        assert "registered as shard" in s.output()


def test_numpy_codec_agrees_with_the_protobuf_runtime():
    """control/remote.py writes GradientUpdate and reads ParameterUpdate itself (float payload as
    one numpy block); both must be what the protobuf runtime writes / reads."""
    from parameter_server_distributed_b200.control import protos, remote
    rng = np.random.default_rng(0)
    grads = [("layer.weight", rng.standard_normal((3, 5)).astype(np.float32)),
             ("scalar", np.float32(2.5) * np.ones((), np.float32)),
             ("empty", np.zeros((0, 4), np.float32)),
             ("", rng.standard_normal(7).astype(np.float32))]
    for wid, it in ((0, 0), (3, 17), (-1, 2 ** 31 - 1)):
        raw = remote.encode_gradient_update(wid, it, grads)
        m = protos.msg("parameter_server.GradientUpdate").FromString(raw)
        assert (m.worker_id, m.iteration, len(m.gradients)) == (wid, it, 4)
        for t, (name, g) in zip(m.gradients, grads):
            assert t.name == name and list(t.shape) == list(g.shape) and t.dtype == 0
            np.testing.assert_array_equal(np.asarray(t.data, np.float32), g.reshape(-1))
        # byte-identical with the runtime's own serialisation of the same message
        assert raw == m.SerializeToString()
    T = protos.msg("parameter_server.Tensor")
    upd = protos.msg("parameter_server.ParameterUpdate")(
        iteration=9, ready=True, parameters=[
            T(name="a", shape=[2, 3], data=[1, 2, 3, 4, 5, 6], dtype=0),
            T(name="b", shape=[], data=[], dtype=1),
            T(name="c", shape=[4], data=[0.5, -1.5, 2.5, 1e-30])])
    it, ready, ts = remote.decode_parameter_update(upd.SerializeToString())
    assert (it, ready) == (9, True) and [t[0] for t in ts] == ["a", "b", "c"]
    assert ts[0][1] == (2, 3) and ts[1][1] == () and ts[1][2].size == 0
    np.testing.assert_array_equal(ts[0][2], np.arange(1, 7, dtype=np.float32))
    np.testing.assert_array_equal(ts[2][2], np.array([0.5, -1.5, 2.5, 1e-30], np.float32))
    # fields in another order (the C++ server sends iteration / ready first, tensors last)
    head = protos.msg("parameter_server.ParameterUpdate")(iteration=4, ready=True).SerializeToString()
    tail = protos.msg("parameter_server.ParameterUpdate")(parameters=upd.parameters).SerializeToString()
    assert remote.decode_parameter_update(head + tail)[0:2] == (4, True)
    with pytest.raises((ValueError, IndexError)):
        remote.decode_parameter_update(upd.SerializeToString()[:-3])
