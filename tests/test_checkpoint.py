"""H8 / K8: checkpoints of the sharded engine in the reference's byte layout (App. B), optimizer
state in the sidecar, resume equals uninterrupted training, file interchange with the C++ server."""
import copy
import os

import grpc
import numpy as np
import torch

from parameter_server_distributed_b200 import OptimConfig
from parameter_server_distributed_b200.control import protos
from parameter_server_distributed_b200.models import MLP, MLPConfig, mse_loss
from parameter_server_distributed_b200.parallel.engine import ShardedParameterServer
from parameter_server_distributed_b200.utils import checkpoint as ck
from tests._native import BIN, Proc, ensure_binaries, free_port, wait_port

CFG = MLPConfig(d_in=16, hidden=(32,), d_out=4)


def test_golden_bytes(tmp_path):
    p = str(tmp_path / "g.ckpt")
    ck.write_reference_file(p, 0, 4, [("weight", (10, 10), np.full(100, -0.03, np.float32))])
    raw = open(p, "rb").read()
    header = bytes.fromhex("00000000" "04000000" "0100000000000000" "0600000000000000" +
                           b"weight".hex() + "0200000000000000" "0a000000" "0a000000"
                           "00000000" "6400000000000000")
    assert len(raw) == 458 and raw == header + bytes.fromhex("8fc2f5bc") * 100
    assert ck.read_reference_file(p)[:2] == (0, 4)


def _train(m, ps, steps, seed):
    g = torch.Generator().manual_seed(seed)
    for _ in range(steps):
        x, y = torch.randn(8, 16, generator=g), torch.randn(8, 4, generator=g)
        mse_loss(m(x), y).backward()
        ps.step()


def test_resume_equals_uninterrupted(tmp_path):
    opt = OptimConfig(kind="adamw", lr=0.01, weight_decay=0.01)
    torch.manual_seed(0)
    a = MLP(CFG)
    b = copy.deepcopy(a)
    psa = ShardedParameterServer(list(a.named_parameters()), opt, bucket_mb=0.001)
    _train(a, psa, 3, seed=1)
    path = str(tmp_path / "mid.ckpt")
    ck.save_checkpoint(psa, path, epoch=1)
    assert os.path.exists(path + ".psb")
    _train(a, psa, 2, seed=2)
    # second run: fresh engine, load, continue
    psb = ShardedParameterServer(list(b.named_parameters()), opt, bucket_mb=0.001)
    assert ck.load_checkpoint(psb, path) == (1, 2)
    _train(b, psb, 2, seed=2)
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.equal(p, q)


def test_file_loads_in_cpp_server(tmp_path):
    ensure_binaries()
    torch.manual_seed(3)
    m = MLP(CFG)
    ps = ShardedParameterServer(list(m.named_parameters()), OptimConfig(kind="sgd", lr=0.1),
                                bucket_mb=0.001)
    _train(m, ps, 2, seed=5)
    path = str(tmp_path / "model.ckpt")
    ck.save_checkpoint(ps, path, epoch=3)
    port = free_port()
    srv = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{port}", "1", "0"], str(tmp_path),
               str(tmp_path / "ps.log"))
    try:
        assert wait_port(port)
        stub = protos.Stub(grpc.insecure_channel(f"127.0.0.1:{port}"),
                           "parameter_server.ParameterServer")
        r = stub.LoadCheckpoint(protos.msg("parameter_server.LoadCheckpointRequest")(path=path),
                                timeout=10)
        assert r.success and r.epoch == 3
        got = {t.name: np.array(t.data, dtype=np.float32).reshape(list(t.shape)) for t in r.parameters}
        for n, p in m.named_parameters():
            assert np.array_equal(got[n], p.detach().numpy())
        # and back: a file written by the C++ server loads into the engine
        assert stub.SaveCheckpoint(protos.msg("parameter_server.SaveCheckpointRequest")(
            epoch=4, path=str(tmp_path / "cpp.ckpt")), timeout=10).success
        m2 = MLP(CFG)
        ps2 = ShardedParameterServer(list(m2.named_parameters()), OptimConfig(kind="sgd", lr=0.1),
                                     bucket_mb=0.001)
        assert ck.load_checkpoint(ps2, str(tmp_path / "cpp.ckpt"))[0] == 4
        for p, q in zip(m.parameters(), m2.parameters()):
            assert torch.equal(p, q)
    finally:
        srv.stop()


def test_trainer_checkpoint_methods_round_trip(tmp_path):
    """Trainer.save_checkpoint / load_checkpoint: the public spelling of the same functions."""
    import parameter_server_distributed_b200 as psd
    from parameter_server_distributed_b200.models import mse_loss
    ctx = psd.init()
    torch.manual_seed(4)
    model = MLP(CFG)
    tr = psd.Trainer(model, mse_loss, ctx, OptimConfig(kind="adam", lr=0.01), bucket_mb=0.001)
    x, y = torch.randn(8, CFG.d_in), torch.randn(8, CFG.d_out)
    for _ in range(2):
        tr.step(x, y).item()
    path = str(tmp_path / "t.ckpt")
    tr.save_checkpoint(path, epoch=3)
    saved = [p.detach().clone() for p in model.parameters()]
    tr.step(x, y).item()
    assert any(not torch.equal(p, s) for p, s in zip(model.parameters(), saved))
    assert tr.load_checkpoint(path) == (3, 1)
    for p, s in zip(model.parameters(), saved):
        assert torch.equal(p, s)
