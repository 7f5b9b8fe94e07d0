"""The C++ binaries with --device: the host-path parameter server keeps its master weights in
HBM and updates them with kernel K3, worker_main fills its synthetic gradients with kernel K7.
Both load libpsb_kernels.so with dlopen; the wire behaviour must not change."""
import subprocess

import grpc
import numpy as np
import pytest
import torch

from parameter_server_distributed_b200.control import protos
from tests._native import BIN, Proc, ensure_binaries, free_port, wait_port

pytestmark = pytest.mark.gpu
M = protos.msg


@pytest.mark.parametrize("kind", ["sgd", "adamw"])
def test_device_resident_ps_matches_host_ps(tmp_path, kind):
    ensure_binaries()
    ports = [free_port(), free_port()]
    flags = [f"--optimizer={kind}", "--lr=0.05", "--weight-decay=0.01"]
    procs = [Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ports[0]}", "2", "0", *flags],
                  str(tmp_path), str(tmp_path / "host.log")),
             Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ports[1]}", "2", "0", *flags,
                   "--device=0"], str(tmp_path), str(tmp_path / "dev.log"))]
    try:
        assert all(wait_port(p) for p in ports)
        stubs = [protos.Stub(grpc.insecure_channel(f"127.0.0.1:{p}"),
                             "parameter_server.ParameterServer") for p in ports]
        rng = np.random.default_rng(0)
        n = 5000
        for it in range(5):
            gs = [rng.standard_normal(n).astype(np.float32) for _ in range(2)]
            for stub in stubs:
                for wid in (0, 1):
                    r = stub.ReceiveGradients(M("parameter_server.GradientUpdate")(
                        worker_id=wid, iteration=it, gradients=[M("parameter_server.Tensor")(
                            name="w", shape=[n], data=gs[wid].tolist())]), timeout=20)
                assert r.aggregation_complete
        got = [np.array(s.ServeParameters(M("parameter_server.PullRequest")(), timeout=20)
                        .parameters[0].data, dtype=np.float32) for s in stubs]
        assert np.allclose(got[0], got[1], atol=2e-5, rtol=1e-4)
        assert "device shard disabled" not in procs[1].output()
    finally:
        for p in procs:
            p.stop()


def test_worker_fills_gradients_on_gpu(tmp_path):
    ensure_binaries()
    ps_port, co_port = free_port(), free_port()
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "1", "0"], str(tmp_path),
              str(tmp_path / "ps.log"))
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"))
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        w = subprocess.run([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", "0", "5", "--device=0"],
                           capture_output=True, text=True, timeout=120)
        assert w.stdout.count("done=true") == 5, w.stdout + w.stderr
        assert "GPU fill unavailable" not in w.stderr
        stub = protos.Stub(grpc.insecure_channel(f"127.0.0.1:{ps_port}"),
                           "parameter_server.ParameterServer")
        p = stub.ServeParameters(M("parameter_server.PullRequest")(), timeout=10)
        assert np.allclose(np.array(p.parameters[0].data), -0.03, atol=1e-6)  # golden trace
    finally:
        ps.stop()
        co.stop()


@pytest.mark.parametrize("kind", ["adam", "momentum"])
def test_device_shard_checkpoint_carries_the_moments(tmp_path, kind):
    """Save -> restart -> load -> continue on a --device server equals an uninterrupted run:
    the optimizer moments live only in HBM and must travel through the checkpoint sidecar
    (without them a resumed Adam runs late-step bias correction on zero moments)."""
    ensure_binaries()
    flags = [f"--optimizer={kind}", "--lr=0.05", "--device=0"]
    rng = np.random.default_rng(1)
    n, steps, cut = 3000, 8, 4
    grads = [rng.standard_normal(n).astype(np.float32) for _ in range(steps)]

    def serve(log):
        port = free_port()
        p = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{port}", "1", "0", *flags], str(tmp_path),
                 str(tmp_path / log))
        assert wait_port(port)
        return p, protos.Stub(grpc.insecure_channel(f"127.0.0.1:{port}"),
                              "parameter_server.ParameterServer")

    def push(stub, it):
        r = stub.ReceiveGradients(M("parameter_server.GradientUpdate")(
            worker_id=0, iteration=it, gradients=[M("parameter_server.Tensor")(
                name="w", shape=[n], data=grads[it].tolist())]), timeout=20)
        assert r.aggregation_complete

    def pull(stub):
        return np.array(stub.ServeParameters(M("parameter_server.PullRequest")(), timeout=20)
                        .parameters[0].data, dtype=np.float32)

    whole, stub = serve("whole.log")
    try:
        for it in range(steps):
            push(stub, it)
        expected = pull(stub)
        assert "device shard disabled" not in whole.output()
    finally:
        whole.stop()
    first, stub = serve("first.log")
    try:
        for it in range(cut):
            push(stub, it)
        assert stub.SaveCheckpoint(M("parameter_server.SaveCheckpointRequest")(
            epoch=1, path="mid.ckpt"), timeout=20).success
    finally:
        first.stop()
    assert (tmp_path / "mid.ckpt.psb").stat().st_size > 2 * n  # the sidecar holds real moments
    second, stub = serve("second.log")
    try:
        assert stub.LoadCheckpoint(M("parameter_server.LoadCheckpointRequest")(path="mid.ckpt"),
                                   timeout=20).success
        for it in range(cut, steps):
            push(stub, it)
        got = pull(stub)
    finally:
        second.stop()
    assert np.allclose(got, expected, atol=2e-5, rtol=1e-4), float(np.abs(got - expected).max())
