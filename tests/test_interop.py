"""T2: wire compatibility proven against a real gRPC C-core on the other end.

grpcio client -> C++ coordinator / parameter_server (all 9 reference RPCs + extensions,
including multi-MiB tensors that need HTTP/2 flow control), and the C++ worker_main ->
grpcio servers implementing the reference semantics."""
import os
import threading
import time
from concurrent import futures

import grpc
import numpy as np
import pytest

from parameter_server_distributed_b200.control import protos
from tests._native import BIN, Proc, ensure_binaries, free_port, wait_port

ensure_binaries()
M = protos.msg


@pytest.fixture()
def servers(tmp_path):
    ps_port, co_port = free_port(), free_port()
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "2", "0"], str(tmp_path),
              str(tmp_path / "ps.log"))
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"),
              env={"PSB_STALE_SCAN_MS": "200", "PSB_STALE_TIMEOUT_S": "1"})
    assert wait_port(ps_port) and wait_port(co_port)
    yield ps_port, co_port, ps, co, tmp_path
    ps.stop()
    co.stop()


def test_grpcio_client_against_cpp_coordinator(servers):
    ps_port, co_port, ps, co, _ = servers
    ch = grpc.insecure_channel(f"127.0.0.1:{co_port}")
    stub = protos.Stub(ch, "coordinator.Coordinator")
    r = stub.GetParameterServerAddress(M("coordinator.GetPSAddressRequest")(), timeout=5)
    assert (r.address, r.port) == ("127.0.0.1", ps_port)
    for wid in (0, 1):
        r = stub.RegisterWorker(M("coordinator.WorkerInfo")(
            worker_id=wid, address="localhost", port=0, hostname=f"worker-{wid}"), timeout=5)
        assert r.success and r.message == "worker registered"
        assert r.parameter_server_address == f"127.0.0.1:{ps_port}"  # dialable (fixes D1)
        assert r.total_workers == wid + 1
    r = stub.ListWorkers(M("coordinator.ListWorkersRequest")(), timeout=5)
    assert r.total_workers == 2 and sorted(w.worker_id for w in r.workers) == [0, 1]
    hb = stub.Heartbeat(M("coordinator.HeartbeatRequest")(worker_id=0, status=1), timeout=5)
    assert hb.success and abs(hb.timestamp - time.time()) < 5
    assert not stub.Heartbeat(M("coordinator.HeartbeatRequest")(worker_id=99), timeout=5).success
    # eviction: worker 1 stays silent, worker 0 keeps beating (timeout 1 s, scan 200 ms)
    end = time.time() + 2.5
    while time.time() < end:
        stub.Heartbeat(M("coordinator.HeartbeatRequest")(worker_id=0, status=0), timeout=5)
        time.sleep(0.2)
    r = stub.ListWorkers(M("coordinator.ListWorkersRequest")(), timeout=5)
    assert [w.worker_id for w in r.workers] == [0]
    assert "coordinator listening on" in co.output() and "parameter server: 127.0.0.1" in co.output()
    # unknown method -> UNIMPLEMENTED through trailers-only response
    bad = ch.unary_unary("/coordinator.Coordinator/Nope", request_serializer=lambda b: b,
                         response_deserializer=lambda b: b)
    with pytest.raises(grpc.RpcError) as e:
        bad(b"", timeout=5)
    assert e.value.code() == grpc.StatusCode.UNIMPLEMENTED


def test_grpcio_client_against_cpp_extension_service(servers):
    _, co_port, _, _, _ = servers
    ch = grpc.insecure_channel(f"127.0.0.1:{co_port}")
    ext = protos.Stub(ch, "psb.CoordinatorExt")
    for s in range(2):
        assert ext.RegisterShard(M("psb.ShardInfo")(shard_id=s, address=f"127.0.0.1:{6000 + s}",
                                                    device=s + 1), timeout=5).success
    sm = ext.GetShardMap(M("psb.ShardMapRequest")(), timeout=5)
    assert [(s.shard_id, s.address, s.device) for s in sm.shards] == [
        (0, "127.0.0.1:6000", 1), (1, "127.0.0.1:6001", 2)]
    blob = bytes(range(64))
    assert ext.PutHandle(M("psb.HandleBlob")(key="heap", rank=1, blob=blob), timeout=5).success
    hl = ext.GetHandles(M("psb.HandleQuery")(key="heap"), timeout=5)
    assert [(h.rank, h.blob) for h in hl.handles] == [(1, blob)]
    assert ext.UpdateClock(M("psb.ClockRequest")(worker_id=0, clock=5), timeout=5).min_clock == 5
    assert ext.UpdateClock(M("psb.ClockRequest")(worker_id=1, clock=3), timeout=5).min_clock == 3
    # barrier with 2 expected participants: the second arrival releases both
    out = {}

    def arrive(wid):
        c = grpc.insecure_channel(f"127.0.0.1:{co_port}")
        out[wid] = protos.Stub(c, "psb.CoordinatorExt").Barrier(M("psb.BarrierRequest")(
            name="b", worker_id=wid, generation=1, expected=2, timeout_ms=5000), timeout=10)

    t = threading.Thread(target=arrive, args=(0,))
    t.start()
    time.sleep(0.3)
    arrive(1)
    t.join()
    assert out[0].released and out[1].released and out[1].arrived == 2
    r = ext.Barrier(M("psb.BarrierRequest")(name="lonely", worker_id=0, generation=1, expected=2,
                                            timeout_ms=200), timeout=5)
    assert not r.released and r.arrived == 1


def test_grpcio_client_against_cpp_parameter_server(servers):
    ps_port, _, ps, _, tmp = servers
    ch = grpc.insecure_channel(f"127.0.0.1:{ps_port}",
                               options=[("grpc.max_receive_message_length", -1),
                                        ("grpc.max_send_message_length", -1)])
    stub = protos.Stub(ch, "parameter_server.ParameterServer")
    T = M("parameter_server.Tensor")
    r = stub.ServeParameters(M("parameter_server.PullRequest")(worker_id=0, iteration=0), timeout=5)
    assert len(r.parameters) == 0 and not r.ready
    grad = lambda v: [T(name="weight", shape=[10, 10], data=[v] * 100, dtype=0)]
    for it in range(5):  # golden trace, SURVEY App. C
        a = stub.ReceiveGradients(M("parameter_server.GradientUpdate")(
            worker_id=0, iteration=it, gradients=grad(0.01)), timeout=5)
        assert a.success and a.message == "gradients received" and not a.aggregation_complete
        assert (a.workers_received, a.total_workers, a.iteration) == (1, 2, it)
        s = stub.CheckSyncStatus(M("parameter_server.SyncStatusRequest")(iteration=it), timeout=5)
        assert not s.ready and s.workers_received == 1
        b = stub.ReceiveGradients(M("parameter_server.GradientUpdate")(
            worker_id=1, iteration=it, gradients=grad(0.01)), timeout=5)
        assert b.aggregation_complete and b.workers_received == 2
        # idempotent re-push after aggregation
        c = stub.ReceiveGradients(M("parameter_server.GradientUpdate")(
            worker_id=0, iteration=it, gradients=grad(123.0)), timeout=5)
        assert c.aggregation_complete
        p = stub.ServeParameters(M("parameter_server.PullRequest")(worker_id=0, iteration=it),
                                 timeout=5)
        assert p.ready and p.iteration == it
        exp = np.float32(0.01) - np.float32(0.01) * it
        assert np.allclose(np.array(p.parameters[0].data), exp, atol=1e-7)
    s = stub.CheckSyncStatus(M("parameter_server.SyncStatusRequest")(iteration=77), timeout=5)
    assert not s.ready and s.workers_received == 0
    # checkpoint: default name in the server's cwd, golden size 458 bytes (App. B)
    sv = stub.SaveCheckpoint(M("parameter_server.SaveCheckpointRequest")(epoch=0), timeout=5)
    assert sv.success and sv.message == "checkpoint saved"
    assert sv.checkpoint_path == "checkpoint_epoch_0.ckpt"
    raw = (tmp / "checkpoint_epoch_0.ckpt").read_bytes()
    assert len(raw) == 458
    header = bytes.fromhex("00000000" "04000000" "0100000000000000" "0600000000000000" +
                           b"weight".hex() + "0200000000000000" "0a000000" "0a000000"
                           "00000000" "6400000000000000")
    assert raw == header + bytes.fromhex("8fc2f5bc") * 100
    assert raw[-4:] == bytes.fromhex("8fc2f5bc")
    ld = stub.LoadCheckpoint(M("parameter_server.LoadCheckpointRequest")(
        path="checkpoint_epoch_0.ckpt"), timeout=5)
    assert ld.success and ld.message == "checkpoint loaded" and ld.epoch == 0
    assert len(ld.parameters) == 1 and len(ld.parameters[0].data) == 100
    ld = stub.LoadCheckpoint(M("parameter_server.LoadCheckpointRequest")(path="missing"), timeout=5)
    assert not ld.success and ld.message == "failed to load checkpoint"
    assert "parameter server listening on" in ps.output()


def test_large_tensors_cross_flow_control_both_ways(tmp_path):
    """24 MiB per message: far beyond the 64 KiB initial windows and the reference's 4 MiB cap."""
    port = free_port()
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{port}", "1", "0"], str(tmp_path),
              str(tmp_path / "ps.log"))
    try:
        assert wait_port(port)
        ch = grpc.insecure_channel(f"127.0.0.1:{port}",
                                   options=[("grpc.max_receive_message_length", -1),
                                            ("grpc.max_send_message_length", -1)])
        stub = protos.Stub(ch, "parameter_server.ParameterServer")
        n = 6 * 1024 * 1024
        data = np.arange(n, dtype=np.float32) * np.float32(1e-3)
        t = M("parameter_server.Tensor")(name="big", shape=[n], dtype=0)
        t.data.extend(data.tolist())
        a = stub.ReceiveGradients(M("parameter_server.GradientUpdate")(
            worker_id=0, iteration=0, gradients=[t]), timeout=60)
        assert a.aggregation_complete  # one worker: bootstraps the model with the gradient
        p = stub.ServeParameters(M("parameter_server.PullRequest")(worker_id=0, iteration=0),
                                 timeout=60)
        got = np.array(p.parameters[0].data, dtype=np.float32)
        assert got.shape == (n,) and np.array_equal(got, data)
    finally:
        ps.stop()


class _RefCoordinator:
    """grpcio server with the REFERENCE's semantics, including the host-only address (D1)."""

    def __init__(self, ps_port):
        self.ps_port = ps_port
        self.registered = []
        self.beats = 0

    def RegisterWorker(self, req, ctx):
        self.registered.append((req.worker_id, req.address, req.hostname))
        return M("coordinator.RegisterResponse")(success=True, message="worker registered",
                                                 parameter_server_address="127.0.0.1",
                                                 total_workers=len(self.registered))

    def Heartbeat(self, req, ctx):
        self.beats += 1
        return M("coordinator.HeartbeatResponse")(success=True, timestamp=int(time.time()))

    def ListWorkers(self, req, ctx):
        return M("coordinator.ListWorkersResponse")()

    def GetParameterServerAddress(self, req, ctx):
        return M("coordinator.GetPSAddressResponse")(address="127.0.0.1", port=self.ps_port)


class _RefPS:
    def __init__(self):
        self.pushes = []
        self.params = None

    def ReceiveGradients(self, req, ctx):
        self.pushes.append((req.worker_id, req.iteration, list(req.gradients[0].data[:3]),
                            req.gradients[0].name, list(req.gradients[0].shape)))
        if self.params is None:
            self.params = list(req.gradients)
        return M("parameter_server.PushResponse")(
            success=True, message="gradients received", iteration=req.iteration,
            aggregation_complete=True, workers_received=1, total_workers=1)

    def ServeParameters(self, req, ctx):
        return M("parameter_server.ParameterUpdate")(iteration=req.iteration,
                                                     parameters=self.params or [], ready=True)

    def CheckSyncStatus(self, req, ctx):
        return M("parameter_server.SyncStatusResponse")(iteration=req.iteration, ready=True)

    def SaveCheckpoint(self, req, ctx):
        return M("parameter_server.SaveCheckpointResponse")(success=True)

    def LoadCheckpoint(self, req, ctx):
        return M("parameter_server.LoadCheckpointResponse")(success=True, epoch=7,
                                                            message="checkpoint loaded")


def test_cpp_worker_against_grpcio_servers(tmp_path):
    ps_impl = _RefPS()
    ps_srv = grpc.server(futures.ThreadPoolExecutor(max_workers=4))
    protos.add_service(ps_srv, "parameter_server.ParameterServer", ps_impl)
    ps_port = ps_srv.add_insecure_port("127.0.0.1:0")
    ps_srv.start()
    co_impl = _RefCoordinator(ps_port)
    co_srv = grpc.server(futures.ThreadPoolExecutor(max_workers=4))
    protos.add_service(co_srv, "coordinator.Coordinator", co_impl)
    co_port = co_srv.add_insecure_port("127.0.0.1:0")
    co_srv.start()
    try:
        w = Proc([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", "3", "4", "", "0", "some.ckpt",
                  "--heartbeat-ms=100"], str(tmp_path), str(tmp_path / "w.log"))
        assert w.p.wait(timeout=60) == 0
        out = w.output()
        assert "worker 3 loaded checkpoint from some.ckpt (epoch 7)" in out
        for it in range(4):
            assert f"worker 3 iter {it} done=true" in out
        assert co_impl.registered == [(3, "localhost", "worker-3")]
        assert [p[:2] for p in ps_impl.pushes] == [(3, i) for i in range(4)]
        # empty server -> the fabricated weight[10,10] with 0.01 gradients
        assert ps_impl.pushes[0][3] == "weight" and ps_impl.pushes[0][4] == [10, 10]
        assert np.allclose(ps_impl.pushes[0][2], 0.01)
    finally:
        ps_srv.stop(0)
        co_srv.stop(0)


def test_barrier_membership_follows_eviction(servers):
    """T7: a worker that dies (stops heart-beating) is evicted and no longer blocks a barrier of
    "all registered workers" — the reference's sync barrier would stall forever."""
    _, co_port, _, _, _ = servers
    ch = grpc.insecure_channel(f"127.0.0.1:{co_port}")
    co = protos.Stub(ch, "coordinator.Coordinator")
    ext = protos.Stub(ch, "psb.CoordinatorExt")
    for wid in (0, 1):
        co.RegisterWorker(M("coordinator.WorkerInfo")(worker_id=wid, address="localhost"), timeout=5)
    stop = threading.Event()

    def beat():  # only worker 0 stays alive (eviction timeout 1 s, scan 200 ms)
        c = protos.Stub(grpc.insecure_channel(f"127.0.0.1:{co_port}"), "coordinator.Coordinator")
        while not stop.is_set():
            c.Heartbeat(M("coordinator.HeartbeatRequest")(worker_id=0, status=1), timeout=5)
            time.sleep(0.2)

    t = threading.Thread(target=beat)
    t.start()
    try:
        t0 = time.time()
        r = ext.Barrier(M("psb.BarrierRequest")(name="epoch", worker_id=0, generation=1, expected=0,
                                                timeout_ms=8000), timeout=15)
        assert r.released and r.arrived == 1
        assert 0.5 < time.time() - t0 < 6.0  # opened when worker 1 was evicted, not by timeout
    finally:
        stop.set()
        t.join()


def test_status_query_is_held_until_the_barrier_opens(tmp_path):
    """--long-poll-ms: a CheckSyncStatus that would say "not ready" waits for the completing
    push instead of making the worker poll again (reference: answers at once,
    /root/reference/src/parameter_server_service.cpp:85-96, worker sleeps 50 ms between polls,
    src/worker.cpp:372-395).  A negative answer still comes back after the hold time."""
    ps_port = free_port()
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "2", "0", "--long-poll-ms=1500"],
              str(tmp_path), str(tmp_path / "ps.log"))
    try:
        assert wait_port(ps_port)
        ch = grpc.insecure_channel(f"127.0.0.1:{ps_port}")
        stub = protos.Stub(ch, "parameter_server.ParameterServer")
        T = M("parameter_server.Tensor")
        grad = [T(name="weight", shape=[4], data=[0.5] * 4, dtype=0)]
        push = lambda w: stub.ReceiveGradients(M("parameter_server.GradientUpdate")(
            worker_id=w, iteration=0, gradients=grad), timeout=5)
        assert not push(0).aggregation_complete
        got = {}

        def waiter():
            t0 = time.time()
            got["resp"] = stub.CheckSyncStatus(
                M("parameter_server.SyncStatusRequest")(iteration=0), timeout=10)
            got["dt"] = time.time() - t0

        th = threading.Thread(target=waiter)
        th.start()
        time.sleep(0.3)
        assert th.is_alive(), "the negative answer was not held"
        assert push(1).aggregation_complete
        th.join(timeout=5)
        assert got["resp"].ready and got["resp"].workers_received == 2
        assert got["dt"] < 1.2, f"woken by the timeout, not by the completion ({got['dt']:.2f}s)"
        # nothing will complete iteration 5: the hold time bounds the wait
        t0 = time.time()
        s = stub.CheckSyncStatus(M("parameter_server.SyncStatusRequest")(iteration=5), timeout=10)
        assert not s.ready and 1.2 < time.time() - t0 < 4.0
    finally:
        ps.stop()


def test_standard_health_check_is_answered(servers):
    """grpc.health.v1.Health/Check (load balancers, grpc_health_probe, Kubernetes gRPC probes):
    both servers answer SERVING.  HealthCheckRequest{service = 1: string},
    HealthCheckResponse{status = 1: enum, SERVING = 1} — encoded by hand, grpcio's health package
    is not installed here."""
    ps_port, co_port, _, _, _ = servers
    for port in (ps_port, co_port):
        ch = grpc.insecure_channel(f"127.0.0.1:{port}")
        check = ch.unary_unary("/grpc.health.v1.Health/Check", request_serializer=lambda b: b,
                               response_deserializer=lambda b: b)
        assert check(b"", timeout=5) == b"\x08\x01"
        assert check(b"\x0a\x10parameter_server", timeout=5) == b"\x08\x01"
