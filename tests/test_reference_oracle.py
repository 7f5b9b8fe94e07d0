"""T1: behaviour of the new parameter-server core against the REFERENCE's own
ParameterServerCore / CoordinatorCore, compiled unmodified from /root/reference into
baseline/_ref (skipped when that install is absent).  Same random push sequences go to the
reference classes (through baseline/ref_harness.cpp) and to the C++ parameter_server binary
(through gRPC); parameters must agree (bit for bit with two workers, up to fp32 re-association beyond)."""
import ctypes as C
import os

import grpc
import numpy as np
import pytest

from parameter_server_distributed_b200.control import protos
from tests._native import BIN, ROOT, Proc, ensure_binaries, free_port, wait_port

LIB = os.path.join(ROOT, "baseline", "_ref", "libref_harness.so")
pytestmark = pytest.mark.skipif(not os.path.exists(LIB), reason="baseline/_ref not built")
ensure_binaries()
M = protos.msg


def _lib():
    lib = C.CDLL(LIB)
    lib.ref_ps_create.restype = C.c_void_p
    lib.ref_coord_create.restype = C.c_void_p
    return lib


def _arrays(names, shapes, bufs):
    n = len(names)
    flat = [d for s in shapes for d in s]
    return (n, (C.c_char_p * n)(*[s.encode() for s in names]),
            (C.c_int * n)(*[len(s) for s in shapes]), (C.c_int * len(flat))(*flat),
            (C.c_void_p * n)(*[b.ctypes.data for b in bufs]),
            (C.c_longlong * n)(*[b.size for b in bufs]))


def test_sync_aggregation_matches_reference(tmp_path):
    lib = _lib()
    W = 3
    ref = C.c_void_p(lib.ref_ps_create(W))
    port = free_port()
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{port}", str(W), "0"], str(tmp_path),
              str(tmp_path / "ps.log"))
    try:
        assert wait_port(port)
        stub = protos.Stub(grpc.insecure_channel(f"127.0.0.1:{port}"),
                           "parameter_server.ParameterServer")
        names, shapes = ["a", "b"], [[7, 5], [11]]
        rng = np.random.default_rng(0)
        # out-of-order iterations and a re-push from the same worker are part of the sequence
        order = [(0, 0), (0, 1), (1, 0), (0, 2), (1, 2), (1, 1), (1, 0), (2, 0), (2, 2), (2, 1)]
        for it, wid in order:
            bufs = [rng.standard_normal(35).astype(np.float32),
                    rng.standard_normal(11).astype(np.float32)]
            r_ref = lib.ref_ps_push(ref, wid, it, *_arrays(names, shapes, bufs))
            r = stub.ReceiveGradients(M("parameter_server.GradientUpdate")(
                worker_id=wid, iteration=it, gradients=[
                    M("parameter_server.Tensor")(name=n, shape=s, data=b.tolist())
                    for n, s, b in zip(names, shapes, bufs)]), timeout=5)
            assert int(r.aggregation_complete) == r_ref, (it, wid)
            got_recv = C.c_int(0)
            ready_ref = lib.ref_ps_sync_status(ref, it, C.byref(got_recv))
            s = stub.CheckSyncStatus(M("parameter_server.SyncStatusRequest")(iteration=it), timeout=5)
            assert (int(s.ready), s.workers_received) == (ready_ref, got_recv.value)
        out = [np.zeros(35, np.float32), np.zeros(11, np.float32)]
        n = lib.ref_ps_pull(ref, 0, 2, (C.c_void_p * 2)(*[o.ctypes.data for o in out]),
                            (C.c_longlong * 2)(35, 11))
        p = stub.ServeParameters(M("parameter_server.PullRequest")(), timeout=5)
        assert n == len(p.parameters) == 2
        for o, t in zip(out, p.parameters):
            # the reference sums in unordered_map iteration order (unspecified), we sum by
            # ascending worker id: identical up to fp32 re-association
            assert np.allclose(o, np.array(t.data, dtype=np.float32), rtol=1e-6, atol=1e-6)
    finally:
        ps.stop()
        lib.ref_ps_destroy(ref)


def test_reference_coordinator_returns_host_only_address_we_return_dialable():
    """Documents SURVEY D1: the reference hands workers a port-less address."""
    lib = _lib()
    c = C.c_void_p(lib.ref_coord_create(b"localhost", 50051))
    buf = C.create_string_buffer(64)
    assert lib.ref_coord_register(c, 0, buf, 64) == 1
    assert buf.value == b"localhost"          # the defect we fix (see test_interop)
    assert lib.ref_coord_register(c, 1, buf, 64) == 2
    assert lib.ref_coord_register(c, 1, buf, 64) == 2  # overwrite allowed
    lib.ref_coord_destroy(c)


def test_threaded_aggregation_is_bit_identical_on_large_tensors(tmp_path):
    """benchmarks/host_core_bench: both cores in one process on a 2.2 M-parameter cut of the
    MLP, two workers — large enough for the threaded, in-place mean + update of ps_core.cpp.
    The binary exits non-zero unless the parameters are bit-identical; the same holds for
    every thread count (each element is computed by the same expression in the same order)."""
    import json
    import subprocess
    ref_src = os.path.join(ROOT, "baseline", "_ref", "src", "parameter_server.cpp")
    if not os.path.exists(ref_src):
        pytest.skip("baseline/_ref sources absent")
    exe = str(tmp_path / "host_core_bench")
    subprocess.run(["g++", "-std=c++17", "-O2", "-pthread", "-I",
                    os.path.join(ROOT, "baseline", "_ref", "include"),
                    os.path.join(ROOT, "benchmarks", "host_core_bench.cpp"), ref_src,
                    os.path.join(ROOT, "csrc", "ps", "ps_core.cpp"),
                    os.path.join(ROOT, "csrc", "ps", "ps_device.cpp"), "-ldl", "-o", exe],
                   check=True, cwd=ROOT)
    for workers in (2, 5):
        r = subprocess.run([exe, str(workers), "4", "0.02"], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        row = json.loads(r.stdout)
        assert row["params"] > 2_000_000
        if workers == 2:
            assert row["bit_identical"] and row["max_abs_diff"] == 0
        else:  # the reference's summation order is that of an unordered_map
            assert row["max_abs_diff"] < 1e-7
