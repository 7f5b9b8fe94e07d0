"""T0: the hand-written proto3 codec (csrc/wire/pb.h) against the protobuf runtime built from
proto/*.proto — golden bytes of SURVEY.md App. A, every message of both reference services and
the extension service, unpacked repeated scalars, unknown fields, negative int32."""
import os
import struct
import subprocess

import pytest

from parameter_server_distributed_b200.control import protos
from tests._native import ensure_binaries

BIN = ensure_binaries()


def roundtrip(name: str, data: bytes) -> bytes:
    r = subprocess.run([f"{BIN}/psb_pbtool", name], input=data, capture_output=True, timeout=30)
    assert r.returncode == 0, (name, r.returncode)
    return r.stdout


def test_golden_bytes_gradient_update():
    T, G = protos.msg("parameter_server.Tensor"), protos.msg("parameter_server.GradientUpdate")
    g = G(worker_id=1, iteration=2,
          gradients=[T(name="weight", shape=[10, 10], data=[0.01, 0.01], dtype=0)])
    golden = bytes.fromhex("080110021a160a0677656967687412020a0a1a080ad7233c0ad7233c")
    assert g.SerializeToString() == golden
    assert roundtrip("parameter_server.GradientUpdate", golden) == golden


def test_golden_bytes_worker_info_and_push_response():
    W = protos.msg("coordinator.WorkerInfo")
    w = W(worker_id=0, address="localhost", port=0, hostname="worker-0")
    golden = bytes.fromhex("12096c6f63616c686f7374" "2208776f726b65722d30")
    assert w.SerializeToString() == golden
    assert roundtrip("coordinator.WorkerInfo", golden) == golden
    P = protos.msg("parameter_server.PushResponse")
    p = P(success=True, message="gradients received", iteration=3, aggregation_complete=True,
          workers_received=2, total_workers=2)
    golden = bytes.fromhex("08011212") + b"gradients received" + bytes.fromhex("180320012802" "3002")
    assert p.SerializeToString() == golden
    assert roundtrip("parameter_server.PushResponse", golden) == golden


def _samples():
    m = protos.msg
    T = m("parameter_server.Tensor")
    t1 = T(name="w", shape=[3, 2], data=[1.5, -2.25, 0.0, 3e-8, 1e30, -0.0], dtype=1)
    t2 = T(name="", shape=[], data=[], dtype=0)
    W = m("coordinator.WorkerInfo")
    return [
        ("coordinator.WorkerInfo", W(worker_id=-7, address="10.0.0.1", port=65535, hostname="h")),
        ("coordinator.RegisterResponse", m("coordinator.RegisterResponse")(
            success=True, message="worker registered", parameter_server_address="ps:50051",
            total_workers=3)),
        ("coordinator.HeartbeatRequest", m("coordinator.HeartbeatRequest")(worker_id=4, status=1)),
        ("coordinator.HeartbeatResponse", m("coordinator.HeartbeatResponse")(
            success=True, timestamp=1793994981)),
        ("coordinator.ListWorkersRequest", m("coordinator.ListWorkersRequest")()),
        ("coordinator.ListWorkersResponse", m("coordinator.ListWorkersResponse")(
            workers=[W(worker_id=0, address="a"), W(worker_id=1, port=9), W()], total_workers=3)),
        ("coordinator.GetPSAddressRequest", m("coordinator.GetPSAddressRequest")()),
        ("coordinator.GetPSAddressResponse", m("coordinator.GetPSAddressResponse")(
            address="localhost", port=50051)),
        ("parameter_server.Tensor", t1),
        ("parameter_server.GradientUpdate", m("parameter_server.GradientUpdate")(
            worker_id=2, iteration=2147483647, gradients=[t1, t2])),
        ("parameter_server.PushResponse", m("parameter_server.PushResponse")(
            success=True, message="gradients received", iteration=-1)),
        ("parameter_server.PullRequest", m("parameter_server.PullRequest")(worker_id=1, iteration=9)),
        ("parameter_server.ParameterUpdate", m("parameter_server.ParameterUpdate")(
            iteration=5, parameters=[t1], ready=True)),
        ("parameter_server.SyncStatusRequest", m("parameter_server.SyncStatusRequest")(iteration=3)),
        ("parameter_server.SyncStatusResponse", m("parameter_server.SyncStatusResponse")(
            iteration=3, ready=True, workers_received=2, total_workers=2)),
        ("parameter_server.SaveCheckpointRequest", m("parameter_server.SaveCheckpointRequest")(
            epoch=2, path="/tmp/x.ckpt")),
        ("parameter_server.SaveCheckpointResponse", m("parameter_server.SaveCheckpointResponse")(
            success=True, message="checkpoint saved", checkpoint_path="/tmp/x.ckpt")),
        ("parameter_server.LoadCheckpointRequest", m("parameter_server.LoadCheckpointRequest")(
            path="c.ckpt")),
        ("parameter_server.LoadCheckpointResponse", m("parameter_server.LoadCheckpointResponse")(
            success=True, message="checkpoint loaded", epoch=1, parameters=[t1, t2])),
        ("psb.ShardInfo", m("psb.ShardInfo")(shard_id=3, address="127.0.0.1:6000", device=4)),
        ("psb.ShardMapResponse", m("psb.ShardMapResponse")(
            shards=[m("psb.ShardInfo")(shard_id=1, address="x")], total_workers=8)),
        ("psb.BarrierRequest", m("psb.BarrierRequest")(
            name="step", worker_id=1, generation=7, expected=8, timeout_ms=1 << 40)),
        ("psb.BarrierResponse", m("psb.BarrierResponse")(released=True, arrived=8)),
        ("psb.ClockRequest", m("psb.ClockRequest")(worker_id=1, clock=12)),
        ("psb.ClockResponse", m("psb.ClockResponse")(min_clock=10)),
        ("psb.HandleBlob", m("psb.HandleBlob")(key="heap", rank=2, blob=bytes(range(64)))),
        ("psb.HandleQuery", m("psb.HandleQuery")(key="heap")),
        ("psb.HandleList", m("psb.HandleList")(
            handles=[m("psb.HandleBlob")(key="k", rank=1, blob=b"\x00\xff")])),
        ("psb.Ack", m("psb.Ack")(success=True)),
        ("psb.TableSpec", m("psb.TableSpec")(name="emb", rows=1 << 33, dim=64, init_std=0.02, seed=-5)),
        ("psb.SparseRows", m("psb.SparseRows")(name="emb", dim=2, indices=[0, -1, 1 << 40],
                                               values=[1.5, -2.0, 0.0, 3e-8, 1e30, -0.0])),
        ("psb.SparsePush", m("psb.SparsePush")(worker_id=2, iteration=9, tables=[
            m("psb.SparseRows")(name="a", dim=1, indices=[7], values=[0.25]),
            m("psb.SparseRows")(name="empty", dim=8)])),
        ("psb.SparsePushResponse", m("psb.SparsePushResponse")(
            success=True, message="sparse gradients received", aggregation_complete=True,
            workers_received=2, total_workers=2)),
        ("psb.RowsRequest", m("psb.RowsRequest")(worker_id=1, iteration=-1, name="emb",
                                                 indices=[5, 5, 0])),
        ("psb.RowsResponse", m("psb.RowsResponse")(ready=True, dim=2, values=[1.0, 2.0, 3.0, 4.0],
                                                   message="m")),
    ]


@pytest.mark.parametrize("name,message", _samples(), ids=[n for n, _ in _samples()])
def test_roundtrip_equals_runtime(name, message):
    data = message.SerializeToString()
    out = roundtrip(name, data)
    assert out == data  # canonical field order + packed encoding, byte for byte
    assert protos.msg(name).FromString(out) == message


def varint(v):
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7f) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def test_decoder_accepts_unpacked_and_unknown_fields():
    # Tensor with shape as two unpacked varints (tag 0x10), data as unpacked fixed32 (tag 0x1d),
    # an unknown varint field 15, an unknown length-delimited field 14 and an unknown fixed64.
    raw = (b"\x0a\x01w" + b"\x10\x02" + b"\x10\x03" + b"\x1d" + struct.pack("<f", 1.5) +
           b"\x1d" + struct.pack("<f", -2.0) + b"\x78\x2a" + b"\x72\x03abc" +
           b"\x69" + b"\x00" * 8 + b"\x20\x01")
    out = roundtrip("parameter_server.Tensor", raw)
    t = protos.msg("parameter_server.Tensor").FromString(out)
    assert t.name == "w" and list(t.shape) == [2, 3] and list(t.data) == [1.5, -2.0] and t.dtype == 1


def test_negative_int32_is_ten_byte_varint():
    P = protos.msg("parameter_server.PullRequest")
    data = P(worker_id=-1, iteration=-2147483648).SerializeToString()
    assert len(data) == 2 * 11
    assert roundtrip("parameter_server.PullRequest", data) == data


def test_malformed_input_is_rejected():
    r = subprocess.run([f"{BIN}/psb_pbtool", "parameter_server.Tensor"], input=b"\x0a\x05ab",
                       capture_output=True)
    assert r.returncode == 3


def test_decoders_never_crash_on_mutated_messages():
    """Bit flips, truncations and random tails of valid encodings: the hand-written decoders
    either accept (exit 0) or reject (exit 3) — never a signal.  The servers feed them bytes
    from the network."""
    import random
    T = protos.msg("parameter_server.Tensor")
    seeds = {
        "parameter_server.GradientUpdate": protos.msg("parameter_server.GradientUpdate")(
            worker_id=3, iteration=7, gradients=[T(name="w", shape=[2, 3], data=[1, 2, 3, 4, 5, 6]),
                                                 T(name="b", shape=[3], data=[0.5, 1.5, 2.5], dtype=1)]
        ).SerializeToString(),
        "parameter_server.ParameterUpdate": protos.msg("parameter_server.ParameterUpdate")(
            iteration=9, ready=True, parameters=[T(name="p", shape=[4], data=[1, 2, 3, 4])]
        ).SerializeToString(),
        "coordinator.ListWorkersResponse": protos.msg("coordinator.ListWorkersResponse")(
            total_workers=2, workers=[protos.msg("coordinator.WorkerInfo")(
                worker_id=i, address="10.0.0.%d" % i, port=5000 + i, hostname="h%d" % i)
                for i in range(2)]).SerializeToString(),
        "psb.SparsePush": protos.msg("psb.SparsePush")(worker_id=1, iteration=3, tables=[
            protos.msg("psb.SparseRows")(name="emb", dim=2, indices=[4, 9], values=[1, 2, 3, 4])]
        ).SerializeToString(),
        "psb.RowsRequest": protos.msg("psb.RowsRequest")(
            worker_id=1, iteration=2, name="emb", indices=[1, 2, 3]).SerializeToString(),
    }
    rng = random.Random(7)
    codes = set()
    for name, good in seeds.items():
        for _ in range(60):
            b = bytearray(good)
            kind = rng.randrange(4)
            if kind == 0:
                for _ in range(rng.randrange(1, 4)):
                    b[rng.randrange(len(b))] ^= 1 << rng.randrange(8)
            elif kind == 1:
                del b[rng.randrange(1, len(b)):]
            elif kind == 2:
                b += bytes(rng.getrandbits(8) for _ in range(rng.randrange(1, 12)))
            else:
                i = rng.randrange(len(b))
                b[i:i] = bytes([0xFF] * rng.randrange(1, 11))  # runaway varints / huge lengths
            r = subprocess.run([f"{BIN}/psb_pbtool", name], input=bytes(b), capture_output=True)
            assert r.returncode in (0, 3), (name, bytes(b).hex(), r.returncode)
            codes.add(r.returncode)
    assert codes == {0, 3}  # the mutations do produce both outcomes


def test_public_headers_are_plain_c(tmp_path):
    """include/psb_kernels.h and include/psb_control.h are the C ABI: they must compile as C11
    (pedantic) and as C++11 without anything else on the include path."""
    src = tmp_path / "abi.c"
    src.write_text('#include "psb_kernels.h"\n#include "psb_control.h"\nint main(void) { return 0; }\n')
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    inc = os.path.join(root, "include")
    for cmd in (["gcc", "-std=c11", "-Wall", "-Wextra", "-Wpedantic", "-Werror"],
                ["g++", "-std=c++11", "-Wall", "-Wextra", "-Werror", "-x", "c++"]):
        r = subprocess.run(cmd + ["-I", inc, "-fsyntax-only", str(src)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
