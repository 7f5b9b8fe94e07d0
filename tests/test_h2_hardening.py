"""The HTTP/2 server under a peer that does not play by the rules (csrc/wire/h2.cpp,
grpc_lite.cpp): advertised and enforced stream limit, bounded header blocks, garbage frames.
Raw sockets on purpose — a well-behaved client library never sends any of this.

The reference gets these properties from the gRPC C++ library it links
(/root/reference/CMakeLists.txt:20); the hand-written transport has to have them itself."""
import random
import socket
import struct
import time

import grpc

from parameter_server_distributed_b200.control import protos
from tests._native import BIN, Proc, ensure_binaries, free_port, wait_port

ensure_binaries()
M = protos.msg
PREFACE = b"PRI * HTTP/2.0\r\n\r\nSM\r\n\r\n"
DATA, HEADERS, RST_STREAM, SETTINGS, GOAWAY, CONTINUATION = 0, 1, 3, 4, 7, 9


def frame(ftype, flags, sid, payload=b""):
    n = len(payload)
    return struct.pack(">BHBBI", n >> 16, n & 0xFFFF, ftype, flags, sid & 0x7FFFFFFF) + payload


def lit(name, value):
    """HPACK literal header field without indexing, new name, no Huffman (RFC 7541 §6.2.2)."""
    assert len(name) < 127 and len(value) < 127
    return b"\x00" + bytes([len(name)]) + name.encode() + bytes([len(value)]) + value.encode()


def request_headers(path):
    return b"".join([lit(":method", "POST"), lit(":scheme", "http"), lit(":path", path),
                     lit(":authority", "x"), lit("content-type", "application/grpc"),
                     lit("te", "trailers")])


def grpc_body(msg):
    b = msg.SerializeToString()
    return b"\x00" + struct.pack(">I", len(b)) + b


def connect(port):
    s = socket.create_connection(("127.0.0.1", port), timeout=5)
    s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
    s.sendall(PREFACE + frame(SETTINGS, 0, 0))
    return s


def read_frames(s, until, timeout=5.0):
    """Collect (type, flags, sid, payload) until `until(frames)` is true, EOF or timeout."""
    buf, out, end = b"", [], time.time() + timeout
    s.settimeout(0.2)
    while time.time() < end and not until(out):
        try:
            chunk = s.recv(65536)
        except socket.timeout:
            continue
        except OSError:
            break
        if not chunk:
            break
        buf += chunk
        while len(buf) >= 9:
            n = (buf[0] << 16) | (buf[1] << 8) | buf[2]
            if len(buf) < 9 + n:
                break
            out.append((buf[3], buf[4], struct.unpack(">I", buf[5:9])[0] & 0x7FFFFFFF, buf[9:9 + n]))
            buf = buf[9 + n:]
    return out


def coordinator(tmp_path):
    port = free_port()
    p = Proc([f"{BIN}/coordinator", f"127.0.0.1:{port}", "127.0.0.1:50051"], str(tmp_path),
             str(tmp_path / "co.log"))
    assert wait_port(port)
    return p, port


def still_serves(port):
    stub = protos.Stub(grpc.insecure_channel(f"127.0.0.1:{port}"), "coordinator.Coordinator")
    r = stub.RegisterWorker(M("coordinator.WorkerInfo")(worker_id=7, address="a", port=1,
                                                        hostname="h"), timeout=5)
    return r.success


def test_settings_advertise_the_stream_limit(tmp_path):
    co, port = coordinator(tmp_path)
    try:
        s = connect(port)
        fr = read_frames(s, lambda f: any(t == SETTINGS and not fl & 1 for t, fl, _, _ in f))
        payload = next(p for t, fl, _, p in fr if t == SETTINGS and not fl & 1)
        settings = {struct.unpack(">H", payload[i:i + 2])[0]: struct.unpack(">I", payload[i + 2:i + 6])[0]
                    for i in range(0, len(payload), 6)}
        assert settings[3] == 128            # MAX_CONCURRENT_STREAMS
        assert settings[4] == 0x7FFFFFFF     # INITIAL_WINDOW_SIZE: never the bottleneck
        s.close()
    finally:
        co.stop()


def test_streams_above_the_limit_are_refused(tmp_path):
    """130 Barrier calls that cannot complete (expected = 2, one participant) on one connection:
    128 are parked on handler threads, the rest get RST_STREAM(REFUSED_STREAM)."""
    co, port = coordinator(tmp_path)
    try:
        s = connect(port)
        hb = request_headers("/psb.CoordinatorExt/Barrier")
        out = b""
        for i in range(130):
            sid = 1 + 2 * i
            body = grpc_body(M("psb.BarrierRequest")(name=f"b{i}", worker_id=0, generation=0,
                                                     expected=2, timeout_ms=4000))
            out += frame(HEADERS, 0x4, sid, hb) + frame(DATA, 0x1, sid, body)
        s.sendall(out)
        fr = read_frames(s, lambda f: sum(t == RST_STREAM for t, _, _, _ in f) >= 2, timeout=3)
        rst = [(sid, struct.unpack(">I", p)[0]) for t, _, sid, p in fr if t == RST_STREAM]
        assert len(rst) == 2 and all(code == 7 for _, code in rst), rst
        assert sorted(sid for sid, _ in rst) == [257, 259]
        # the parked ones answer when their barrier times out: released = false, grpc-status 0
        fr = read_frames(s, lambda f: sum(t == HEADERS and fl & 1 for t, fl, _, _ in f) >= 128,
                         timeout=10)
        assert sum(t == HEADERS and fl & 1 for t, fl, _, _ in fr) == 128
        s.close()
        assert still_serves(port)
    finally:
        co.stop()


def test_oversized_header_block_ends_the_connection(tmp_path):
    co, port = coordinator(tmp_path)
    try:
        s = connect(port)
        chunk = lit("x-filler", "y" * 100) * 140  # ~15 KiB per frame
        s.sendall(frame(HEADERS, 0, 1, chunk))
        try:
            for _ in range(6):  # > 64 KiB in CONTINUATION frames, END_HEADERS never set
                s.sendall(frame(CONTINUATION, 0, 1, chunk))
        except OSError:
            pass  # the server may already have hung up
        fr = read_frames(s, lambda f: any(t == GOAWAY for t, _, _, _ in f), timeout=3)
        go = [p for t, _, _, p in fr if t == GOAWAY]
        assert go and struct.unpack(">I", go[0][4:8])[0] == 11  # ENHANCE_YOUR_CALM
        s.settimeout(2)
        try:
            assert s.recv(1) == b""  # closed
        except OSError:
            pass
        assert still_serves(port)
    finally:
        co.stop()


def test_garbage_does_not_take_the_server_down(tmp_path):
    co, port = coordinator(tmp_path)
    try:
        rng = random.Random(1234)
        for c in range(60):
            s = socket.create_connection(("127.0.0.1", port), timeout=5)
            try:
                kind = c % 3
                if kind == 0:    # no preface at all
                    s.sendall(bytes(rng.getrandbits(8) for _ in range(rng.randrange(1, 400))))
                else:            # valid preface, then random frames (kind 2: plausible frame types)
                    out = PREFACE + frame(SETTINGS, 0, 0)
                    for _ in range(rng.randrange(1, 12)):
                        t = rng.randrange(0, 10) if kind == 2 else rng.randrange(0, 256)
                        payload = bytes(rng.getrandbits(8) for _ in range(rng.randrange(0, 300)))
                        out += frame(t, rng.getrandbits(8), rng.choice([0, 1, 3, 5, 2, 99]), payload)
                    if rng.random() < 0.3:  # a frame header that promises more than ever arrives
                        out += struct.pack(">BHBBI", 0x7F, 0xFFFF, 0, 0, 1)
                    s.sendall(out)
                s.settimeout(0.05)
                try:
                    s.recv(4096)
                except OSError:
                    pass
            except OSError:
                pass
            finally:
                s.close()
        assert co.p.poll() is None, co.output()[-500:]
        assert still_serves(port)
    finally:
        co.stop()


def test_burst_of_interdependent_calls_after_the_pool_has_idle_threads(tmp_path):
    """40 Barrier calls of ONE barrier (expected = 40) in a single write, on a server whose
    handler pool already has idle threads: every call needs its own thread at the same time —
    a request queued behind another one would deadlock the barrier until its timeout."""
    co, port = coordinator(tmp_path)
    try:
        for _ in range(3):
            assert still_serves(port)  # leaves idle pooled threads behind
        s = connect(port)
        hb = request_headers("/psb.CoordinatorExt/Barrier")
        out = b""
        k = 40
        for i in range(k):
            sid = 1 + 2 * i
            body = grpc_body(M("psb.BarrierRequest")(name="all", worker_id=i, generation=0,
                                                     expected=k, timeout_ms=6000))
            out += frame(HEADERS, 0x4, sid, hb) + frame(DATA, 0x1, sid, body)
        t0 = time.time()
        s.sendall(out)
        fr = read_frames(s, lambda f: sum(t == HEADERS and fl & 1 for t, fl, _, _ in f) >= k,
                         timeout=12)
        dt = time.time() - t0
        bodies = [p for t, _, _, p in fr if t == DATA and len(p) >= 5]
        released = [M("psb.BarrierResponse").FromString(p[5:]).released for p in bodies]
        assert len(released) == k and all(released), (len(released), sum(released))
        assert dt < 4.0, f"the barrier opened only after {dt:.1f}s"
        s.close()
    finally:
        co.stop()
