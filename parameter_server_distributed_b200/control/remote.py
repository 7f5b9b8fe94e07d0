"""Worker-side client of the host plane: train a torch model through the gRPC parameter server.

The reference's worker is a C++ process whose "gradients" are a constant
(/root/reference/src/worker.cpp:316-329) — its protocol, though, is a complete synchronous
data-parallel step: register with the coordinator (:159-186), discover the parameter server
(:141-157), then per iteration pull (:240-252) -> compute -> push (:254-272) -> poll the sync
status (:274-287, :366-395).  This module speaks that protocol from Python (grpcio, message
classes from ``protos``; no codegen) against ``build/bin/coordinator`` and
``build/bin/parameter_server`` — or against the reference's own binaries, the wire is the
same — with real gradients of a ``torch.nn.Module``.

Two layers:

* :class:`RemoteParameterServer` — registration, heartbeats, ``pull`` / ``push`` / ``wait`` /
  checkpoints as numpy dictionaries;
* :class:`RemoteTrainer` — ``step(*inputs)`` on a model: pull, load, forward, backward, push,
  wait.  ``mode`` is the server's mode: ``sync`` waits for the barrier, ``async`` never waits,
  ``ssp`` waits at the pull until the slowest worker is within the server's bound.

This is the slow path by construction (parameters cross TCP as ``repeated float`` every step);
the HBM data plane is ``psd.Trainer``.  It exists so that a deployment built around a central
parameter-server process keeps working, with any worker that can produce gradients.
"""
from __future__ import annotations

import socket
import threading
import time
from typing import Callable, Dict, Iterable, List, Optional, Tuple

import numpy as np

from . import protos

IDLE, TRAINING, CHECKPOINTING, ERROR = 0, 1, 2, 3


# ---------------------------------------------------------------------- tensor messages in numpy
# The protobuf runtime moves `repeated float` through Python one element at a time: a 109 M-
# parameter model takes 17 s to pull and 33 s to push that way.  The two tensor-carrying messages
# are therefore written and read here, with the float payload going through numpy as one block
# (same bytes: proto3 packed floats are little-endian IEEE, proto/parameter_server.proto:13-24;
# tests/test_remote_trainer.py checks both directions against the protobuf runtime).
def _varint(v: int) -> bytes:
    v &= (1 << 64) - 1  # negative int32 are sign-extended to 10 bytes, as protobuf does
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _read_varint(buf, pos: int) -> Tuple[int, int]:
    v = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            return v, pos
        shift += 7
        if shift > 63:
            raise ValueError("varint too long")


def _int32(v: int) -> int:
    v &= 0xFFFFFFFF
    return v - (1 << 32) if v & 0x80000000 else v


def encode_gradient_update(worker_id: int, iteration: int,
                           grads: Iterable[Tuple[str, np.ndarray]]) -> bytes:
    """parameter_server.GradientUpdate{worker_id=1, iteration=2, repeated Tensor gradients=3}."""
    parts: List[bytes] = []
    if worker_id:
        parts += [b"\x08", _varint(worker_id)]
    if iteration:
        parts += [b"\x10", _varint(iteration)]
    for name, g in grads:
        g = np.require(g, dtype="<f4", requirements="C")  # keeps 0-d arrays 0-d
        t: List[bytes] = []
        nm = name.encode()
        if nm:
            t += [b"\x0a", _varint(len(nm)), nm]
        if g.ndim:
            shape = b"".join(_varint(d) for d in g.shape)
            t += [b"\x12", _varint(len(shape)), shape]
        if g.size:
            t += [b"\x1a", _varint(g.size * 4), g.reshape(-1).data.cast("B")]
        parts += [b"\x1a", _varint(sum(len(x) for x in t))] + t
    return b"".join(parts)


def decode_parameter_update(buf: bytes) -> Tuple[int, bool, List[Tuple[str, Tuple[int, ...], np.ndarray]]]:
    """parameter_server.ParameterUpdate -> (iteration, ready, [(name, shape, fp32 array)]).  The
    arrays are read-only views into ``buf`` (no copy); unknown fields are skipped."""
    mv = memoryview(buf)
    iteration, ready, tensors = 0, False, []

    def skip(wt: int, pos: int) -> int:
        if wt == 0:
            return _read_varint(mv, pos)[1]
        if wt == 1:
            return pos + 8
        if wt == 2:
            n, pos = _read_varint(mv, pos)
            return pos + n
        if wt == 5:
            return pos + 4
        raise ValueError(f"unsupported wire type {wt}")

    pos, end = 0, len(mv)
    while pos < end:
        tag, pos = _read_varint(mv, pos)
        field, wt = tag >> 3, tag & 7
        if field == 1 and wt == 0:
            v, pos = _read_varint(mv, pos)
            iteration = _int32(v)
        elif field == 3 and wt == 0:
            v, pos = _read_varint(mv, pos)
            ready = bool(v)
        elif field == 2 and wt == 2:
            n, pos = _read_varint(mv, pos)
            tend = pos + n
            name, shape, chunks = "", [], []
            while pos < tend:
                ttag, pos = _read_varint(mv, pos)
                tf, twt = ttag >> 3, ttag & 7
                if tf == 1 and twt == 2:
                    k, pos = _read_varint(mv, pos)
                    name = bytes(mv[pos:pos + k]).decode()
                    pos += k
                elif tf == 2 and twt == 2:
                    k, pos = _read_varint(mv, pos)
                    send = pos + k
                    while pos < send:
                        d, pos = _read_varint(mv, pos)
                        shape.append(_int32(d))
                elif tf == 2 and twt == 0:
                    d, pos = _read_varint(mv, pos)
                    shape.append(_int32(d))
                elif tf == 3 and twt == 2:
                    k, pos = _read_varint(mv, pos)
                    if k % 4:
                        raise ValueError("packed float field of odd length")
                    chunks.append(np.frombuffer(mv, dtype="<f4", count=k // 4, offset=pos))
                    pos += k
                elif tf == 3 and twt == 5:
                    chunks.append(np.frombuffer(mv, dtype="<f4", count=1, offset=pos))
                    pos += 4
                else:
                    pos = skip(twt, pos)
            if pos != tend:
                raise ValueError("tensor message overruns its length")
            data = chunks[0] if len(chunks) == 1 else (
                np.concatenate(chunks) if chunks else np.zeros(0, dtype="<f4"))
            tensors.append((name, tuple(shape), data))
        else:
            pos = skip(wt, pos)
    if pos != end:
        raise ValueError("message overruns its length")
    return iteration, ready, tensors


def encode_sparse_push(worker_id: int, iteration: int,
                       tables: Iterable[Tuple[str, int, np.ndarray, np.ndarray]]) -> bytes:
    """psb.SparsePush from (name, dim, int64 ids, fp32 values[len(ids), dim]) entries; ids and
    values go out as packed sfixed64 / float blocks straight from numpy."""
    chunks: List[bytes] = []  # joined once: no repeated copies of the value blocks
    if worker_id:
        chunks.append(b"\x08" + _varint(worker_id))
    if iteration:
        chunks.append(b"\x10" + _varint(iteration))
    for name, dim, ids, vals in tables:
        ids = np.ascontiguousarray(ids, dtype="<i8").ravel()
        vals = np.ascontiguousarray(vals, dtype="<f4").ravel()
        if vals.size != ids.size * dim:
            raise ValueError(f"table {name!r}: {vals.size} values for {ids.size} rows of {dim}")
        nb = name.encode()
        head = b"\x0a" + _varint(len(nb)) + nb
        if dim:
            head += b"\x10" + _varint(dim)
        body = [head]
        if ids.size:
            body += [b"\x1a" + _varint(ids.nbytes), ids.data.cast("B"),
                     b"\x22" + _varint(vals.nbytes), vals.data.cast("B")]
        chunks.append(b"\x1a" + _varint(sum(len(c) for c in body)))
        chunks += body
    return b"".join(chunks)


def encode_rows_request(worker_id: int, iteration: int, name: str, ids: np.ndarray) -> bytes:
    """psb.RowsRequest with the ids as one packed sfixed64 block."""
    ids = np.ascontiguousarray(ids, dtype="<i8").ravel()
    nb = name.encode()
    parts = []
    if worker_id:
        parts.append(b"\x08" + _varint(worker_id))
    if iteration:
        parts.append(b"\x10" + _varint(iteration))
    if nb:
        parts.append(b"\x1a" + _varint(len(nb)) + nb)
    if ids.size:
        parts += [b"\x22" + _varint(ids.nbytes), ids.data.cast("B")]
    return b"".join(parts)


def decode_rows_response(buf: bytes) -> Tuple[bool, int, np.ndarray, str]:
    """psb.RowsResponse -> (ready, dim, fp32 values (flat, read-only view), message)."""
    view = memoryview(buf)
    pos, ready, dim, msg = 0, False, 0, ""
    vals = np.empty(0, dtype="<f4")
    singles: List[float] = []
    while pos < len(buf):
        tag, pos = _read_varint(view, pos)
        field, wt = tag >> 3, tag & 7
        if wt == 0:
            v, pos = _read_varint(view, pos)
            if field == 1:
                ready = bool(v)
            elif field == 2:
                dim = _int32(v)
        elif wt == 2:
            n, pos = _read_varint(view, pos)
            if field == 3:
                vals = np.frombuffer(view[pos:pos + n], dtype="<f4")
            elif field == 4:
                msg = bytes(view[pos:pos + n]).decode(errors="replace")
            pos += n
        elif wt == 5:
            if field == 3:  # an unpacked repeated float
                singles.append(float(np.frombuffer(view[pos:pos + 4], dtype="<f4")[0]))
            pos += 4
        elif wt == 1:
            pos += 8
        else:
            raise ValueError(f"unsupported wire type {wt}")
    if singles:
        vals = np.concatenate([vals, np.asarray(singles, dtype="<f4")])
    return ready, dim, vals, msg


class RemoteParameterServer:
    """One worker's connection to the coordinator and the parameter server."""

    def __init__(self, coordinator: str, worker_id: int, *, address: str = "",
                 port: int = 0, hostname: Optional[str] = None, heartbeat_s: float = 5.0,
                 rpc_timeout_s: float = 30.0, ps_address: Optional[str] = None, shards: int = 0):
        """``shards > 1``: the model is split over that many parameter servers, the ones started
        with ``--shard-id=0..N-1 --coordinator=...`` (tensor k lives on shard k % N, as in the C++
        worker's ``--shards=N``); their addresses come from the coordinator's shard map."""
        import grpc
        self.worker_id = int(worker_id)
        self.timeout = rpc_timeout_s
        unlimited = [("grpc.max_receive_message_length", -1), ("grpc.max_send_message_length", -1)]
        self._coord_ch = grpc.insecure_channel(coordinator)
        self.coord = protos.Stub(self._coord_ch, "coordinator.Coordinator")
        self._info = protos.msg("coordinator.WorkerInfo")(
            worker_id=self.worker_id, address=address or "127.0.0.1", port=port,
            hostname=hostname or socket.gethostname())
        reg = self._retry(lambda: self.coord.RegisterWorker(self._info, timeout=self.timeout))
        if not reg.success:
            raise RuntimeError(f"registration refused: {reg.message}")
        self.total_workers = reg.total_workers
        if ps_address is None:
            where = self.coord.GetParameterServerAddress(
                protos.msg("coordinator.GetPSAddressRequest")(), timeout=self.timeout)
            # the reference coordinator answers host-only addresses with port 0
            ps_address = f"{where.address}:{where.port or 50051}"
        self.ps_address = ps_address
        if shards > 1:
            addrs = self._discover_shards(shards)
        else:
            addrs = [ps_address]
        self.shard_addresses = addrs
        self._ps_chs = [grpc.insecure_channel(a, options=unlimited) for a in addrs]
        self._stubs = [protos.Stub(ch, "parameter_server.ParameterServer") for ch in self._ps_chs]
        raw = dict(request_serializer=lambda b: b, response_deserializer=lambda b: b)
        base = "/parameter_server.ParameterServer/"
        self._raw_pull = [ch.unary_unary(base + "ServeParameters", **raw) for ch in self._ps_chs]
        self._raw_push = [ch.unary_unary(base + "ReceiveGradients", **raw) for ch in self._ps_chs]
        ext = "/psb.ParameterServerExt/"
        self._raw_push_sparse = [ch.unary_unary(ext + "PushSparse", **raw) for ch in self._ps_chs]
        self._raw_pull_rows = [ch.unary_unary(ext + "PullRows", **raw) for ch in self._ps_chs]
        self._ext_stubs = [protos.Stub(ch, "psb.ParameterServerExt") for ch in self._ps_chs]
        self._tables: Dict[str, Tuple[int, int, int]] = {}  # name -> (rows, dim, rows per shard)
        self.ps = self._stubs[0]
        self._pool = None
        if len(addrs) > 1:
            from concurrent.futures import ThreadPoolExecutor
            self._pool = ThreadPoolExecutor(max_workers=len(addrs))
        self.status = IDLE
        self.evictions = 0
        self._stop = threading.Event()
        self._hb = threading.Thread(target=self._heartbeat_loop, args=(heartbeat_s,), daemon=True)
        self._hb.start()

    # ------------------------------------------------------------------ plumbing
    def _discover_shards(self, n: int, timeout_s: float = 30.0) -> List[str]:
        ext = protos.Stub(self._coord_ch, "psb.CoordinatorExt")
        end = time.monotonic() + timeout_s
        while True:
            sm = ext.GetShardMap(protos.msg("psb.ShardMapRequest")(), timeout=self.timeout)
            found = {s.shard_id: s.address for s in sm.shards
                     if 0 <= s.shard_id < n and not s.address.startswith("hbm://")}
            if len(found) == n:
                return [found[i] for i in range(n)]
            if time.monotonic() >= end:
                raise RuntimeError(f"{len(found)} of {n} parameter-server shards registered")
            time.sleep(0.1)

    def _each_shard(self, fn):
        """[fn(shard_index, stub) for every shard], in parallel when there are several."""
        if self._pool is None:
            return [fn(0, self._stubs[0])]
        return list(self._pool.map(lambda a: fn(*a), enumerate(self._stubs)))

    def _retry(self, fn: Callable, attempts: int = 5):
        """Exponential back-off like the reference's query_with_retry (src/worker.cpp:129-139)."""
        import grpc
        for a in range(attempts):
            try:
                return fn()
            except grpc.RpcError:
                if a + 1 == attempts:
                    raise
                time.sleep(0.1 * 2 ** a)

    def _heartbeat_loop(self, period: float) -> None:
        import grpc
        HB = protos.msg("coordinator.HeartbeatRequest")
        while not self._stop.wait(period):
            try:
                r = self.coord.Heartbeat(HB(worker_id=self.worker_id, status=self.status),
                                         timeout=self.timeout)
                if not r.success:  # evicted (a long stall): come back, as the C++ worker does
                    self.evictions += 1
                    self.coord.RegisterWorker(self._info, timeout=self.timeout)
            except grpc.RpcError:
                pass  # coordinator away: keep trying at the next beat

    def close(self) -> None:
        self._stop.set()
        self._hb.join(timeout=2)
        self._coord_ch.close()
        for ch in self._ps_chs:
            ch.close()
        if self._pool is not None:
            self._pool.shutdown(wait=False)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ------------------------------------------------------------------ protocol
    def pull(self, iteration: int) -> Tuple[Dict[str, np.ndarray], bool]:
        """ServeParameters: ({name: fp32 array shaped like the server's tensor}, ready).  The
        arrays are read-only views of the response buffers."""
        req = protos.msg("parameter_server.PullRequest")(
            worker_id=self.worker_id, iteration=iteration).SerializeToString()
        resps = self._each_shard(lambda i, stub: decode_parameter_update(
            self._raw_pull[i](req, timeout=self.timeout)))
        n = len(resps)
        total = sum(len(r[2]) for r in resps)
        out = {}
        for k in range(total):  # tensor k is shard (k % n)'s tensor k // n
            name, shape, a = resps[k % n][2][k // n]
            out[name] = a.reshape(shape) if shape else a
        return out, all(r[1] for r in resps)

    def push(self, iteration: int, grads: Iterable[Tuple[str, np.ndarray]]) -> Tuple[bool, int, int]:
        """ReceiveGradients: (aggregation_complete, workers_received, total_workers).  Order
        matters: the server matches tensors by position and checks name + shape."""
        n = len(self._stubs)
        parts: List[List[Tuple[str, np.ndarray]]] = [[] for _ in range(n)]
        for k, item in enumerate(grads):
            parts[k % n].append(item)
        reqs = [encode_gradient_update(self.worker_id, iteration, p) for p in parts]
        Resp = protos.msg("parameter_server.PushResponse")
        rs = self._each_shard(lambda i, stub: Resp.FromString(self._retry(
            lambda: self._raw_push[i](reqs[i], timeout=self.timeout), attempts=3)))
        for r in rs:
            if not r.success:
                raise RuntimeError(f"push refused: {r.message}")
        return (all(bool(r.aggregation_complete) for r in rs),
                min(r.workers_received for r in rs), rs[0].total_workers)

    def server_iteration(self) -> int:
        """Newest iteration any worker has pushed (shard 0); this framework's server answers a
        pull with a negative iteration that way, the reference's echoes the request (-1)."""
        req = protos.msg("parameter_server.PullRequest")(
            worker_id=self.worker_id, iteration=-1).SerializeToString()
        return decode_parameter_update(self._raw_pull[0](req, timeout=self.timeout))[0]

    def wait(self, iteration: int, timeout_s: float = 10.0, poll_s: float = 0.0) -> bool:
        """CheckSyncStatus until the iteration is aggregated.  This framework's server holds a
        negative answer (``--long-poll-ms``), so polling back to back costs nothing; against the
        reference's server pass ``poll_s=0.05`` (its worker's own interval)."""
        Req = protos.msg("parameter_server.SyncStatusRequest")
        end = time.monotonic() + timeout_s
        while True:
            if all(stub.CheckSyncStatus(Req(iteration=iteration), timeout=self.timeout).ready
                   for stub in self._stubs):
                return True
            if time.monotonic() >= end:
                return False
            if poll_s:
                time.sleep(poll_s)

    def save_checkpoint(self, epoch: int, path: str = "") -> str:
        req = protos.msg("parameter_server.SaveCheckpointRequest")(epoch=epoch, path=path)
        rs = self._each_shard(lambda i, stub: stub.SaveCheckpoint(req, timeout=max(self.timeout, 120)))
        for r in rs:  # every shard writes `path` relative to its own working directory
            if not r.success:
                raise RuntimeError(r.message)
        return rs[0].checkpoint_path

    def load_checkpoint(self, path: str) -> int:
        req = protos.msg("parameter_server.LoadCheckpointRequest")(path=path)
        rs = self._each_shard(lambda i, stub: stub.LoadCheckpoint(req, timeout=max(self.timeout, 120)))
        for r in rs:
            if not r.success:
                raise RuntimeError(r.message)
        return rs[0].epoch

    # ------------------------------------------------------------------ sparse embedding tables
    # No counterpart in the reference (dense Tensor only, proto/parameter_server.proto:19-24).
    # A table is row-sharded over the parameter servers of the job: shard k holds rows
    # [k * R, (k + 1) * R), R = ceil(rows / shards) — the layout of the GPU plane's
    # ``ShardedEmbedding``.  With one server that is the whole table.
    def create_table(self, name: str, rows: int, dim: int, init_std: float = 0.0,
                     seed: int = 0) -> None:
        """Idempotent: every worker may call it; the first call creates the table."""
        n = len(self._stubs)
        per = (rows + n - 1) // n
        Spec = protos.msg("psb.TableSpec")

        def one(i, _stub):
            local = max(0, min(per, rows - i * per))
            if local == 0:
                return True
            return self._ext_stubs[i].CreateTable(
                Spec(name=name, rows=local, dim=dim, init_std=init_std, seed=seed + i),
                timeout=self.timeout).success
        if not all(self._each_shard(one)):
            raise RuntimeError(f"table {name!r} exists with another shape")
        self._tables[name] = (rows, dim, per)

    def _split_rows(self, name: str, ids: np.ndarray):
        rows, dim, per = self._tables[name]
        ids = np.asarray(ids, dtype=np.int64).ravel()
        valid = (ids >= 0) & (ids < rows)
        shard = np.where(valid, ids // per, -1)
        return rows, dim, per, ids, shard

    def push_sparse(self, iteration: int,
                    grads: Iterable[Tuple[str, np.ndarray, np.ndarray]]) -> bool:
        """PushSparse of (table, ids, values[len(ids), dim]) entries; every worker calls it once
        per iteration (sync mode counts pushers like the dense push), empty ids are fine.
        Returns whether the iteration is applied on every shard."""
        n = len(self._stubs)
        parts: List[list] = [[] for _ in range(n)]
        for name, ids, vals in grads:
            rows, dim, per, ids, shard = self._split_rows(name, ids)
            vals = np.asarray(vals, dtype=np.float32).reshape(-1, dim)
            for k in range(n):
                pick = shard == k
                parts[k].append((name, dim, ids[pick] - k * per, vals[pick]))
        reqs = [encode_sparse_push(self.worker_id, iteration, p) for p in parts]
        Resp = protos.msg("psb.SparsePushResponse")
        rs = self._each_shard(lambda i, stub: Resp.FromString(self._retry(
            lambda: self._raw_push_sparse[i](reqs[i], timeout=self.timeout), attempts=3)))
        for r in rs:
            if not r.success:
                raise RuntimeError(f"sparse push refused: {r.message}")
        return all(bool(r.aggregation_complete) for r in rs)

    def pull_rows(self, name: str, ids: np.ndarray, after_iteration: int = -1,
                  timeout_s: float = 30.0) -> np.ndarray:
        """Rows ``ids`` of ``name`` as fp32 ``[len(ids), dim]`` (zero rows for ids outside the
        table), once the sparse pushes of ``after_iteration`` are applied (-1: at once)."""
        rows, dim, per, ids, shard = self._split_rows(name, ids)
        n = len(self._stubs)
        out = np.zeros((ids.size, dim), dtype=np.float32)
        end = time.monotonic() + timeout_s

        def one(k, _stub):
            pick = np.nonzero(shard == k)[0]
            if pick.size == 0 and after_iteration < 0:
                return
            body = encode_rows_request(self.worker_id, after_iteration, name, ids[pick] - k * per)
            while True:
                ready, d, vals, msg = decode_rows_response(
                    self._raw_pull_rows[k](body, timeout=self.timeout))
                if msg:
                    raise RuntimeError(msg)
                if ready:
                    out[pick] = vals.reshape(-1, dim)
                    return
                if time.monotonic() >= end:
                    raise TimeoutError(f"sparse iteration {after_iteration} not applied on shard {k}")
        self._each_shard(one)
        return out

    def live_workers(self) -> List[int]:
        r = self.coord.ListWorkers(protos.msg("coordinator.ListWorkersRequest")(),
                                   timeout=self.timeout)
        return sorted(w.worker_id for w in r.workers)


class RemoteEmbedding:
    """An embedding table on the host-plane parameter server(s), with autograd.

    ``lookup(ids)`` pulls the rows (fp32) and is differentiable; ``apply_gradients(iteration)``
    pushes the row gradients the backward passes collected — once per iteration on every worker,
    like the dense push — and makes the next ``lookup`` wait for that iteration's update.  The
    host-plane sibling of ``parallel.sparse.ShardedEmbedding``; the reference has no sparse path
    (dense ``Tensor`` only, /root/reference/proto/parameter_server.proto:19-24)."""

    def __init__(self, ps: "RemoteParameterServer", name: str, vocab: int, dim: int,
                 init_std: float = 0.02, seed: int = 0):
        self.ps, self.name, self.vocab, self.dim = ps, name, int(vocab), int(dim)
        ps.create_table(name, vocab, dim, init_std=init_std, seed=seed)
        self.after = -1          # iteration whose sparse pushes the next pull must include
        self._pending: List[Tuple[np.ndarray, np.ndarray]] = []

    def lookup(self, ids):
        import torch
        emb = self

        class _Lookup(torch.autograd.Function):
            @staticmethod
            def forward(ctx, anchor):
                ctx.ids = ids.detach().reshape(-1).cpu().numpy()
                rows = emb.ps.pull_rows(emb.name, ctx.ids, after_iteration=emb.after)
                return torch.from_numpy(rows).view(*ids.shape, emb.dim)

            @staticmethod
            def backward(ctx, g):
                emb._pending.append((ctx.ids, g.reshape(-1, emb.dim).float().cpu().numpy()))
                return None

        return _Lookup.apply(torch.empty(0, requires_grad=True))

    def apply_gradients(self, iteration: int) -> bool:
        if self._pending:
            ids = np.concatenate([i for i, _ in self._pending])
            g = np.concatenate([g for _, g in self._pending])
        else:
            ids, g = np.empty(0, np.int64), np.empty((0, self.dim), np.float32)
        self._pending = []
        done = self.ps.push_sparse(iteration, [(self.name, ids, g)])
        self.after = iteration
        return done


class RemoteTrainer:
    """Data-parallel training of ``model`` with the parameter server as the only optimizer.

    The server owns the weights and applies ``--optimizer`` / ``--lr`` to the mean gradient of
    an iteration; this process holds a replica that it overwrites from every pull.  An empty
    server adopts the first aggregate as its model (reference semantics): every
    worker therefore pushes its initial weights as iteration 0 — with identically seeded
    workers their mean is that initialisation — and training starts at iteration 1.
    """

    def __init__(self, model, loss_fn, ps: RemoteParameterServer, *, mode: str = "sync",
                 sync_timeout_s: float = 10.0, start_iteration: Optional[int] = None,
                 tables: Iterable["RemoteEmbedding"] = ()):
        """``tables``: embedding tables the model looks rows up in (``RemoteEmbedding``); their
        row gradients are pushed after the dense ones, under the same iteration number."""
        import torch
        self.tables = list(tables)
        if mode not in ("sync", "async", "ssp"):
            raise ValueError(mode)
        self.torch = torch
        self.model, self.loss_fn, self.ps, self.mode = model, loss_fn, ps, mode
        self.sync_timeout_s = sync_timeout_s
        self.names = [n for n, p in model.named_parameters() if p.requires_grad]
        self.params = [p for _, p in model.named_parameters() if p.requires_grad]
        self.skipped = 0  # iterations whose barrier did not open in time (reference: moves on)
        have, _ = ps.pull(0)
        if start_iteration is not None:
            self.iteration = start_iteration
        elif have:
            # the server already has a model (other workers, or a checkpoint): join where the job
            # is — a pull with a negative iteration is answered with the newest pushed iteration
            self.iteration = max(1, ps.server_iteration())
        else:
            init = [(n, p.detach().float().cpu().numpy()) for n, p in zip(self.names, self.params)]
            done, _, _ = ps.push(0, init)
            if not done and mode == "sync":
                ps.wait(0, sync_timeout_s)
            self.iteration = 1

    def _load(self, pulled: Dict[str, np.ndarray]) -> None:
        import warnings
        torch = self.torch
        with torch.no_grad(), warnings.catch_warnings():
            warnings.simplefilter("ignore")  # the views are read-only; they are only read from
            for n, p in zip(self.names, self.params):
                a = pulled.get(n)
                if a is not None and a.size == p.numel():
                    p.copy_(torch.from_numpy(a).view_as(p))

    def step(self, *inputs):
        """One iteration; returns the loss tensor of this worker's batch."""
        ps, it = self.ps, self.iteration
        ps.status = TRAINING
        pulled, ready = ps.pull(it)
        if self.mode == "ssp":  # the server's staleness gate (held server-side, no sleep needed)
            end = time.monotonic() + self.sync_timeout_s
            while not ready and time.monotonic() < end:
                pulled, ready = ps.pull(it)
        self._load(pulled)
        for p in self.params:
            p.grad = None
        loss = self.loss_fn(self.model, *inputs)
        loss.backward()
        grads = [(n, (p.grad if p.grad is not None else self.torch.zeros_like(p))
                  .detach().float().cpu().numpy()) for n, p in zip(self.names, self.params)]
        done, _, _ = ps.push(it, grads)
        if not done and self.mode == "sync" and not ps.wait(it, self.sync_timeout_s):
            self.skipped += 1
        for t in self.tables:  # the next lookup waits (server side) for this iteration's rows
            t.apply_gradients(it)
        ps.status = IDLE
        self.iteration = it + 1
        return loss.detach()

    def sync_model(self) -> None:
        """Overwrite the local replica with the server's current weights."""
        self._load(self.ps.pull(self.iteration)[0])
