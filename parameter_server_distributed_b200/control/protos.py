"""Runtime protobuf classes + grpcio stubs for the three schemas in ``proto/``.

There is no ``grpc_tools`` on the box and protoc-3.13 generated ``_pb2.py`` files do not load
under protobuf 6, so the classes are built at import time from a serialized descriptor set
(``proto/descriptor_set.bin``, regenerated with ``torch/bin/protoc --descriptor_set_out`` when a
``.proto`` file is newer).  Python workers and the tests talk to the C++ binaries through
these; the binaries themselves use the hand-written codec in ``csrc/wire/pb.h``.
"""
from __future__ import annotations

import os
import subprocess
from pathlib import Path
from typing import Dict

ROOT = Path(__file__).resolve().parents[2]
PROTO_DIR = ROOT / "proto"
DESC = PROTO_DIR / "descriptor_set.bin"
FILES = ["coordinator.proto", "parameter_server.proto", "psb_ext.proto"]

SERVICES = {
    "coordinator.Coordinator": {
        "RegisterWorker": ("coordinator.WorkerInfo", "coordinator.RegisterResponse"),
        "Heartbeat": ("coordinator.HeartbeatRequest", "coordinator.HeartbeatResponse"),
        "ListWorkers": ("coordinator.ListWorkersRequest", "coordinator.ListWorkersResponse"),
        "GetParameterServerAddress": ("coordinator.GetPSAddressRequest",
                                      "coordinator.GetPSAddressResponse"),
    },
    "parameter_server.ParameterServer": {
        "ReceiveGradients": ("parameter_server.GradientUpdate", "parameter_server.PushResponse"),
        "ServeParameters": ("parameter_server.PullRequest", "parameter_server.ParameterUpdate"),
        "CheckSyncStatus": ("parameter_server.SyncStatusRequest",
                            "parameter_server.SyncStatusResponse"),
        "SaveCheckpoint": ("parameter_server.SaveCheckpointRequest",
                           "parameter_server.SaveCheckpointResponse"),
        "LoadCheckpoint": ("parameter_server.LoadCheckpointRequest",
                           "parameter_server.LoadCheckpointResponse"),
    },
    "psb.CoordinatorExt": {
        "RegisterShard": ("psb.ShardInfo", "psb.Ack"),
        "GetShardMap": ("psb.ShardMapRequest", "psb.ShardMapResponse"),
        "Barrier": ("psb.BarrierRequest", "psb.BarrierResponse"),
        "UpdateClock": ("psb.ClockRequest", "psb.ClockResponse"),
        "PutHandle": ("psb.HandleBlob", "psb.Ack"),
        "GetHandles": ("psb.HandleQuery", "psb.HandleList"),
    },
    "psb.ParameterServerExt": {
        "CreateTable": ("psb.TableSpec", "psb.Ack"),
        "PushSparse": ("psb.SparsePush", "psb.SparsePushResponse"),
        "PullRows": ("psb.RowsRequest", "psb.RowsResponse"),
    },
}


def _find_protoc():
    try:
        import torch
        p = Path(torch.__file__).parent / "bin" / "protoc"
        if p.exists():
            return str(p)
    except Exception:
        pass
    import shutil
    return shutil.which("protoc")


def _ensure_descriptor_set() -> bytes:
    newest = max((PROTO_DIR / f).stat().st_mtime for f in FILES)
    if not DESC.exists() or DESC.stat().st_mtime < newest:
        protoc = _find_protoc()
        if protoc is None:
            if DESC.exists():
                return DESC.read_bytes()
            raise RuntimeError("no protoc available to build proto/descriptor_set.bin")
        subprocess.run([protoc, f"--descriptor_set_out={DESC}", f"-I{PROTO_DIR}", *FILES],
                       check=True, cwd=PROTO_DIR)
    return DESC.read_bytes()


_CLASSES: Dict[str, type] = {}


def _load() -> None:
    if _CLASSES:
        return
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    fds = descriptor_pb2.FileDescriptorSet.FromString(_ensure_descriptor_set())
    pool = descriptor_pool.DescriptorPool()
    for fd in fds.file:
        pool.Add(fd)
    for fd in fds.file:
        for m in fd.message_type:
            full = f"{fd.package}.{m.name}"
            _CLASSES[full] = message_factory.GetMessageClass(pool.FindMessageTypeByName(full))


def msg(full_name: str):
    """Message class by full name, e.g. ``msg('parameter_server.Tensor')``."""
    _load()
    return _CLASSES[full_name]


class Stub:
    """grpcio client stub for one of the services above: ``Stub(channel, service).Method(req)``."""

    def __init__(self, channel, service: str):
        _load()
        for method, (req, resp) in SERVICES[service].items():
            setattr(self, method, channel.unary_unary(
                f"/{service}/{method}",
                request_serializer=lambda m: m.SerializeToString(),
                response_deserializer=_CLASSES[resp].FromString))


def add_service(server, service: str, impl) -> None:
    """Register ``impl`` (an object with one method per RPC) on a grpcio server."""
    import grpc
    _load()
    handlers = {}
    for method, (req, resp) in SERVICES[service].items():
        fn = getattr(impl, method)
        handlers[method] = grpc.unary_unary_rpc_method_handler(
            fn, request_deserializer=_CLASSES[req].FromString,
            response_serializer=lambda m: m.SerializeToString())
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(service, handlers),))
