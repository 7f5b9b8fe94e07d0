"""parameter_server_distributed_b200 — a B200-native parameter-server training framework.

Capabilities of araju6/parameter-server-distributed (coordinator, parameter-server shards,
workers, gRPC-compatible control plane, checkpoints) rebuilt Blackwell-first: shards live in
HBM, one per GPU; gradient push is a fused reduce-scatter + optimizer kernel over NVLink
peer memory, parameter pull an all-gather fused into the first tcgen05 GEMM.
"""
from . import _lib
from .ops.native import OptimConfig
from .parallel.sparse import ShardedEmbedding
from .trainer import Context, Trainer, init


def __getattr__(name):  # host-plane client: needs grpcio, so imported on first use
    if name in ("RemoteParameterServer", "RemoteTrainer", "RemoteEmbedding"):
        from .control import remote
        return getattr(remote, name)
    raise AttributeError(name)


__version__ = "0.1.0"
__all__ = ["OptimConfig", "Context", "Trainer", "init", "ShardedEmbedding",
           "RemoteParameterServer", "RemoteTrainer", "RemoteEmbedding", "_lib"]
