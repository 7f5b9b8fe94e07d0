#!/usr/bin/env python
"""A worker written in Python against the C++ coordinator and parameter server, over the same
gRPC wire schema the reference uses (proto/coordinator.proto, proto/parameter_server.proto):
register, find the parameter server, then per iteration pull -> compute -> push -> poll the
sync status, as the reference's worker does (/root/reference/src/worker.cpp:331-406).

    build/bin/coordinator 0.0.0.0:50052 127.0.0.1:50051 &
    build/bin/parameter_server 0.0.0.0:50051 1 0 --optimizer=sgd --lr=0.1 &
    python examples/python_worker.py --coordinator 127.0.0.1:50052 --iterations 20

The "model" is a least-squares fit y = x·w: the parameter server owns w, this process only
computes gradients.  Any grpcio client generated from the reference's .proto files works the
same way; here the message classes come from `control/protos.py` (no protoc codegen needed).
"""
from __future__ import annotations

import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import grpc
import numpy as np

from parameter_server_distributed_b200.control.protos import Stub, msg


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--coordinator", default="127.0.0.1:50052")
    ap.add_argument("--worker-id", type=int, default=0)
    ap.add_argument("--iterations", type=int, default=20)
    ap.add_argument("--dim", type=int, default=16)
    args = ap.parse_args()

    coord = Stub(grpc.insecure_channel(args.coordinator), "coordinator.Coordinator")
    reg = coord.RegisterWorker(msg("coordinator.WorkerInfo")(
        worker_id=args.worker_id, address="127.0.0.1", port=0, hostname="python-worker"))
    where = coord.GetParameterServerAddress(msg("coordinator.GetPSAddressRequest")())
    ps_addr = f"{where.address}:{where.port}"
    print(f"registered ({reg.message}); parameter server at {ps_addr}", flush=True)
    ps = Stub(grpc.insecure_channel(ps_addr), "parameter_server.ParameterServer")

    Tensor = msg("parameter_server.Tensor")
    rng = np.random.default_rng(args.worker_id)
    target = np.linspace(-1.0, 1.0, args.dim, dtype=np.float32)
    w = np.zeros(args.dim, dtype=np.float32)
    loss = float("nan")
    for it in range(args.iterations):
        coord.Heartbeat(msg("coordinator.HeartbeatRequest")(worker_id=args.worker_id, status=1))
        pulled = ps.ServeParameters(msg("parameter_server.PullRequest")(
            worker_id=args.worker_id, iteration=it))
        for t in pulled.parameters:  # empty before the first push: the server adopts the first
            if t.name == "w":        # gradient's names and shapes (bootstrap, like the reference)
                w = np.asarray(t.data, dtype=np.float32)
        x = rng.standard_normal((64, args.dim)).astype(np.float32)
        err = x @ w - x @ target
        loss = float(np.mean(err ** 2))
        grad = (2.0 / len(x)) * (x.T @ err)
        push = ps.ReceiveGradients(msg("parameter_server.GradientUpdate")(
            worker_id=args.worker_id, iteration=it,
            gradients=[Tensor(name="w", shape=[args.dim], data=grad.tolist(), dtype=0)]))
        for _ in range(200):  # 50 ms x 200, the reference's poll budget
            st = ps.CheckSyncStatus(msg("parameter_server.SyncStatusRequest")(iteration=it))
            if st.ready:
                break
            time.sleep(0.05)
        print(f"worker {args.worker_id} iter {it} done={str(bool(st.ready)).lower()} "
              f"loss={loss:.5f} received={push.workers_received}/{push.total_workers}", flush=True)
    return 0 if np.isfinite(loss) else 1


if __name__ == "__main__":
    sys.exit(main())
