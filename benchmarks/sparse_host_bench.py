#!/usr/bin/env python
"""Host plane: what a sparse table buys over the reference's only way to train an embedding.

The reference's wire format is dense: a worker that trains a ``rows x dim`` embedding pushes a
gradient of the whole table and pulls the whole table back every iteration
(/root/reference/src/worker.cpp:240-312, proto/parameter_server.proto:19-24).  This script times
that path and the sparse one (``psb.ParameterServerExt``: PushSparse + PullRows of the rows a
batch touches) against the same ``build/bin/parameter_server``, one Python worker, localhost:

    python benchmarks/sparse_host_bench.py [--rows 1000000] [--dim 64] [--batch-rows 8192] [--out FILE]

One JSON line per path: median wall time of an iteration (push, wait for the update, pull) and
the bytes that crossed the socket.  No GPU involved.
"""
import argparse
import json
import os
import statistics
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np

import parameter_server_distributed_b200 as psd
from tests._native import BIN, Proc, ensure_binaries, free_port, wait_port


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--batch-rows", type=int, default=8192)
    ap.add_argument("--iters", type=int, default=7)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    ensure_binaries()
    tmp = tempfile.mkdtemp(prefix="psb_sparse_host_")
    ps_port, co_port = free_port(), free_port()
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "1", "0", "--optimizer=adam",
               "--lr=0.001"], tmp, os.path.join(tmp, "ps.log"))
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"], tmp,
              os.path.join(tmp, "co.log"))
    lines = []
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        rng = np.random.default_rng(0)
        with psd.RemoteParameterServer(f"127.0.0.1:{co_port}", 0, rpc_timeout_s=600) as w:
            # ---- sparse: the rows of a batch
            w.create_table("emb", args.rows, args.dim, init_std=0.02)
            ts = []
            for it in range(args.iters):
                ids = rng.integers(0, args.rows, args.batch_rows)
                g = rng.standard_normal((args.batch_rows, args.dim), dtype=np.float32)
                t0 = time.perf_counter()
                w.push_sparse(it, [("emb", ids, g)])
                w.pull_rows("emb", ids, after_iteration=it)
                ts.append(time.perf_counter() - t0)
            nbytes = args.batch_rows * (8 + 4 * args.dim) + args.batch_rows * (8 + 4 * args.dim)
            lines.append({"path": "sparse (PushSparse + PullRows)", "rows": args.rows, "dim": args.dim,
                          "rows_per_iteration": args.batch_rows,
                          "ms_per_iteration": round(statistics.median(ts) * 1e3, 2),
                          "wire_bytes_per_iteration": nbytes})
            # ---- dense: the whole table both ways (the first push bootstraps the model)
            table = rng.standard_normal((args.rows, args.dim), dtype=np.float32)
            w.push(0, [("emb", table)])
            w.wait(0)
            ts = []
            for it in range(1, 1 + max(3, args.iters // 2)):
                t0 = time.perf_counter()
                w.push(it, [("emb", table)])  # a dense gradient: zero except the touched rows
                w.wait(it)
                w.pull(it + 1)
                ts.append(time.perf_counter() - t0)
            lines.append({"path": "dense (ReceiveGradients + ServeParameters, the reference's only path)",
                          "rows": args.rows, "dim": args.dim,
                          "ms_per_iteration": round(statistics.median(ts) * 1e3, 2),
                          "wire_bytes_per_iteration": 2 * table.nbytes})
        lines.append({"speedup_sparse_over_dense":
                      round(lines[1]["ms_per_iteration"] / lines[0]["ms_per_iteration"], 1),
                      "wire_bytes_ratio": round(lines[1]["wire_bytes_per_iteration"] /
                                                lines[0]["wire_bytes_per_iteration"], 1)})
    finally:
        ps.stop()
        co.stop()
    text = "\n".join(json.dumps(x) for x in lines)
    print(text)
    if args.out:
        with open(args.out, "w") as f:
            f.write(text + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
