#!/usr/bin/env python
"""Short single-GPU program that launches each hand-written kernel a few times: the target of
`ncu --set full` captures (profiles/).  Run plainly first; never take timings from a profiled
run."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

from parameter_server_distributed_b200 import OptimConfig
from parameter_server_distributed_b200.ops import native
from parameter_server_distributed_b200.parallel.buckets import plan_buckets
from parameter_server_distributed_b200.parallel.device_backend import DeviceBackend
from parameter_server_distributed_b200.parallel.fabric import Fabric

DEV = torch.device("cuda:0")
torch.cuda.set_device(DEV)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3

# K2f: tcgen05 GEMM (first MLP layer shape of bench.py: M=4096 tokens, N=8192, K=4096)
x = (torch.randn(4096, 4096, device=DEV) * 0.1).to(torch.bfloat16)
w = (torch.randn(8192, 4096, device=DEV) * 0.1).to(torch.bfloat16)
y = torch.empty(4096, 8192, dtype=torch.bfloat16, device=DEV)
from parameter_server_distributed_b200 import _lib
for variant in (1, 0):  # cta_group::2 (default), then cta_group::1
    _lib.check(_lib.load().psb_gemm_select(variant))
    for _ in range(reps):
        native.gemm_bf16(x, w, out=y)
_lib.check(_lib.load().psb_gemm_select(1))

# K3: AdamW on a 64M-element shard
n = 64 << 20
master = torch.randn(n, device=DEV)
m = torch.zeros(n, device=DEV)
v = torch.zeros(n, device=DEV)
g = torch.randn(n, device=DEV).to(torch.bfloat16)
out = torch.empty(n, dtype=torch.bfloat16, device=DEV)
cfg = OptimConfig(kind="adamw", lr=1e-3)
for _ in range(reps):
    native.optim_step(cfg, master, g, m, v, out)

# K1: fused push on one rank (reduce over 1 peer + AdamW + replica store), 64M-element bucket
plan = plan_buckets([("vec", (n,))], 1, bucket_elems=1 << 62)
fab = Fabric(0, 1, DEV, plan.total_elems * 4 + (8 << 20))
be = DeviceBackend(fab, plan, OptimConfig(kind="adamw", lr=1e-3))
be.grad_flat.normal_()
for _ in range(reps):
    be.push(0)
for _ in range(reps):
    be.push(0, tma=True, grid=32)  # bulk-copy variant on 32 SMs

# K4 / K6
for _ in range(reps):
    native.sumsq(g)
bucket = torch.empty(n, dtype=torch.bfloat16, device=DEV)
srcs = [g[i * (n // 8):(i + 1) * (n // 8)] for i in range(8)]
for _ in range(reps):
    native.gather_to_bucket(srcs, bucket, [i * (n // 8) for i in range(8)])
# training GEMMs of the MLP (cluster-launch-control tile walk): forward + GELU, dX * dGELU with an
# MN-major B operand, dW with both operands MN-major; bias gradient; fused MSE loss
if len(sys.argv) > 2 and sys.argv[2] == "train":
    B, K_in, N_out = 4096, 8192, 8192
    a_in = (torch.randn(B, K_in, device=DEV) * 0.1).to(torch.bfloat16)
    wt = (torch.randn(N_out, K_in, device=DEV) * 0.02).to(torch.bfloat16)
    bias = (torch.randn(N_out, device=DEV) * 0.1).to(torch.bfloat16)
    z = torch.empty(B, N_out, dtype=torch.bfloat16, device=DEV)
    gz = (torch.randn(B, N_out, device=DEV) * 0.1).to(torch.bfloat16)
    zin = (torch.randn(B, K_in, device=DEV)).to(torch.bfloat16)
    for _ in range(reps):
        native.gemm(a_in, wt, bias=bias, epilogue="gelu", aux=z)
        native.gemm(gz, wt, trans_b=True, epilogue="dgelu", aux=zin)
        native.gemm(gz, a_in, trans_a=True, trans_b=True)
        native.colsum(gz)
        loss = native.mse_fwd(a_in, zin)
        native.mse_bwd(a_in, zin, None)
# K5: sparse embedding pull and push (mark / accumulate / apply), a 1 Mi x 128 table, 64 Ki rows
if "--sparse" in sys.argv:
    from parameter_server_distributed_b200 import ShardedEmbedding
    table = ShardedEmbedding(1 << 20, 128, OptimConfig(kind="adam", lr=1e-3), capacity=64 << 10,
                             device=DEV)
    ids = torch.randint(0, 1 << 20, (64 << 10,), device=DEV)
    gr = torch.randn(64 << 10, 128, device=DEV).to(torch.bfloat16)
    for _ in range(reps):
        table.pull(ids)
        table.push(ids, gr)
torch.cuda.synchronize()
print("ncu_target ok")
