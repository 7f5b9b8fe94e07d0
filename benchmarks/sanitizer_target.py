#!/usr/bin/env python
"""Small-size pass over every kernel for `compute-sanitizer --tool memcheck` (T8): sizes with
ragged tails, emulated multi-rank K1 / K2 / fused GEMM on one GPU (ranks launched one after the
other, never concurrently spinning kernels)."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

from parameter_server_distributed_b200 import OptimConfig, _lib
from parameter_server_distributed_b200.ops import native
from parameter_server_distributed_b200.parallel.buckets import plan_buckets
from parameter_server_distributed_b200.parallel.device_backend import DeviceBackend
from parameter_server_distributed_b200.parallel.fabric import Fabric

DEV = torch.device("cuda:0")
torch.cuda.set_device(DEV)
n = 4096 + 13
f = torch.empty(n, device=DEV)
native.fill_(f[3:], 0.5)
h = torch.empty(n, dtype=torch.bfloat16, device=DEV)
native.cast_scale(f, h, 2.0)
native.cast_scale(h, f, 0.5)
native.sumsq(h)
native.sumsq(f)
bucket = torch.zeros(n + 256, dtype=torch.bfloat16, device=DEV)
native.gather_to_bucket([h[:1000].clone(), h[:77].clone()], bucket, [0, 1024])
for kind in ("sgd", "momentum", "adam", "adamw"):
    m, v = torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    native.optim_step(OptimConfig(kind=kind, lr=0.01), f.clone(), h, m, v, torch.empty_like(h))
for variant in (1, 0):
    _lib.check(_lib.load().psb_gemm_select(variant))
    x = torch.randn(300, 72, device=DEV).to(torch.bfloat16)
    w = torch.randn(264, 72, device=DEV).to(torch.bfloat16)
    native.gemm_bf16(x, w)
_lib.check(_lib.load().psb_gemm_select(1))
world = 4
plan = plan_buckets([("a", (3000,)), ("b", (70, 33))], world, bucket_elems=4000)
fabs = Fabric.local_group(world, DEV, plan.total_elems * 8 + (4 << 20))
for tma in (False, True):
    bes = [DeviceBackend(fb, plan, OptimConfig(kind="adamw", lr=0.01, grad_scale=0.25),
                         push_tma=tma, clip=True) for fb in fabs]
    for be in bes:
        be.grad_flat.normal_()
    for b in range(len(plan.buckets)):
        for be in bes:
            be.push(b, no_wait=True)
        for be in bes:
            be.push(b, no_wait=True, reduce_only=True)
        for be in bes:
            be.push(b, from_reduced=True, clip=True)
        for be in bes:
            be.pull(b)
torch.cuda.synchronize()
print("sanitizer_target ok")
