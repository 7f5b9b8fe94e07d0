#!/usr/bin/env python
"""ncu target: the short-K dX*dGELU GEMM of the MLP's last layer (4096 x 8192 x 1024), a few launches."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from parameter_server_distributed_b200.ops import native

dev = torch.device("cuda:0")
torch.manual_seed(0)
M, N, K = 4096, 8192, 1024
dy = (torch.randn(M, K, device=dev) * 0.1).to(torch.bfloat16)
w = (torch.randn(K, N, device=dev) * 0.1).to(torch.bfloat16)
z = torch.randn(M, N, device=dev).to(torch.bfloat16)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 6):
    out = native.gemm(dy, w, trans_b=True, epilogue="dgelu", aux=z)
    out2 = native.gemm(dy, w, trans_b=True)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
