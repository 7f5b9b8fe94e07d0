#!/usr/bin/env python
"""ncu target: the training GEMMs of the MLP on the CTA-pair kernel with the staged epilogue —
forward + GELU (4096 x 8192 x 8192), dX * dGELU with long and short K, dW.  A few launches each."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from parameter_server_distributed_b200.ops import native

dev = torch.device("cuda:0")
torch.manual_seed(0)
B, H, O = 4096, 8192, 1024
r = lambda *s: (torch.randn(*s, device=dev) * 0.1).to(torch.bfloat16)
x, w1, b1 = r(B, H), r(H, H), r(H)
dy_o, w2 = r(B, O), r(O, H)
dy_h, z = r(B, H), r(B, H)
zbuf = torch.empty(B, H, dtype=torch.bfloat16, device=dev)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    native.gemm(x, w1, bias=b1, epilogue="gelu", aux=zbuf)              # forward + GELU
    native.gemm(dy_h, w1, trans_b=True, epilogue="dgelu", aux=z)        # dX * dGELU, K = 8192
    native.gemm(dy_o, w2, trans_b=True, epilogue="dgelu", aux=z)        # dX * dGELU, K = 1024
    native.gemm(dy_h, x, trans_a=True, trans_b=True)                    # dW
torch.cuda.synchronize()
print("ok")
