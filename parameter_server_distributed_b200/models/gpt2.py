"""GPT-2 (BASELINE.json config 4: "GPT-2 medium with Adam update fused into the server-side
reduce-scatter").  Architecture of the public GPT-2: learned positions, pre-LayerNorm blocks,
tanh-GELU MLP, tied input/output embeddings; random init (no checkpoints offline).  Attention
uses ``F.scaled_dot_product_attention`` (library kernel); the parameter-server kernels of this
repo handle every gradient push / optimizer update / parameter pull.  With ``own_gemms`` the
four Linear layers of every block run as links of :mod:`..ops.linear` (tcgen05 GEMMs for
forward, dX and dW, GELU / dGELU fused into the c_fc / c_proj2 epilogues); the tied LM head
(50 257 rows, not a multiple of 8) stays a library GEMM.
The reference has no model (src/worker.cpp:316-320 fills gradients with 0.01).
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import linear


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    n_positions: int = 1024
    n_embd: int = 1024
    n_layer: int = 24
    n_head: int = 16
    # block Linears on the repo's tcgen05 kernel (CUDA bf16 only; PSB_GPT2_OWN_GEMMS=0/1 overrides)
    own_gemms: bool = False

    @staticmethod
    def medium() -> "GPT2Config":
        return GPT2Config()

    @staticmethod
    def small() -> "GPT2Config":
        return GPT2Config(n_embd=768, n_layer=12, n_head=12)

    @staticmethod
    def tiny() -> "GPT2Config":
        return GPT2Config(vocab_size=512, n_positions=64, n_embd=64, n_layer=2, n_head=4)

    def num_params(self) -> int:
        d = self.n_embd
        per_block = 12 * d * d + 13 * d
        return self.vocab_size * d + self.n_positions * d + self.n_layer * per_block + 2 * d

    def flops_per_token(self, seq_len: int) -> int:
        d, L = self.n_embd, self.n_layer
        matmul = 6 * (12 * d * d * L + self.vocab_size * d)
        attn = 6 * 2 * L * seq_len * d  # QK^T and PV, forward + backward (causal not discounted)
        return matmul + attn


class Block(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        d = cfg.n_embd
        self.n_head = cfg.n_head
        env = os.environ.get("PSB_GPT2_OWN_GEMMS")
        self.own_gemms = cfg.own_gemms if env is None else env != "0"
        self.ln_1 = nn.LayerNorm(d)
        self.c_attn = nn.Linear(d, 3 * d)
        self.c_proj = nn.Linear(d, d)
        self.ln_2 = nn.LayerNorm(d)
        self.c_fc = nn.Linear(d, 4 * d)
        self.c_proj2 = nn.Linear(4 * d, d)

    def _lin(self, layer: nn.Linear, x: torch.Tensor) -> torch.Tensor:
        if self.own_gemms:
            return linear.link(x, None, layer.weight, layer.bias, gelu_in=False, gelu_out=False)[0]
        return layer(x)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        B, T, D = x.shape
        q, k, v = self._lin(self.c_attn, self.ln_1(x)).split(D, dim=2)
        h = self.n_head
        q = q.view(B, T, h, D // h).transpose(1, 2)
        k = k.view(B, T, h, D // h).transpose(1, 2)
        v = v.view(B, T, h, D // h).transpose(1, 2)
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        x = x + self._lin(self.c_proj, a.transpose(1, 2).reshape(B, T, D))
        if self.own_gemms:
            z, act = linear.link(self.ln_2(x), None, self.c_fc.weight, self.c_fc.bias,
                                 gelu_in=False, gelu_out=True)
            return x + linear.link(z, act, self.c_proj2.weight, self.c_proj2.bias,
                                   gelu_in=True, gelu_out=False)[0]
        x = x + self.c_proj2(F.gelu(self.c_fc(self.ln_2(x)), approximate="tanh"))
        return x


class GPT2(nn.Module):
    def __init__(self, cfg: GPT2Config = GPT2Config()):
        super().__init__()
        self.cfg = cfg
        self.wte = nn.Embedding(cfg.vocab_size, cfg.n_embd)
        self.wpe = nn.Embedding(cfg.n_positions, cfg.n_embd)
        self.h = nn.ModuleList([Block(cfg) for _ in range(cfg.n_layer)])
        self.ln_f = nn.LayerNorm(cfg.n_embd)
        self.apply(self._init)
        for name, p in self.named_parameters():
            if name.endswith("c_proj.weight") or name.endswith("c_proj2.weight"):
                nn.init.normal_(p, mean=0.0, std=0.02 / math.sqrt(2 * cfg.n_layer))

    @staticmethod
    def _init(m: nn.Module) -> None:
        if isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)
            if m.bias is not None:
                nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Embedding):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)

    def forward(self, idx: torch.Tensor) -> torch.Tensor:
        B, T = idx.shape
        pos = torch.arange(T, device=idx.device)
        x = self.wte(idx) + self.wpe(pos)
        for blk in self.h:
            x = blk(x)
        x = self.ln_f(x)
        return F.linear(x, self.wte.weight)  # tied output embedding


def lm_loss(logits: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
    """Next-token cross entropy; ``targets`` are the input ids shifted by the caller."""
    return F.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), targets.reshape(-1))
