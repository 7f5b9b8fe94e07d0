"""Sparse-feature classifier: a sharded embedding table (K5) under a small dense tower.

The reference ships no model; this is the model family its dense wire format cannot train — a
table too large to push whole (``parallel/sparse.py``).  ``fields`` categorical ids per sample
are looked up on the parameter server (only those rows move), concatenated and classified by an
MLP whose parameters go through the dense engine like any other model.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..parallel.sparse import ShardedEmbedding


@dataclass
class EmbeddingMLPConfig:
    vocab: int = 100_000
    dim: int = 64
    fields: int = 8
    hidden: Tuple[int, ...] = (256, 128)
    classes: int = 2


class EmbeddingMLP(nn.Module):
    """``forward(ids[B, fields]) -> logits[B, classes]``.  The table is not an ``nn.Parameter``:
    its rows live with their owners; gradients of the looked-up rows are pushed by
    ``table.apply_gradients()`` after ``backward()``."""

    def __init__(self, cfg: EmbeddingMLPConfig, table: ShardedEmbedding):
        super().__init__()
        if table.vocab != cfg.vocab or table.dim != cfg.dim:
            raise ValueError("table and config disagree on vocab / dim")
        self.cfg, self.table = cfg, table
        dims = (cfg.fields * cfg.dim,) + tuple(cfg.hidden) + (cfg.classes,)
        self.tower = nn.ModuleList([nn.Linear(dims[i], dims[i + 1]) for i in range(len(dims) - 1)])

    def forward(self, ids: torch.Tensor) -> torch.Tensor:
        x = self.table.lookup(ids).reshape(ids.shape[0], -1)
        x = x.to(self.tower[0].weight.dtype)
        for i, layer in enumerate(self.tower):
            x = layer(x)
            if i + 1 < len(self.tower):
                x = F.gelu(x)
        return x
