"""Checkpoints in the reference's binary format for the sharded engine.

The reference PS writes its whole ``vector<tensor>`` as raw little-endian structs
(src/parameter_server.cpp:112-144): ``int32 epoch, int32 iteration, uint64 n`` then per tensor
``uint64 name_len, name, uint64 ndim, int32 shape[], int32 dtype, uint64 numel, float32 data[]``.
Here the fp32 master weights are gathered from the shards (K8 of the survey: shard -> host ->
file) and written in exactly that layout, so the file loads in the reference's
``load_checkpoint`` and in ``parameter_server`` of this repo.  Everything the format cannot hold
(optimizer moments, step counter, optimizer kind) goes to a ``<path>.psb`` sidecar with the same
layout the C++ server uses (csrc/ps/ps_core.cpp).  Writes are atomic (temp file + rename).
"""
from __future__ import annotations

import os
import struct
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch


def write_reference_file(path: str, epoch: int, iteration: int,
                         tensors: List[Tuple[str, Tuple[int, ...], np.ndarray]]) -> None:
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(struct.pack("<iiQ", epoch, iteration, len(tensors)))
        for name, shape, data in tensors:
            nb = name.encode()
            f.write(struct.pack("<Q", len(nb)))
            f.write(nb)
            f.write(struct.pack("<Q", len(shape)))
            f.write(struct.pack(f"<{len(shape)}i", *shape))
            f.write(struct.pack("<i", 0))  # dtype 0 = float32
            flat = np.ascontiguousarray(data, dtype="<f4").reshape(-1)
            f.write(struct.pack("<Q", flat.size))
            f.write(flat.tobytes())
    os.replace(tmp, path)


def read_reference_file(path: str):
    """-> (epoch, iteration, [(name, shape, float32 array)])"""
    with open(path, "rb") as f:
        raw = f.read()
    pos = 0

    def take(fmt):
        nonlocal pos
        vals = struct.unpack_from(fmt, raw, pos)
        pos += struct.calcsize(fmt)
        return vals

    epoch, iteration, n = take("<iiQ")
    out = []
    for _ in range(n):
        (nl,) = take("<Q")
        name = raw[pos:pos + nl].decode()
        pos += nl
        (nd,) = take("<Q")
        shape = take(f"<{nd}i")
        take("<i")
        (ne,) = take("<Q")
        data = np.frombuffer(raw, dtype="<f4", count=ne, offset=pos).copy()
        pos += ne * 4
        out.append((name, tuple(shape), data))
    if pos != len(raw):
        raise ValueError(f"{path}: {len(raw) - pos} trailing bytes")
    return epoch, iteration, out


def _write_sidecar(path: str, kind: str, step: int, states: List[Tuple[np.ndarray, np.ndarray]]):
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(b"PSBS")
        f.write(struct.pack("<Iq", 1, step))
        kb = kind.encode()
        f.write(struct.pack("<Q", len(kb)))
        f.write(kb)
        f.write(struct.pack("<Q", len(states)))
        for m, v in states:
            for s in (m, v):
                flat = np.ascontiguousarray(s, dtype="<f4").reshape(-1)
                f.write(struct.pack("<Q", flat.size))
                f.write(flat.tobytes())
    os.replace(tmp, path)


def _read_sidecar(path: str):
    with open(path, "rb") as f:
        raw = f.read()
    if raw[:4] != b"PSBS":
        raise ValueError(f"{path}: bad magic")
    ver, step = struct.unpack_from("<Iq", raw, 4)
    pos = 16
    (nk,) = struct.unpack_from("<Q", raw, pos)
    pos += 8
    kind = raw[pos:pos + nk].decode()
    pos += nk
    (n,) = struct.unpack_from("<Q", raw, pos)
    pos += 8
    states = []
    for _ in range(n):
        pair = []
        for _ in range(2):
            (ne,) = struct.unpack_from("<Q", raw, pos)
            pos += 8
            pair.append(np.frombuffer(raw, dtype="<f4", count=ne, offset=pos).copy())
            pos += ne * 4
        states.append(tuple(pair))
    return kind, step, states


def save_checkpoint(ps, path: str, epoch: int = 0, iteration: Optional[int] = None,
                    validate=None) -> None:
    """Collective: every rank calls it; rank 0 writes ``path`` (+ ``path.psb`` when the optimizer
    has state).  Tensors are written in the model's own parameter order.  ``validate`` is
    called after the state has been gathered and before anything is written; it raises if the
    gathered data cannot be trusted (a peer was lost while the kernels ran)."""
    masters = ps.gather_state("master")
    m = ps.gather_state("m")
    v = ps.gather_state("v")
    if validate is not None:
        validate()
    if ps.rank != 0:
        return
    names = [n for n, _ in ps.named_params]
    tensors = [(n, tuple(masters[n].shape), masters[n].detach().float().cpu().contiguous().numpy())
               for n in names]
    it = ps.steps_done - 1 if iteration is None else iteration
    write_reference_file(path, epoch, max(it, 0), tensors)
    side = path + ".psb"
    if m:
        step = int(ps.backend.step_word.item())
        states = [(m[n].detach().float().cpu().contiguous().numpy(),
                   v[n].detach().float().cpu().contiguous().numpy() if v else np.zeros(0, "<f4"))
                  for n in names]
        _write_sidecar(side, ps.optim.kind, step, states)
    elif os.path.exists(side):
        os.remove(side)


def load_checkpoint(ps, path: str) -> Tuple[int, int]:
    """Collective: every rank reads the file and keeps its own shard.  Returns (epoch, iteration)."""
    epoch, iteration, tensors = read_reference_file(path)
    known = dict(ps.named_params)
    masters: Dict[str, torch.Tensor] = {}
    for name, shape, data in tensors:
        if name not in known:
            continue
        if tuple(known[name].shape) != tuple(shape):
            raise ValueError(f"{path}: shape of {name} is {shape}, model has {tuple(known[name].shape)}")
        masters[name] = torch.from_numpy(data).view(*shape)
    ps.scatter_state(masters, "master")
    side = path + ".psb"
    if os.path.exists(side) and ps.backend.m is not None:
        kind, step, states = _read_sidecar(side)
        if kind == ps.optim.kind and len(states) == len(tensors):
            ms = {t[0]: torch.from_numpy(s[0]).view(*t[1]) for t, s in zip(tensors, states)
                  if t[0] in known and s[0].size}
            ps.scatter_state(ms, "m")
            if ps.backend.v is not None:
                vs = {t[0]: torch.from_numpy(s[1]).view(*t[1]) for t, s in zip(tensors, states)
                      if t[0] in known and s[1].size}
                ps.scatter_state(vs, "v")
            ps.backend.step_word.fill_(step)
    ps.steps_done = iteration + 1
    return epoch, iteration
